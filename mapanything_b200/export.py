"""Dense exports of the prediction dicts on the GPU.

Host-side mirror of `mapanything.utils.geometry.depthmap_to_world_frame` (geometry.py:76-114) and of
the mask-filtered point cloud every demo assembles on the CPU from `infer()`'s output
(scripts/demo_local_weight.py:122-140, mapanything/utils/viz.py:233-239,275-279), both as ONE pass of
libmab200's `mab_export_points` over device-resident predictions.  No CPU fallback.
"""
import ctypes as C
from typing import Dict, List, Optional, Tuple

import torch

from . import _lib


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def _run(depth, intr, pose, mask, rgb, want_dense, want_cloud, capacity=None):
    lib = _lib.load()
    assert depth.is_cuda, "mapanything_b200.export works on device tensors (no CPU path)"
    V, H, W = depth.shape
    dev = depth.device
    depth = depth.contiguous().float()
    intr = intr.contiguous().float()
    pose = pose.contiguous().float() if pose is not None else None
    mask = mask.contiguous().view(torch.uint8) if mask is not None else None
    rgb = rgb.contiguous().float() if rgb is not None else None
    ws_bytes = lib.mab_export_workspace_bytes(V, H, W)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
    counts = torch.empty(V + 1, dtype=torch.int64, device=dev)
    dense = torch.empty((V, H, W, 3), dtype=torch.float32, device=dev) if want_dense else None
    cap = V * H * W if capacity is None else int(capacity)
    xyz = torch.empty((cap, 3), dtype=torch.float32, device=dev) if want_cloud else None
    col = torch.empty((cap, 3), dtype=torch.uint8, device=dev) if want_cloud and rgb is not None else None
    st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    _lib.check(lib.mab_export_points(_ptr(depth), _ptr(intr), _ptr(pose), _ptr(mask), _ptr(rgb), V, H, W,
                                     _ptr(dense), _ptr(xyz), _ptr(col), cap, _ptr(counts), _ptr(ws), ws_bytes, st),
               "mab_export_points")
    return dense, xyz, col, counts


def depthmap_to_world_frame(depthmap: torch.Tensor, intrinsics: torch.Tensor,
                            camera_pose: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """geometry.py:76-114: depth map (HxW or BxHxW) + intrinsics (+ cam2world pose) -> point map in the
    world (or camera) frame and the `depth > 0` mask.  Same argument meaning and return shapes."""
    squeeze = depthmap.dim() == 2
    d = depthmap[None] if squeeze else depthmap
    K = intrinsics[None] if intrinsics.dim() == 2 else intrinsics
    P = None if camera_pose is None else (camera_pose[None] if camera_pose.dim() == 2 else camera_pose)
    dense, _, _, _ = _run(d, K, P, None, None, True, False, capacity=0)
    valid = d > 0.0
    return (dense[0], valid[0]) if squeeze else (dense, valid)


def export_point_cloud(predictions: List[Dict[str, torch.Tensor]], use_mask: bool = True):
    """The demos' cloud from `MapAnything.infer()` output (demo_local_weight.py:122-140): per view
    depth_z + intrinsics + camera_poses -> world points, kept where `mask & (depth > 0)`; colours
    `(img_no_norm * 255).astype(uint8)` (viz.py:239).  Returns (points [n,3] f32, colors [n,3] u8,
    per_view_counts [views] int64), all on the device, in view-major / row-major order."""
    d = torch.cat([p["depth_z"][..., 0] for p in predictions], 0)
    K = torch.cat([p["intrinsics"] for p in predictions], 0)
    P = torch.cat([p["camera_poses"] for p in predictions], 0)
    m = torch.cat([p["mask"][..., 0] for p in predictions], 0) if use_mask else None
    rgb = torch.cat([p["img_no_norm"] for p in predictions], 0)
    _, xyz, col, counts = _run(d, K, P, m, rgb, False, True)
    n = int(counts[-1].item())  # the one host sync: the caller needs the size
    return xyz[:n], col[:n], counts[:-1]
