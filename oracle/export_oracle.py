"""TEST INFRASTRUCTURE ONLY -- CPU restatement (numpy, fp32) of the reference's dense export path.

`depthmap_to_world_frame`: mapanything/utils/geometry.py:24-114 (camera-frame unprojection
`(x - cx) * z / fx`, then `pose @ [x, y, z, 1]`); the cloud: scripts/demo_local_weight.py:122-140
(`mask &= depth > 0`) and mapanything/utils/viz.py:233-239,275-279 (`points[mask]`,
`(rgb * 255).astype(uint8)[mask]`, frame by frame).  Pinned by tests/golden/export_*.npz, generated
from the LIVE reference functions by oracle/make_golden_export.py.  Only tests/ may import this file.
"""
import numpy as np


def depthmap_to_world_frame(depth, intrinsics, pose=None):
    """depth (B,H,W) f32, intrinsics (B,3,3), pose (B,4,4) or None -> (pts (B,H,W,3), valid (B,H,W))."""
    depth = np.asarray(depth, np.float32)
    K = np.asarray(intrinsics, np.float32)
    B, H, W = depth.shape
    xg, yg = np.meshgrid(np.arange(W, dtype=np.float32), np.arange(H, dtype=np.float32), indexing="xy")
    fx, fy = K[:, 0, 0].reshape(-1, 1, 1), K[:, 1, 1].reshape(-1, 1, 1)
    cx, cy = K[:, 0, 2].reshape(-1, 1, 1), K[:, 1, 2].reshape(-1, 1, 1)
    xx = (xg[None] - cx) * depth / fx  # geometry.py:60
    yy = (yg[None] - cy) * depth / fy
    pts = np.stack([xx, yy, depth], -1).astype(np.float32)
    valid = depth > 0.0
    if pose is not None:
        P = np.asarray(pose, np.float32)
        # geometry.py:98-103, summed left to right in fp32
        out = np.empty_like(pts)
        for i in range(3):
            out[..., i] = ((P[:, i, 0].reshape(-1, 1, 1) * pts[..., 0] + P[:, i, 1].reshape(-1, 1, 1) * pts[..., 1])
                           + P[:, i, 2].reshape(-1, 1, 1) * pts[..., 2]) + P[:, i, 3].reshape(-1, 1, 1)
        pts = out
    return pts, valid


def point_cloud(depth, intrinsics, pose, mask, rgb):
    """-> (points [n,3] f32, colors [n,3] u8, per-view counts)."""
    pts, valid = depthmap_to_world_frame(depth, intrinsics, pose)
    keep = valid if mask is None else (np.asarray(mask).astype(bool) & valid)  # demo_local_weight.py:131-132
    col = (np.asarray(rgb, np.float32).reshape(len(depth), -1, 3) * 255).astype(np.uint8)  # viz.py:239
    P, Cc, n = [], [], []
    for v in range(len(depth)):  # viz.py:275-279, frame by frame
        k = keep[v].reshape(-1)
        P.append(pts[v].reshape(-1, 3)[k])
        Cc.append(col[v][k])
        n.append(int(k.sum()))
    return np.concatenate(P, 0), np.concatenate(Cc, 0), np.asarray(n, np.int64)
