"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/export_*.npz from the LIVE reference.

    python -m oracle.make_golden_export       (needs /root/reference; run in the build container)

Runs mapanything.utils.geometry.depthmap_to_world_frame (geometry.py:76-114) and the point-cloud
assembly of scripts/demo_local_weight.py:122-140 / mapanything/utils/viz.py:233-239,275-279 (plain
numpy there; restated verbatim here because viz.py imports the viewer stack) on seeded inputs.
"""
import os

import numpy as np
import torch

from oracle import ref_shim

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    ref_shim.install()
    from mapanything.utils.geometry import depthmap_to_world_frame

    for name, (V, H, W, seed) in {"export_v3_70x98": (3, 70, 98, 11), "export_v2_37x330": (2, 37, 330, 12)}.items():
        g = torch.Generator().manual_seed(seed)
        depth = torch.rand(V, H, W, generator=g) * 4 + 0.5
        depth[torch.rand(V, H, W, generator=g) < 0.2] = 0.0          # invalid pixels
        depth[0, :3] = -1.0                                             # negative depth is invalid too
        K = torch.eye(3)[None].repeat(V, 1, 1)
        K[:, 0, 0] = 90.0 + torch.arange(V); K[:, 1, 1] = 88.0; K[:, 0, 2] = W / 2 - 0.5; K[:, 1, 2] = H / 2 + 0.25
        ang = torch.arange(V).float() * 0.3 + 0.1
        pose = torch.eye(4)[None].repeat(V, 1, 1)
        pose[:, 0, 0] = ang.cos(); pose[:, 0, 2] = ang.sin(); pose[:, 2, 0] = -ang.sin(); pose[:, 2, 2] = ang.cos()
        pose[:, :3, 3] = torch.randn(V, 3, generator=g)
        mask = torch.rand(V, H, W, generator=g) < 0.7
        mask[V - 1] = False                                             # a view that contributes nothing
        rgb = torch.rand(V, H, W, 3, generator=g)
        pts_world, valid = depthmap_to_world_frame(depth, K, pose)
        pts_cam, _ = depthmap_to_world_frame(depth, K, None)
        keep = mask.numpy() & valid.numpy()
        colors = (rgb.numpy().reshape(V, -1, 3) * 255).astype(np.uint8)
        cloud = np.concatenate([pts_world[v].numpy().reshape(-1, 3)[keep[v].reshape(-1)] for v in range(V)], 0)
        cols = np.concatenate([colors[v][keep[v].reshape(-1)] for v in range(V)], 0)
        path = os.path.join(ROOT, "tests", "golden", name + ".npz")
        np.savez_compressed(path, depth=depth.numpy(), intrinsics=K.numpy(), pose=pose.numpy(), mask=mask.numpy(),
                            rgb=rgb.numpy(), pts_world=pts_world.numpy(), pts_cam=pts_cam.numpy(),
                            valid=valid.numpy(), cloud=cloud, colors=cols,
                            counts=keep.reshape(V, -1).sum(1).astype(np.int64))
        print("wrote", path, os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main()
