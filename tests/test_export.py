"""Dense exports: the oracle against live-reference golden vectors
(CPU) and mab_export_points / mapanything_b200.export against oracle + golden (GPU)."""
import glob
import os

import numpy as np
import pytest
import torch

from oracle import export_oracle

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "export_*.npz")))
IDS = [os.path.basename(p) for p in GOLDEN]
# points: fp32 with the reference's operand order; the reference's einsum may contract with FMA and
# sum in another order, hence a relative tolerance.  Masks, counts, order, colours: bit-exact.
RTOL = 2e-6


def _close(a, b):
    return np.abs(a - b).max() <= RTOL * max(1.0, np.abs(b).max())


def test_golden_files_exist():
    assert len(GOLDEN) == 2


@pytest.mark.parametrize("path", GOLDEN, ids=IDS)
def test_oracle_matches_reference_golden(path):
    g = np.load(path)
    pw, valid = export_oracle.depthmap_to_world_frame(g["depth"], g["intrinsics"], g["pose"])
    pc, _ = export_oracle.depthmap_to_world_frame(g["depth"], g["intrinsics"], None)
    assert np.array_equal(valid, g["valid"])
    assert _close(pw, g["pts_world"]) and _close(pc, g["pts_cam"])
    pts, col, n = export_oracle.point_cloud(g["depth"], g["intrinsics"], g["pose"], g["mask"], g["rgb"])
    assert np.array_equal(n, g["counts"]) and np.array_equal(col, g["colors"])
    assert pts.shape == g["cloud"].shape and _close(pts, g["cloud"])


@pytest.mark.gpu
@pytest.mark.parametrize("path", GOLDEN, ids=IDS)
def test_cuda_matches_golden_and_oracle(path):
    from mapanything_b200 import export

    g = np.load(path)
    d, K, P = (torch.from_numpy(g[k]).cuda() for k in ("depth", "intrinsics", "pose"))
    pw, valid = export.depthmap_to_world_frame(d, K, P)
    pc, _ = export.depthmap_to_world_frame(d, K)
    assert np.array_equal(valid.cpu().numpy(), g["valid"])
    assert _close(pw.cpu().numpy(), g["pts_world"]) and _close(pc.cpu().numpy(), g["pts_cam"])
    # unbatched form (geometry.py:85-91,108-110)
    p1, v1 = export.depthmap_to_world_frame(d[1], K[1], P[1])
    assert p1.shape == d[1].shape + (3,) and torch.equal(p1, pw[1]) and torch.equal(v1, valid[1])
    preds = [{"depth_z": d[v:v + 1, ..., None], "intrinsics": K[v:v + 1], "camera_poses": P[v:v + 1],
              "mask": torch.from_numpy(g["mask"][v:v + 1, ..., None]).cuda(),
              "img_no_norm": torch.from_numpy(g["rgb"][v:v + 1]).cuda()} for v in range(d.shape[0])]
    pts, col, n = export.export_point_cloud(preds)
    assert np.array_equal(n.cpu().numpy(), g["counts"])            # bit-exact: counts
    assert np.array_equal(col.cpu().numpy(), g["colors"])          # bit-exact: order + colours
    assert pts.shape == tuple(g["cloud"].shape) and _close(pts.cpu().numpy(), g["cloud"])
    ora = export_oracle.point_cloud(g["depth"], g["intrinsics"], g["pose"], g["mask"], g["rgb"])
    assert np.array_equal(pts.cpu().numpy(), ora[0])               # same operand order as the oracle: bit-exact


@pytest.mark.gpu
def test_cuda_compaction_properties_at_full_size():
    """8 views 392x518: compaction is a stable filter -- count == mask.sum(), all-ones mask is the identity,
    empty mask gives zero points, capacity clips the stores but not the counts."""
    from mapanything_b200 import export

    torch.manual_seed(5)
    V, H, W = 8, 392, 518
    d = torch.rand(V, H, W, device="cuda") + 0.1
    K = torch.tensor([[440.0, 0, 259], [0, 440.0, 196], [0, 0, 1]], device="cuda")[None].repeat(V, 1, 1)
    P = torch.eye(4, device="cuda")[None].repeat(V, 1, 1)
    rgb = torch.rand(V, H, W, 3, device="cuda")
    m = torch.rand(V, H, W, device="cuda") < 0.37
    dense, xyz, col, counts = export._run(d, K, P, m, rgb, True, True)
    n = counts.cpu().numpy()
    assert np.array_equal(n[:-1], m.reshape(V, -1).sum(1).cpu().numpy()) and n[-1] == int(m.sum())
    assert torch.equal(xyz[:n[-1]], dense[m])                       # stable order == boolean indexing
    assert torch.equal(col[:n[-1]], (rgb * 255).to(torch.uint8)[m])
    _, xyz1, _, c1 = export._run(d, K, P, torch.ones_like(m), rgb, True, True)
    assert int(c1[-1]) == V * H * W and torch.equal(xyz1, dense.reshape(-1, 3))
    _, _, _, c0 = export._run(d, K, P, torch.zeros_like(m), rgb, False, True)
    assert int(c0[-1]) == 0
    _, xyz2, _, c2 = export._run(d, K, P, m, rgb, False, True, capacity=1000)
    assert int(c2[-1]) == n[-1] and torch.equal(xyz2, xyz[:1000])
