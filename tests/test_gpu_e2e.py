"""End-to-end parity of the CUDA path (through MapAnything.infer -> C ABI) against
 (a) golden vectors generated from the LIVE reference (tests/golden, oracle/make_golden.py) and
 (b) the oracle's stage taps, at the tolerances BASELINE.json states for bf16:
     pointmap/depth relative L2 <= 1e-2, pose angular error <= 0.1 deg."""
import pytest
import torch

pytestmark = pytest.mark.gpu

TOL_DENSE = 1e-2
TOL_POSE_DEG = 0.1


@pytest.fixture(scope="module")
def model(synthetic_sd):
    from mapanything_b200 import MapAnything

    m = MapAnything(device="cuda:0", head_chunk=2)
    m.load_state_dict(synthetic_sd)
    return m


@pytest.mark.parametrize("name", ["small_n3_70x98", "c1_loose_n2_392x518", "c1_n2_392x518"])
def test_infer_matches_reference_golden(name, model):
    from oracle import make_golden
    from tests import helpers

    n, H, W, mm, stride, kw = make_golden.CASES[name]
    views = make_golden.case_views(name)
    preds = model.infer([dict(v) for v in views], **kw)
    torch.cuda.synchronize()
    g = helpers.load_golden(name)
    rep = helpers.compare_with_golden(preds, g, stride, tol_dense=TOL_DENSE, tol_pose_deg=TOL_POSE_DEG,
                                      tol_small=TOL_DENSE, min_mask_agree=0.97)
    print({k: round(v, 5) for k, v in rep.items()})
    p = preds[0]
    assert p["pts3d"].shape == (1, H, W, 3) and p["pts3d"].dtype == torch.float32
    assert p["mask"].shape == (1, H, W, 1) and p["mask"].dtype == torch.bool
    assert p["non_ambiguous_mask"].shape == (1, H, W) and p["non_ambiguous_mask"].dtype == torch.bool
    assert p["metric_scaling_factor"].shape == (1, 1) and p["camera_poses"].shape == (1, 4, 4)
    assert p["conf"].shape == (1, H, W) and p["depth_z"].shape == (1, H, W, 1)
    assert p["intrinsics"].shape == (1, 3, 3) and p["img_no_norm"].shape == (1, H, W, 3)
    assert all(t.device.type == "cuda" for t in p.values())


def test_forward_raw_matches_oracle(model, synthetic_sd):
    """raw forward dicts (model.py:1907-1923) vs the fp32 oracle, incl. the views-chunked dense head"""
    from oracle import mapanything_oracle as orc, weights
    from tests import helpers

    views = weights.synthetic_views(5, 56, 84, seed=7)
    raw = model.forward([dict(v) for v in views])
    torch.cuda.synchronize()
    with torch.inference_mode():
        ref = orc.forward(synthetic_sd, orc.preprocess_views([dict(v) for v in views]))
    assert len(raw) == 5
    for p, r in zip(raw, ref):
        for k in ("pts3d", "pts3d_cam", "ray_directions", "depth_along_ray", "conf", "non_ambiguous_mask_logits",
                  "cam_trans", "metric_scaling_factor"):
            assert p[k].shape == r[k].shape, k
            e = helpers.rel_l2(p[k].float().cpu().numpy(), r[k].numpy())
            assert e < TOL_DENSE, (k, e)
        assert helpers.quat_angle_deg(p["cam_quats"].cpu().numpy(), r["cam_quats"].numpy()) < TOL_POSE_DEG
        agree = float((p["non_ambiguous_mask"].cpu() == r["non_ambiguous_mask"]).float().mean())
        assert agree > 0.99


def test_validation_errors_match_reference(model):
    v = {"img": torch.zeros(1, 3, 28, 28), "data_norm_type": ["dinov2"]}
    with pytest.raises(ValueError):
        model.infer([])
    with pytest.raises(ValueError):
        model.infer([{"img": v["img"]}])
    with pytest.raises(ValueError):
        model.infer([dict(v, depth_z=torch.zeros(1, 28, 28))])
    with pytest.raises(ValueError):
        model.infer([v, dict(v, camera_poses=torch.eye(4)[None])])
    with pytest.raises(AssertionError):
        model.infer([dict(v, img=torch.zeros(1, 3, 30, 28))])
