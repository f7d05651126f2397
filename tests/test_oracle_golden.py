"""Pins the CPU restatement (oracle/mapanything_oracle.py) to the LIVE reference's outputs.

tests/golden/*.npz were produced by oracle/make_golden.py from the unmodified reference
`MapAnything.infer` (fp32, CPU) in the build container; this test needs only the fixtures."""
import pytest
import torch

from oracle import make_golden, mapanything_oracle as orc
from tests import helpers

CASES = ["small_n1_56x84", "small_n3_70x98", "small_mm_n3_70x98", "small_conf_n2_56x84", "c1_loose_n2_392x518",
         "small_escaling_n3_70x98", "small_noref_n3_70x98", "small_48l_n2_56x84",
         "small_viewpe_rand_n3_70x98", "small_viewpe_seq_n3_70x98", "small_gat_n3_70x98"]


@pytest.mark.parametrize("name", CASES)
def test_restatement_matches_reference_golden(name, synthetic_sd, monkeypatch):
    n, H, W, mm, stride, kw = make_golden.CASES[name]
    views = make_golden.case_views(name)
    margs = make_golden.CASE_MODULE_ARGS.get(name, {})
    # the 48-layer trunk and the view-PE table have their own shapes
    own = "dim" in margs or margs.get("use_pe_for_non_reference_views") or "max_num_views" in margs
    sd = make_golden.case_state_dict(name) if own else synthetic_sd
    g = helpers.load_golden(name)
    if "view_pe_rows" in g:  # the rows the reference drew for this fixture (torch.randint, aat.py:631)
        monkeypatch.setattr(torch, "randint", lambda *a, **k: torch.from_numpy(g["view_pe_rows"]))
    preds = orc.infer(sd, views, aat_opts=orc.aat_opts_for(make_golden.case_info_sharing_config(name)), **kw)
    monkeypatch.undo()
    rep = helpers.compare_with_golden(preds, g, stride, tol_dense=2e-5, tol_pose_deg=1e-3,
                                      min_mask_agree=0.9995)
    assert len(preds) == n
    # every key/shape/dtype of the reference's prediction dict
    p = preds[0]
    assert p["pts3d"].shape == (1, H, W, 3) and p["pts3d"].dtype == torch.float32
    assert p["conf"].shape == (1, H, W)
    assert p["non_ambiguous_mask"].dtype == torch.bool and p["non_ambiguous_mask"].shape == (1, H, W)
    assert p["mask"].dtype == torch.bool and p["mask"].shape == (1, H, W, 1)
    assert p["metric_scaling_factor"].shape == (1, 1)
    assert p["camera_poses"].shape == (1, 4, 4) and p["intrinsics"].shape == (1, 3, 3)
    assert rep


def test_validation_errors():
    v = {"img": torch.zeros(1, 3, 28, 28), "data_norm_type": ["dinov2"]}
    with pytest.raises(ValueError):
        orc.validate_views([])
    with pytest.raises(ValueError):
        orc.validate_views([{"img": v["img"]}])
    with pytest.raises(ValueError):
        orc.validate_views([dict(v, intrinsics=torch.eye(3)[None], ray_directions=torch.zeros(1, 28, 28, 3))])
    with pytest.raises(ValueError):
        orc.validate_views([dict(v, depth_z=torch.zeros(1, 28, 28))])
    with pytest.raises(ValueError):
        orc.validate_views([v, dict(v, camera_poses=torch.eye(4)[None])])
