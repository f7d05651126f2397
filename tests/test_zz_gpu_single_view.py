"""Single-view scene through the CUDA path (the smallest scene the reference accepts: global attention
over one view's 24 tokens + the scale token, one dense-head pass shorter than head_chunk) against a
golden of the LIVE reference.  Runs last (file name) so that it cannot mask another test."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_single_view_matches_reference_golden(synthetic_sd):
    from mapanything_b200 import MapAnything
    from oracle import make_golden
    from oracle import mapanything_oracle as orc
    from tests import helpers

    name = "small_n1_56x84"
    m = MapAnything(device="cuda:0", head_chunk=2)
    m.load_state_dict(synthetic_sd)
    n, H, W, mm, stride, kw = make_golden.CASES[name]
    views = make_golden.case_views(name)
    preds = m.infer([dict(v) for v in views], **kw)
    torch.cuda.synchronize()
    g = helpers.load_golden(name)
    rep = helpers.compare_with_golden(preds, g, stride, None, None)
    sd_cuda = {k: v.cuda() for k, v in synthetic_sd.items()}
    views_cuda = [{k: (v.cuda() if torch.is_tensor(v) else v) for k, v in view.items()} for view in views]
    yard = helpers.compare_with_golden(orc.infer(sd_cuda, views_cuda, amp=True, **kw), g, stride, None, None)
    print("libmab200  :", {k: round(v, 5) for k, v in rep.items()})
    print("ref bf16   :", {k: round(v, 5) for k, v in yard.items()})
    cam_frame = {k: v for k, v in rep.items() if k.split(".", 1)[1] not in helpers.POSE_DEPENDENT_KEYS
                 and k.split(".", 1)[1] != "non_ambiguous_mask_logits"}
    bad = helpers.out_of_tolerance(cam_frame, 1e-2, 0.1, 1e-2, min_mask_agree=helpers.MIN_MASK_AGREE)
    assert not bad, f"camera-frame keys out of absolute tolerance: {bad}"
    helpers.record_parity("golden:" + name, rep, yard)
    bad = helpers.out_of_tolerance(rep, 1e-2, 0.1, 1e-2, min_mask_agree=helpers.MIN_MASK_AGREE, yardstick=yard, tol_intrinsics=helpers.tol_intrinsics_for(H))
    assert not bad, f"further from the fp32 reference than the reference's bf16 path: {bad}\nyardstick {yard}"
    assert len(preds) == 1 and preds[0]["pts3d"].shape == (1, H, W, 3)
