"""Parity of the CUDA path on the configurations bench.py actually times (BASELINE.json):
configs[1] = 8 views 518x392 image-only on one B200 (the BENCH workload) and configs[2] = 32 views
multi-modal (images + intrinsics + poses + sparse depth), through MapAnything.infer -> C ABI, against the
fp32 oracle run on the same GPU (helpers.strict_fp32: no TF32 in the checker) with the oracle's
reference-precision run (bf16 autocast trunk, TF32 heads) as the yardstick.  Every per-key error goes
into the parity report (helpers.record_parity -> gpurun_out/parity_report.json, committed copy
profiles/r2_parity_report.json)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

TOL_DENSE = 1e-2     # BASELINE.json: pointmap / depth relative L2
TOL_POSE_DEG = 0.1   # BASELINE.json: pose angular error


def _to_cuda(views):
    return [{k: (v.cuda() if torch.is_tensor(v) else v) for k, v in view.items()} for view in views]


def _run_case(case, views, synthetic_sd, head_chunk=8):
    from mapanything_b200 import MapAnything
    from oracle import mapanything_oracle as orc
    from tests import helpers

    m = MapAnything(device="cuda:0", head_chunk=head_chunk)
    m.load_state_dict(synthetic_sd)
    sd_cuda = {k: v.cuda() for k, v in synthetic_sd.items()}
    preds = m.infer(_to_cuda(views))
    torch.cuda.synchronize()
    with helpers.strict_fp32():
        ref = orc.infer(sd_cuda, _to_cuda(views))
    amp = orc.infer(sd_cuda, _to_cuda(views), amp=True)
    rep = helpers.compare_pred_lists(preds, ref)
    yard = helpers.compare_pred_lists(amp, ref)
    entry = helpers.record_parity(case, rep, yard, extra={"views": len(views), "H": 392, "W": 518})
    print(case, "cuda     :", {k: round(v, 5) for k, v in entry["cuda"].items()})
    print(case, "ref bf16 :", {k: round(v, 5) for k, v in entry["ref_bf16"].items()})
    del m
    torch.cuda.empty_cache()
    return rep, yard, entry


def _assert_case(rep, yard, entry):
    from tests import helpers

    # (1) BASELINE.json's absolute tolerances on every key whose error is not the pose head's
    # amplification of trunk noise (DESIGN.md section 5 lists which keys those are and by how much)
    cam_frame = {k: v for k, v in rep.items() if k.split(".", 1)[1] not in helpers.POSE_DEPENDENT_KEYS
                 and k.split(".", 1)[1] != "non_ambiguous_mask_logits"}  # a near-cancelling sum: rule (2)
    bad = helpers.out_of_tolerance(cam_frame, TOL_DENSE, TOL_POSE_DEG, TOL_DENSE, min_mask_agree=helpers.MIN_MASK_AGREE)
    assert not bad, f"camera-frame keys out of absolute tolerance: {bad}"
    # (2) every key, pose-dependent ones included: inside the BASELINE tolerance, or no further from fp32
    # than helpers.YARDSTICK_SLACK (POSE_SLACK for pose-dependent keys) x the reference's own bf16 path on the same scene
    bad = helpers.out_of_tolerance(rep, TOL_DENSE, TOL_POSE_DEG, TOL_DENSE, min_mask_agree=helpers.MIN_MASK_AGREE,
                                   yardstick=yard)
    assert not bad, f"further from fp32 than the reference's own bf16 path allows: {bad}\n{entry}"


def test_config1_8_views_518x392_image_only(synthetic_sd):
    """BASELINE.json configs[1], the workload of BENCH_r*.json: seed-1 synthetic views (bench.py)."""
    from oracle import weights

    views = weights.synthetic_views(8, 392, 518, seed=1)
    _assert_case(*_run_case("config1_8views_518x392_image_only", views, synthetic_sd))


def test_config2_32_views_multimodal(synthetic_sd):
    """BASELINE.json configs[2] (`bench.py --views 32 --multimodal`): images + intrinsics + poses +
    sparse depth, 32 views."""
    from oracle import weights

    views = weights.add_multimodal_inputs(weights.synthetic_views(32, 392, 518, seed=1), seed=2)
    _assert_case(*_run_case("config2_32views_518x392_multimodal", views, synthetic_sd))


def test_fixed_scene_64_views_518x392_image_only(synthetic_sd):
    """The `fixed_scene` leg of every bench line (64 views, one scene: global attention over 66 305 tokens,
    519 K/V tiles per query tile, the two-tile attention kernel) against the fp32 oracle on the same GPU."""
    from oracle import weights

    free, _ = torch.cuda.mem_get_info()
    if free < 100 << 30:
        pytest.skip("the fp32 oracle needs ~60 GB for 64 views")
    views = weights.synthetic_views(64, 392, 518, seed=1)
    _assert_case(*_run_case("fixed_scene_64views_518x392_image_only", views, synthetic_sd))


def test_unit_gain_checkpoint_tracks_the_reference_bf16_path(monkeypatch):
    """The parity checkpoints damp every residual branch (synthetic.BRANCH_GAIN = 0.1, as trained checkpoints'
    LayerScale does) so that BASELINE.json's absolute tolerances mean something.  With the reference's OWN
    initialiser scale (LayerScale gamma = 1, un-damped xavier branches) the 48-block trunk
    amplifies bf16 rounding to ~1e-2 at the trunk outputs and > 1 deg at the pose in the reference's own
    autocast path, so no absolute tolerance holds for either implementation.  What must still hold, and what
    this records in the parity report: the CUDA path is no further from fp32 than a small multiple (5x, the
    two paths carry independent noise draws) of the reference's own bf16 path, key by key."""
    from mapanything_b200 import MapAnything, synthetic
    from oracle import mapanything_oracle as orc
    from tests import helpers

    monkeypatch.setattr(synthetic, "BRANCH_GAIN", 1.0)
    sd = synthetic.synthetic_state_dict()
    monkeypatch.undo()
    views = synthetic.synthetic_views(4, 392, 518, seed=1)
    m = MapAnything(device="cuda:0", head_chunk=4)
    m.load_state_dict(sd)
    sd_cuda = {k: v.cuda() for k, v in sd.items()}
    preds = m.infer(_to_cuda(views))
    torch.cuda.synchronize()
    with helpers.strict_fp32():
        ref = orc.infer(sd_cuda, _to_cuda(views))
    amp = orc.infer(sd_cuda, _to_cuda(views), amp=True)
    entry = helpers.record_parity("unit_gain_checkpoint_4views_518x392", helpers.compare_pred_lists(preds, ref),
                                  helpers.compare_pred_lists(amp, ref),
                                  extra={"views": 4, "H": 392, "W": 518, "branch_gain": 1.0})
    print("unit gain cuda     :", {k: round(v, 5) for k, v in entry["cuda"].items()})
    print("unit gain ref bf16 :", {k: round(v, 5) for k, v in entry["ref_bf16"].items()})
    bad = []
    for k, v in entry["cuda"].items():
        y = entry["ref_bf16"][k]
        if k.endswith("_agree"):
            if (1.0 - v) > max(1.0 - helpers.MIN_MASK_AGREE, 5.0 * (1.0 - y)):
                bad.append((k, v, y))
        elif v > max(TOL_POSE_DEG if k.endswith("_deg") else TOL_DENSE, 5.0 * y):
            bad.append((k, v, y))
    assert not bad, bad
    del m
    torch.cuda.empty_cache()
