"""TEST INFRASTRUCTURE ONLY -- run in the build container (needs /root/reference):

    python -m oracle.check_batch_independence

Shows on the LIVE reference that the samples of a per-view batch (img (B,3,H,W), B > 1) are independent
scenes: `infer` on the batched views equals the per-scene `infer` results concatenated along dim 0 (fp32
noise only; metric_scaling_factor (B,1)).  mapanything_b200.MapAnything.infer relies on this to run a
batch as B scenes (tests/test_gpu_e2e.py::test_batched_views_are_independent_scenes).
Last run: max |diff| 2.3e-6 on pts3d, 5.3e-5 on intrinsics, 0 differing mask pixels.
"""
import sys, torch
sys.path.insert(0, __import__('os').path.dirname(__import__('os').path.dirname(__import__('os').path.abspath(__file__))))
from oracle import ref_shim, weights
sd = weights.synthetic_state_dict()
model = ref_shim.build_reference_model(sd)
va = weights.synthetic_views(2, 42, 56, seed=1)
vb = weights.synthetic_views(2, 42, 56, seed=5)
batched = [{"img": torch.cat([a["img"], b["img"]], 0), "data_norm_type": ["dinov2"]} for a, b in zip(va, vb)]
pa = model.infer([dict(v) for v in va], use_amp=False)
pb = model.infer([dict(v) for v in vb], use_amp=False)
pp = model.infer(batched, use_amp=False)
for k in pp[0]:
    for i in range(2):
        ref = torch.cat([pa[i][k], pb[i][k]], 0)
        got = pp[i][k]
        if got.shape != ref.shape: print(k, i, "shape", got.shape, ref.shape); continue
        d = (got.float()-ref.float()).abs().max().item() if got.dtype!=torch.bool else (got!=ref).sum().item()
        print(k, i, tuple(got.shape), d)
