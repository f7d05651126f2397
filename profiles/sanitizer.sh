#!/bin/bash
# compute-sanitizer over a tiny forward + post-processing of the whole hot path (2 views 56x84, image-only
# and multi-modal, both head precisions): memcheck, synccheck, racecheck.  The kernels hand-roll mbarrier,
# named-barrier, cluster and TMEM protocols (SURVEY.md section 5 "Race detection").
#   bash profiles/sanitizer.sh        (on the GPU box; writes gpurun_out/sanitizer_*.log)
set -u
mkdir -p gpurun_out
cat > /tmp/mab_sanitize_case.py <<'PY'
import sys, torch
sys.path.insert(0, ".")
from mapanything_b200 import MapAnything, synthetic
sd = synthetic.synthetic_state_dict()
views = synthetic.synthetic_views(2, 56, 84, seed=1)
mm = synthetic.add_multimodal_inputs(synthetic.synthetic_views(2, 56, 84, seed=3), seed=4)
for prec in ("tf32", "bf16"):
    m = MapAnything(device="cuda:0", head_chunk=1, enc_chunk=1, head_precision=prec)
    m.load_state_dict(sd)
    a = m.infer([dict(v) for v in views])
    b = m.infer([dict(v) for v in mm], apply_confidence_mask=True)
    torch.cuda.synchronize()
    assert all(bool(torch.isfinite(p["pts3d"]).all()) for p in a + b)
print("case ok")
PY
for tool in memcheck synccheck racecheck; do
  timeout 1500 compute-sanitizer --tool $tool --launch-timeout 0 --print-limit 20 \
    python /tmp/mab_sanitize_case.py > gpurun_out/sanitizer_$tool.log 2>&1
  echo "exit=$?" >> gpurun_out/sanitizer_$tool.log
  tail -n 4 gpurun_out/sanitizer_$tool.log
done
