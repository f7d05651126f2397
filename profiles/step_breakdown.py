"""Where one step goes, operator by operator: runs the bench workload (N views 518x392, default 8) with the
library's per-launch CUDA-event profiler (mab_prof_enable / mab_prof_records) over K steps and groups the
launches by (category, algorithmic FLOPs | bytes) = operator shape, in order of first appearance.

    python profiles/step_breakdown.py [views] [steps] > profiles/r2_step_breakdown_n8.json
"""
import ctypes as C
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mapanything_b200 import MapAnything, synthetic  # noqa: E402

CATS = ["gemm", "conv3", "conv3_regress", "attention", "layernorm", "elementwise", "postprocess", "gemm_tf32",
        "conv3_tf32"]


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
    prec = sys.argv[3] if len(sys.argv) > 3 else "tf32"
    dev = torch.device("cuda", 0)
    model = MapAnything(device=dev, head_chunk=8, head_precision=prec)
    model.load_state_dict(synthetic.synthetic_state_dict())
    views = synthetic.synthetic_views(min(n, 16), 392, 518, seed=1)
    imgs = torch.cat([views[i % len(views)]["img"] for i in range(n)], 0).to(dev)
    lib = model._lib
    for _ in range(3):
        model.run_resident(imgs, 392, 518)
    lib.mab_prof_enable(1)
    for _ in range(steps):
        model.run_resident(imgs, 392, 518)
    total = lib.mab_prof_records(None, None, None, None, 0)
    cat = (C.c_int * total)(); ms = (C.c_double * total)(); fl = (C.c_double * total)(); by = (C.c_double * total)()
    assert lib.mab_prof_records(cat, ms, fl, by, total) == total
    lib.mab_prof_enable(0)
    groups, order = {}, []
    for i in range(total):
        key = (CATS[cat[i]], fl[i], by[i])
        if key not in groups:
            groups[key] = [0, 0.0]
            order.append(key)
        groups[key][0] += 1
        groups[key][1] += ms[i]
    rows, tot = [], 0.0
    for key in order:
        cnt, t = groups[key]
        tot += t / steps
        rows.append({"category": key[0], "gflop_per_launch": round(key[1] / 1e9, 3) if key[1] else None,
                     "mb_per_launch": round(key[2] / 1e6, 2) if key[2] else None, "launches_per_step": cnt / steps,
                     "us_per_launch": round(1e3 * t / cnt, 2), "ms_per_step": round(t / steps, 4),
                     "tflops": round(key[1] / (t / cnt * 1e-3) / 1e12, 1) if key[1] else None,
                     "gbs": round(key[2] / (t / cnt * 1e-3) / 1e9, 1) if key[2] else None})
    print(json.dumps({"views": n, "steps": steps, "head_precision": prec, "kernel_ms_per_step": round(tot, 3),
                      "how": "library-recorded CUDA events around every launch on the launch stream (warm L2, in step order)",
                      "operators": rows}, indent=1))


if __name__ == "__main__":
    main()
