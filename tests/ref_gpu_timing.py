"""TEST INFRASTRUCTURE -- SURVEY.md 8(d) "reference GPU baseline": the oracle (the reference's
algorithm in eager PyTorch: library SDPA / cuBLASLt / cuDNN) on the B200 with the reference's
precision regions (bf16 autocast trunk, fp32/TF32 heads), same synthetic weights and inputs, timed
with the protocol of bench.py.  Prints one JSON line; its output is kept under profiles/.

    python -m tests.ref_gpu_timing [views] [steps]
"""
import json
import sys
import time

import torch

from oracle import mapanything_oracle as orc
from oracle import weights


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
    sd = {k: v.cuda() for k, v in weights.synthetic_state_dict().items()}
    views = [{k: (v.cuda() if torch.is_tensor(v) else v) for k, v in view.items()}
             for view in weights.synthetic_views(n, 392, 518, seed=1)]
    pv = orc.preprocess_views(views)

    def step():
        with torch.inference_mode():
            return orc.forward(sd, pv, amp=True)

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    torch.cuda.reset_peak_memory_stats()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    t0 = time.perf_counter()
    with torch.inference_mode():
        orc.infer(sd, views, amp=True)
    torch.cuda.synchronize()
    infer_ms = (time.perf_counter() - t0) * 1e3
    print(json.dumps({"what": "oracle forward (eager PyTorch, bf16 autocast trunk + TF32 heads) on this GPU",
                      "views": n, "ms_per_forward": ms, "views_per_s_forward": n / (ms * 1e-3),
                      "infer_ms_incl_cpu_postprocess": infer_ms, "views_per_s_infer": n / (infer_ms * 1e-3),
                      "peak_mem_gb": torch.cuda.max_memory_allocated() / 2 ** 30,
                      "gpu": torch.cuda.get_device_name(0), "torch": torch.__version__}))


if __name__ == "__main__":
    main()
