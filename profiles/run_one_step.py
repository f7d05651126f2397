"""One profiled step of the bench workload (N views 518x392, default 8): model setup and three warm-up
steps run outside the profiler range; use with `ncu --profile-from-start off ...`."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mapanything_b200 import MapAnything, synthetic  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    model = MapAnything(device=dev, head_chunk=8)
    model.load_state_dict(synthetic.synthetic_state_dict())
    views = synthetic.synthetic_views(min(n, 16), 392, 518, seed=1)
    imgs = torch.cat([views[i % len(views)]["img"] for i in range(n)], 0).to(dev)
    for _ in range(3):
        model.run_resident(imgs, 392, 518)
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStart()
    model.run_resident(imgs, 392, 518)
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStop()
    print("done")


if __name__ == "__main__":
    main()
