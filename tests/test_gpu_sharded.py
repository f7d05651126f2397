"""View-sharded infer (global attention by K|V gather AND by ring, over NCCL) == single-GPU infer.  Needs >= 2 GPUs
(`gpurun --gpus 2`); skipped on a 1-GPU box, where tests/test_gpu_attention.py::test_attention_ring_carry
and the gloo tests of tests/test_parallel_cpu.py cover the same pieces separately."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("n_views,H,W,mode", [(5, 70, 98, "img"), (4, 392, 518, "img"), (5, 70, 98, "mm"), (5, 70, 98, "viewpe")])
def test_sharded_equals_single(n_views, H, W, mode):
    ngpu = torch.cuda.device_count()
    if ngpu < 2:
        pytest.skip("needs >= 2 GPUs")
    world = 2 if n_views < 8 else min(ngpu, 4)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", "29731",
           os.path.join(ROOT, "tests", "dist_sharded_check.py"), str(n_views), str(H), str(W), mode]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    print(res.stdout[-2000:], res.stderr[-3000:])
    assert res.returncode == 0
