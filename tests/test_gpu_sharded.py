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


@pytest.mark.parametrize("n_views,H,W,mode", [(5, 70, 98, "img"), (4, 392, 518, "img"), (5, 70, 98, "mm"), (5, 70, 98, "viewpe"),
                                                 (48, 392, 518, "img")])
def test_sharded_equals_single(n_views, H, W, mode):
    ngpu = torch.cuda.device_count()
    if ngpu < 2:
        pytest.skip("needs >= 2 GPUs")
    world = 2 if n_views < 8 else min(ngpu, 4)  # 48 views: 24 or 12 per rank, the two-tile attention kernel
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", "29731",
           os.path.join(ROOT, "tests", "dist_sharded_check.py"), str(n_views), str(H), str(W), mode]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    print(res.stdout[-2000:], res.stderr[-3000:])
    assert res.returncode == 0


def test_forward_sharded_with_one_rank_equals_forward(synthetic_sd):
    """mab_set_comm + mab_forward_sharded on ONE GPU (a one-rank NCCL communicator made with ctypes, no process
    group): the one-call sharded forward must reproduce mab_forward bit for bit when the shard is the whole
    scene -- the entry points the multi-GPU runs use, exercised where the driver's 1-GPU box can see them."""
    import ctypes as C

    from mapanything_b200 import MapAnything, _lib, nccl_comm
    from oracle import weights

    torch.cuda.set_device(0)
    m = MapAnything(device="cuda:0", head_chunk=2)
    m.load_state_dict(synthetic_sd)
    lib = m._lib
    H, W, V = 56, 84, 3
    imgs = torch.cat([v["img"] for v in weights.synthetic_views(V, H, W, seed=61)], 0).cuda()
    want = m._run_forward(imgs, H, W)                      # mab_forward
    torch.cuda.synchronize()

    path = nccl_comm.nccl_library_path()
    nccl = C.CDLL(path)
    uid = nccl_comm._UniqueId()
    assert nccl.ncclGetUniqueId(C.byref(uid)) == 0
    comm = C.c_void_p()
    nccl.ncclCommInitRank.argtypes = [C.POINTER(C.c_void_p), C.c_int, nccl_comm._UniqueId, C.c_int]
    assert nccl.ncclCommInitRank(C.byref(comm), 1, uid, 0) == 0
    _lib.check(lib.mab_set_comm(m._handle, comm, 0, 1, path.encode()), "mab_set_comm")
    # argument checks: a shard that does not match the table is refused before any launch
    outs = m._alloc_outputs(V, H, W)
    inp, ost = m._input_struct(imgs, H, W), m._out_struct(outs)
    ws = m._get_workspace(lib.mab_workspace_bytes(m._handle, V, V, H, W))
    kv = torch.empty(lib.mab_kv_gather_bytes(m._handle, V, H, W), dtype=torch.uint8, device="cuda")
    st = m._stream_ptr()
    assert lib.mab_forward_sharded(m._handle, C.byref(inp), V + 1, C.byref(ost), ws.data_ptr(), ws.numel(),
                                   kv.data_ptr(), kv.numel(), 1, st) == -1
    _lib.check(lib.mab_forward_sharded(m._handle, C.byref(inp), V, C.byref(ost), ws.data_ptr(), ws.numel(),
                                       kv.data_ptr(), kv.numel(), 1, st), "mab_forward_sharded")
    torch.cuda.synchronize()
    for k, t in want.items():
        assert torch.equal(outs[k], t), k
    nccl.ncclCommDestroy.argtypes = [C.c_void_p]
    nccl.ncclCommDestroy(comm)
