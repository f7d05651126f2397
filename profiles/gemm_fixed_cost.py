"""Fixed (K-independent) cost of one GEMM launch: time back-to-back launches for a K sweep."""
import ctypes as C, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mapanything_b200 import _lib
lib = _lib.load()
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
p = lambda t: C.c_void_p(t.data_ptr()) if t is not None else None
def run(M, N, K, residual, act=0):
    A = torch.randn(M, K, device="cuda").bfloat16(); B = (torch.randn(N, K, device="cuda") * .03).bfloat16()
    bias = torch.randn(N, device="cuda")
    ob = None if residual else torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    x = torch.randn(M, N, device="cuda") if residual else None
    g = torch.ones(N, device="cuda") if residual else None
    call = lambda: _lib.check(lib.mab_gemm(p(A), p(B), M, N, K, p(bias), p(g), p(x), act, p(x), p(ob), st))
    for _ in range(3): call()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(50): call()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / 50 * 1e3
    print(f"M={M:6d} N={N:5d} K={K:5d} res={int(residual)} act={act}: {us:7.1f} us")
for res in (False, True):
    for (M, N) in ((256, 256), (8296, 1024), (8296, 3072)):
        for K in (64, 256, 1024, 4096):
            run(M, N, K, res)
