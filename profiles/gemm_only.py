import ctypes as C, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mapanything_b200 import _lib
lib = _lib.load()
M, N, K = (int(a) for a in sys.argv[1:4]) if len(sys.argv) > 3 else (66305, 3072, 1024)
act = int(sys.argv[4]) if len(sys.argv) > 4 else 0
A = torch.randn(M, K, device="cuda").bfloat16(); B = torch.randn(N, K, device="cuda").bfloat16()
bias = torch.randn(N, device="cuda"); ob = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
p = lambda t: C.c_void_p(t.data_ptr())
for _ in range(3):
    _lib.check(lib.mab_gemm(p(A), p(B), M, N, K, p(bias), None, None, act, None, p(ob), st))
torch.cuda.synchronize()
print("done")
