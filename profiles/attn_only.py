import ctypes as C, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mapanything_b200 import _lib
lib = _lib.load()
heads, length = 12, int(sys.argv[1]) if len(sys.argv) > 1 else 16577
D = heads * 64
qkv = torch.randn(length, 3 * D, device="cuda").bfloat16()
out = torch.empty(length, D, device="cuda", dtype=torch.bfloat16)
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
for _ in range(3):
    _lib.check(lib.mab_attention(C.c_void_p(qkv.data_ptr()), length, C.c_void_p(qkv.data_ptr()), length, heads, 1, length, length,
                                 length, length, 0.125, C.c_void_p(out.data_ptr()), None, st))
torch.cuda.synchronize()
print("done")
