"""Debug: per-iteration timeline of one softmax warp per tile (library built with -DMAB_ATT_TRACE and
trace points inserted by the developer; not part of the product build)."""
import ctypes as C, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mapanything_b200 import _lib
lib = _lib.load()
heads, length = 12, int(sys.argv[1]) if len(sys.argv) > 1 else 8289
D = heads * 64
qkv = torch.randn(length, 3 * D, device="cuda").bfloat16()
out = torch.empty(length, D, device="cuda", dtype=torch.bfloat16)
trace = torch.zeros(length * heads, dtype=torch.float32, device="cuda")
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
p = lambda t: C.c_void_p(t.data_ptr())
for _ in range(2):
    trace.zero_()
    _lib.check(lib.mab_attention(p(qkv), length, p(qkv), length, heads, 1, length, length, length, length, 0.125, p(out), p(trace), st))
torch.cuda.synchronize()
tr = trace.view(torch.int64)[:2 * 16 * 8].view(2, 16, 8).cpu()
base = int(tr[0, 0, 0])
names = ["top", "s_full", "S_regs", "turn", "exp_done", "pv_done", "p_arrive"]
for t in range(2):
    print(f"tile {t}: {names}; absolute(delta)")
    for j in range(2, 6):
        row = [int(tr[t, j, k]) - base for k in range(7)]
        d = [0] + [row[k] - row[k - 1] for k in range(1, 7)]
        print(f"  j={j+8:2d} " + " ".join(f"{row[k]:6d}({d[k]:4d})" for k in range(7)))
