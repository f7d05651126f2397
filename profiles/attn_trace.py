"""Debug: per-iteration timeline of the softmax warps (library built with -DMAB_ATT_TRACE)."""
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
all64 = trace.view(torch.int64)
tr = all64[:2 * 16 * 8].view(2, 16, 8).cpu()
mt = all64[256:256 + 2 * 16 * 8].view(2, 16, 8).cpu()
base = int(tr[0, 0, 0])
names = ["loop_top", "s_full", "S_in_regs", "exp_done", "pv_done_seen", "p_arrive"]
for t in range(2):
    print(f"tile {t}: columns = {names} (cycles since tile0 iter8 loop_top); deltas in ()")
    for j in range(16):
        row = [int(tr[t, j, k]) - base for k in range(6)]
        d = [row[0]] + [row[k] - row[k - 1] for k in range(1, 6)]
        print(f"  j={j+8:2d} " + " ".join(f"{row[k]:7d}({d[k]:5d})" for k in range(6)))

print("MMA warp: [top, s_free seen, QK issued, p_full seen, PV issued]")
for j in range(2, 8):
    for t in range(2):
        row = [int(mt[t, j, k]) - base for k in range(5)]
        print(f"  j={j+8:2d} t={t} " + " ".join(f"{x:7d}" for x in row))
