"""Attention / GEMM kernel microbenchmarks through the C ABI (CUDA events, L2 flushed between reps)."""
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mapanything_b200 import _lib  # noqa: E402

lib = _lib.load()


def ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def timeit(fn, reps=5):
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return min(ts), sorted(ts)[len(ts) // 2]


def attn(heads, seqs, length):
    D = heads * 64
    rows = seqs * length
    qkv = torch.randn(rows, 3 * D, device="cuda").bfloat16()
    out = torch.empty(rows, D, device="cuda", dtype=torch.bfloat16)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def fn():
        _lib.check(lib.mab_attention(ptr(qkv), rows, ptr(qkv), rows, heads, seqs, length, length, length, length,
                                     0.125, ptr(out), None, st))
    best, med = timeit(fn)
    fl = 4.0 * length * length * 64 * heads * seqs
    print(f"attention heads={heads} seqs={seqs} len={length}: best {best:.3f} ms med {med:.3f} ms  {fl / best / 1e9:.1f} TFLOP/s")


def gemm(M, N, K, act=0, res=False):
    A = torch.randn(M, K, device="cuda").bfloat16()
    B = torch.randn(N, K, device="cuda").bfloat16()
    bias = torch.randn(N, device="cuda")
    of = torch.randn(M, N, device="cuda") if res else None
    ob = None if res else torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def fn():
        _lib.check(lib.mab_gemm(ptr(A), ptr(B), M, N, K, ptr(bias), None, ptr(of), act, ptr(of), ptr(ob), st))
    best, med = timeit(fn)
    print(f"gemm M={M} N={N} K={K} act={act} res={res}: best {best:.3f} ms  {2.0 * M * N * K / best / 1e9:.1f} TFLOP/s")


if __name__ == "__main__":
    attn(16, 8, 1037)      # encoder, 8 views
    attn(12, 8, 1036)      # frame, 8 views
    attn(12, 1, 8289)      # global, 8 views
    attn(12, 1, 66305)     # global, 64 views
    for M in (8296, 66305):
        gemm(M, 3072, 1024)
        gemm(M, 1024, 1024, res=True)
        gemm(M, 4096, 1024, act=1)
        gemm(M, 1024, 4096, res=True)
        gemm(M, 2304, 768)
        gemm(M, 768, 768, res=True)
        gemm(M, 3072, 768, act=1)
        gemm(M, 768, 3072, res=True)
