"""Micro-benchmark of the linear-layer shapes of one step (through the C ABI, CUDA events).

    python profiles/gemm_shapes.py [views]

Prints TFLOP/s per (shape, epilogue) for the encoder (D=1024) and info-sharing (D=768) blocks.
"""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mapanything_b200 import _lib  # noqa: E402

lib = _lib.load()
views = int(sys.argv[1]) if len(sys.argv) > 1 else 8
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)


def p(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def run(name, M, N, K, act, residual):
    A = torch.randn(M, K, device="cuda").bfloat16()
    B = (torch.randn(N, K, device="cuda") * 0.03).bfloat16()
    bias = torch.randn(N, device="cuda")
    ob = None if residual else torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    x = torch.randn(M, N, device="cuda") if residual else None
    gamma = torch.ones(N, device="cuda") if residual else None

    def call():
        _lib.check(lib.mab_gemm(p(A), p(B), M, N, K, p(bias), p(gamma), p(x), act, p(x), p(ob), st))

    for _ in range(3):
        call()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    iters = 20
    e0.record()
    for _ in range(iters):
        call()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    print(f"{name:28s} M={M:6d} N={N:5d} K={K:5d}  {ms * 1e3:8.1f} us  {2.0 * M * N * K / ms / 1e9:7.1f} TFLOP/s")


for D, T in ((1024, 1037), (768, 1036)):
    M = views * T + (0 if D == 1024 else 1)
    tag = "enc" if D == 1024 else "aat"
    run(f"{tag} qkv", M, 3 * D, D, 0, False)
    run(f"{tag} proj (+res, gamma)", M, D, D, 0, True)
    run(f"{tag} fc1 (gelu)", M, 4 * D, D, 1, False)
    run(f"{tag} fc2 (+res, gamma)", M, D, 4 * D, 0, True)
