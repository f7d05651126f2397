"""Per-operator A/B on the same box: the library kernels the reference calls on this path
(F.scaled_dot_product_attention at transformer_blocks.py:198-201 / encoders/dinov2.py:136, cuBLASLt
nn.Linear at transformer_blocks.py:75-78,146-154) against mab_attention / mab_gemm on the exact shapes
of the 8-view and 64-view steps.  CUDA events, L2 flushed between repetitions, best and median of 7.

    python profiles/library_ab.py > profiles/r2_library_ab.json
"""
import ctypes as C
import json
import os
import sys

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mapanything_b200 import _lib  # noqa: E402

lib = _lib.load()
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def timeit(fn, flop, reps=7):
    for _ in range(2):
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
    ts.sort()
    b, m = ts[0], ts[len(ts) // 2]
    return {"best_ms": b, "median_ms": m, "tflops_best": flop / b / 1e9, "tflops_median": flop / m / 1e9}


def attention_case(seqs, heads, length):
    from torch.nn.attention import SDPBackend, sdpa_kernel

    D = heads * 64
    rows = seqs * length
    qkv = torch.randn(rows, 3 * D, device="cuda").bfloat16()
    out = torch.empty(rows, D, device="cuda", dtype=torch.bfloat16)
    fl = 4.0 * length * length * 64 * heads * seqs

    def ours():
        _lib.check(lib.mab_attention(ptr(qkv), rows, ptr(qkv), rows, heads, seqs, length, length, length, length,
                                     0.125, ptr(out), None, st))

    # the reference's layout: qkv.reshape(B, N, 3, H, hd).permute(2, 0, 3, 1, 4) -> (B, H, N, hd) views
    t = qkv.view(seqs, length, 3, heads, 64).permute(2, 0, 3, 1, 4)
    q, k, v = t[0], t[1], t[2]
    qc, kc, vc = q.contiguous(), k.contiguous(), v.contiguous()
    res = {"shape": [seqs, heads, length, 64], "flop": fl, "mab_attention": timeit(ours, fl)}
    for name, backends, args in (("sdpa_default_strided_views", None, (q, k, v)),
                                 ("sdpa_default_contiguous", None, (qc, kc, vc)),
                                 ("sdpa_flash", [SDPBackend.FLASH_ATTENTION], (qc, kc, vc)),
                                 ("sdpa_cudnn", [SDPBackend.CUDNN_ATTENTION], (qc, kc, vc)),
                                 ("sdpa_efficient", [SDPBackend.EFFICIENT_ATTENTION], (qc, kc, vc))):
        try:
            fn = lambda: F.scaled_dot_product_attention(*args)  # noqa: E731
            if backends is None:
                res[name] = timeit(fn, fl)
            else:
                with sdpa_kernel(backends):
                    res[name] = timeit(fn, fl)
        except Exception as e:  # a backend that refuses the shape
            res[name] = {"error": str(e).splitlines()[0][:200]}
    return res


def gemm_case(name, M, N, K, act, residual):
    A = torch.randn(M, K, device="cuda").bfloat16()
    B = (torch.randn(N, K, device="cuda") * 0.03).bfloat16()
    bias = torch.randn(N, device="cuda")
    bias_bf = bias.bfloat16()
    ob = None if residual else torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    x = torch.randn(M, N, device="cuda") if residual else None
    gamma = torch.ones(N, device="cuda") if residual else None
    fl = 2.0 * M * N * K

    def ours():
        _lib.check(lib.mab_gemm(ptr(A), ptr(B), M, N, K, ptr(bias), ptr(gamma), ptr(x), act, ptr(x), ptr(ob), st))

    def lib_linear():
        return F.linear(A, B, bias_bf)

    def lib_block():  # what the reference's eager block runs around the GEMM (autocast: bf16 linear, fp32 stream)
        y = F.linear(A, B, bias_bf)
        if act:
            y = F.gelu(y)
        return x + gamma * y.float() if residual else y

    return {"name": name, "M": M, "N": N, "K": K, "flop": fl,
            "epilogue": ("bias+GELU" if act else "bias") + ("+LayerScale+fp32 residual" if residual else ""),
            "mab_gemm_fused_epilogue": timeit(ours, fl), "cublaslt_linear_bias_only": timeit(lib_linear, fl),
            "eager_linear_plus_epilogue_ops": timeit(lib_block, fl)}


def main():
    out = {"gpu": torch.cuda.get_device_name(0), "torch": torch.__version__,
           "how": "CUDA events on the current stream, 256 MB L2 flush before every repetition, best / median of 7",
           "attention": [], "gemm": []}
    for seqs, heads, length in ((8, 16, 1037), (8, 12, 1036), (1, 12, 8289), (1, 12, 66305)):
        out["attention"].append(attention_case(seqs, heads, length))
    for D, T in ((1024, 1037), (768, 1036)):
        M = 8 * T + (0 if D == 1024 else 1)
        tag = "enc" if D == 1024 else "aat"
        out["gemm"].append(gemm_case(f"{tag} qkv", M, 3 * D, D, 0, False))
        out["gemm"].append(gemm_case(f"{tag} proj", M, D, D, 0, True))
        out["gemm"].append(gemm_case(f"{tag} fc1", M, 4 * D, D, 1, False))
        out["gemm"].append(gemm_case(f"{tag} fc2", M, D, 4 * D, 0, True))
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
