"""Thin torch <-> C-ABI glue for the `-m gpu` parity tests (calls go through include/mab200.h)."""
import ctypes as C

import torch

from mapanything_b200 import _lib


def lib():
    return _lib.load()


def ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def check(st, what=""):
    _lib.check(st, what)


def rel(a, b):
    a, b = a.double(), b.double()
    return float((a - b).norm() / (b.norm() + 1e-30))


def gemm(A, B, bias=None, gamma=None, res=None, act=0, out_f32=False, out_bf16=True):
    """A [M,K] bf16, B [N,K] bf16 -> dict(f32=?, bf16=?)"""
    M, K = A.shape
    N = B.shape[0]
    of = torch.empty((M, N), dtype=torch.float32, device="cuda") if out_f32 else None
    ob = torch.empty((M, N), dtype=torch.bfloat16, device="cuda") if out_bf16 else None
    check(lib().mab_gemm(ptr(A), ptr(B), M, N, K, ptr(bias), ptr(gamma), ptr(res), act, ptr(of),
                         ptr(ob), stream()), "mab_gemm")
    torch.cuda.synchronize()
    return of, ob


def attention(qkv, heads, num_seqs, q_len, kv_len, q_stride, kv_stride, kv=None, scale=0.125, lse=False):
    D = heads * 64
    kvt = qkv if kv is None else kv
    out = torch.zeros((qkv.shape[0], D), dtype=torch.bfloat16, device="cuda")
    l = torch.zeros((qkv.shape[0], heads), dtype=torch.float32, device="cuda") if lse else None
    check(lib().mab_attention(ptr(qkv), qkv.shape[0], ptr(kvt), kvt.shape[0], heads, num_seqs, q_len,
                              kv_len, q_stride, kv_stride, scale, ptr(out), ptr(l), stream()),
          "mab_attention")
    torch.cuda.synchronize()
    return (out, l) if lse else out


def conv3x3(x_nhwc, w_oihw, bias=None, res1=None, res2=None, act=0, want_relu=False):
    """x [B,H,W,C] bf16, w [Cout,Cin,3,3] fp32"""
    Bn, H, W, Cc = x_nhwc.shape
    Cout, Cin = w_oihw.shape[:2]
    Cpad = (Cin + 63) // 64 * 64
    wp = torch.empty((Cout, 9 * Cpad), dtype=torch.bfloat16, device="cuda")
    check(lib().mab_pack_conv3_weight(ptr(w_oihw.contiguous()), Cout, Cin, ptr(wp), stream()), "pack")
    out = torch.empty((Bn, H, W, Cout), dtype=torch.bfloat16, device="cuda")
    outr = torch.empty_like(out) if want_relu else None
    check(lib().mab_conv3x3(ptr(x_nhwc), Bn, H, W, Cc, ptr(wp), Cout, ptr(bias), ptr(res1), ptr(res2),
                            act, ptr(out), ptr(outr), stream()), "mab_conv3x3")
    torch.cuda.synchronize()
    return (out, outr) if want_relu else out
