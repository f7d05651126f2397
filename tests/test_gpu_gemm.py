"""tcgen05 GEMM vs fp32 matmul of the same bf16-rounded operands (parity gate of SURVEY 7.2 A)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref(A, B, bias, gamma, res, act):
    y = A.float() @ B.float().t()
    if bias is not None:
        y = y + bias
    if act == 1:
        y = torch.nn.functional.gelu(y)
    elif act == 2:
        y = torch.relu(y)
    if gamma is not None:
        y = y * gamma
    if res is not None:
        y = y + res
    return y


@pytest.mark.parametrize("M,N,K,act,use_res", [
    (128, 256, 64, 0, False),      # one tile, one k-block
    (128, 256, 128, 0, False),     # two k-blocks
    (300, 768, 1024, 0, True),     # M tail, LayerScale + fp32 residual (attn.proj / fc2)
    (1037, 3072, 1024, 0, False),  # encoder qkv
    (2073, 2304, 768, 0, False),   # info-sharing qkv, odd row count (N*T+1)
    (777, 3072, 768, 1, False),    # fc1 + exact GELU
    (500, 768, 3072, 0, True),     # fc2
    (1036, 1024, 640, 0, False),   # patch embed (K padded 588 -> 640)
    (8289, 768, 768, 0, True),     # info-sharing proj at 8 views: 192-wide tiles, two waves
    (8289, 2304, 768, 0, False),   # info-sharing qkv at 8 views: 192-wide tiles
    (1036, 96, 1024, 0, False),    # DPT 1x1 to 96 channels (N < tile)
    (1036, 1536, 96, 0, False),    # convT as GEMM, K tail (96 = 1.5 k-blocks)
    (1036, 784, 768, 2, False),    # pose head proj + ReLU, N tail (784 = 3*256 + 16)
    (1036, 784, 784, 2, False),    # pose head res conv, K and N tails
    (266, 768, 6912, 0, False),    # 3x3/s2 conv as im2col GEMM
    (40000, 256, 256, 0, False),   # many tiles per CTA (persistent loop, TMEM double buffering)
])
def test_gemm(M, N, K, act, use_res):
    from tests import gpu_helpers as g

    torch.manual_seed(M * 7 + N + K)
    A = (torch.randn(M, K, device="cuda") * 0.5).bfloat16()
    B = (torch.randn(N, K, device="cuda") / K ** 0.5).bfloat16()
    bias = torch.randn(N, device="cuda") * 0.1
    gamma = 1 + 0.1 * torch.randn(N, device="cuda") if use_res else None
    res = torch.randn(M, N, device="cuda") if use_res else None
    of, ob = g.gemm(A, B, bias, gamma, res, act, out_f32=True, out_bf16=True)
    ref = _ref(A, B, bias, gamma, res, act)
    e32 = g.rel(of, ref)
    e16 = g.rel(ob.float(), ref)
    # GELU is evaluated as x*sigmoid(2u), erf ~ tanh(u) (|error| <= 2.6e-5 absolute, gemm.cuh)
    assert e32 < (1e-4 if act == 1 else 2e-5), f"fp32 output rel-L2 {e32}"
    assert e16 < 4e-3, f"bf16 output rel-L2 {e16}"
    # in-place residual update (res aliases out) as used for the residual stream
    if use_res:
        x = res.clone()
        import ctypes as C
        g.check(g.lib().mab_gemm(g.ptr(A), g.ptr(B), M, N, K, g.ptr(bias), g.ptr(gamma), g.ptr(x), act,
                                 g.ptr(x), None, g.stream()))
        torch.cuda.synchronize()
        assert g.rel(x, ref) < 2e-5


def test_layernorm():
    from tests import gpu_helpers as g

    for C_ in (768, 1024):
        x = torch.randn(1037, C_, device="cuda") * 3 + 0.5
        w = torch.randn(C_, device="cuda")
        b = torch.randn(C_, device="cuda")
        ob = torch.empty((1037, C_), dtype=torch.bfloat16, device="cuda")
        of = torch.empty((1037, C_), dtype=torch.float32, device="cuda")
        g.check(g.lib().mab_layernorm(g.ptr(x), 1037, C_, g.ptr(w), g.ptr(b), 1e-6, g.ptr(ob), g.ptr(of), g.stream()))
        torch.cuda.synchronize()
        ref = torch.nn.functional.layer_norm(x, (C_,), w, b, 1e-6)
        assert g.rel(of, ref) < 1e-6
        assert g.rel(ob.float(), ref) < 3e-3
