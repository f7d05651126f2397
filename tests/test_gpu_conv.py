"""Implicit-GEMM 3x3 conv (TMA-shifted taps) and the DPT element-wise kernels vs torch fp32."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("B,H,W,Cin,Cout", [
    (1, 8, 16, 64, 256),     # exactly one tile
    (2, 14, 19, 768, 256),   # layer4_rn level (tile tails in both dims)
    (1, 28, 37, 384, 256),   # layer3_rn
    (1, 40, 50, 96, 256),    # Cin = 1.5 k-blocks (zero-padded taps)
    (1, 56, 74, 192, 256),
    (2, 60, 70, 256, 128),   # regressor conv1 (Cout 128 tile)
])
def test_conv3x3(B, H, W, Cin, Cout):
    from tests import gpu_helpers as g

    torch.manual_seed(H * W + Cin)
    x = torch.randn(B, H, W, Cin, device="cuda").bfloat16()
    w = torch.randn(Cout, Cin, 3, 3, device="cuda") / (9 * Cin) ** 0.5
    b = torch.randn(Cout, device="cuda") * 0.1
    out = g.conv3x3(x, w, b)
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), w.bfloat16().float(), b, padding=1).permute(0, 2, 3, 1)
    e = g.rel(out.float(), ref)
    assert e < 4e-3, f"conv rel-L2 {e}"


def test_conv3x3_rcu_epilogue():
    """bias + two bf16 residuals + dual (value, relu(value)) outputs, as the ResidualConvUnits use"""
    from tests import gpu_helpers as g

    torch.manual_seed(0)
    B, H, W, Cc = 1, 30, 41, 256
    x = torch.randn(B, H, W, Cc, device="cuda").bfloat16()
    r1 = torch.randn(B, H, W, Cc, device="cuda").bfloat16()
    r2 = torch.randn(B, H, W, Cc, device="cuda").bfloat16()
    w = torch.randn(Cc, Cc, 3, 3, device="cuda") / (9 * Cc) ** 0.5
    b = torch.randn(Cc, device="cuda") * 0.1
    out, outr = g.conv3x3(x, w, b, r1, r2, act=0, want_relu=True)
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), w.bfloat16().float(), b, padding=1).permute(0, 2, 3, 1)
    ref = ref + r1.float() + r2.float()
    assert g.rel(out.float(), ref) < 4e-3
    assert g.rel(outr.float(), torch.relu(ref)) < 4e-3
    out2 = g.conv3x3(x, w, b, act=2)
    ref2 = torch.relu(F.conv2d(x.float().permute(0, 3, 1, 2), w.bfloat16().float(), b, padding=1)).permute(0, 2, 3, 1)
    assert g.rel(out2.float(), ref2) < 4e-3


@pytest.mark.parametrize("Hi,Wi,Hv,Wv,Ho,Wo,Cc", [(14, 19, 28, 38, 28, 37, 256), (28, 37, 56, 74, 56, 74, 256),
                                                    (32, 40, 55, 71, 55, 71, 128)])
def test_resize_bilinear(Hi, Wi, Hv, Wv, Ho, Wo, Cc):
    from tests import gpu_helpers as g

    x = torch.randn(2, Hi, Wi, Cc, device="cuda").bfloat16()
    out = torch.empty(2, Ho, Wo, Cc, device="cuda", dtype=torch.bfloat16)
    outr = torch.empty_like(out)
    g.check(g.lib().mab_resize_bilinear(g.ptr(x), 2, Hi, Wi, Cc, Hv, Wv, Ho, Wo, g.ptr(out), g.ptr(outr), g.stream()))
    torch.cuda.synchronize()
    ref = F.interpolate(x.float().permute(0, 3, 1, 2), size=(Hv, Wv), mode="bilinear", align_corners=True)
    ref = ref[:, :, :Ho, :Wo].permute(0, 2, 3, 1)
    assert g.rel(out.float(), ref) < 3e-3
    assert g.rel(outr.float(), torch.relu(ref)) < 3e-3


def test_posembed_bicubic():
    from oracle import mapanything_oracle as orc
    from tests import gpu_helpers as g

    pos = torch.randn(1, 1 + 37 * 37, 1024, device="cuda") * 0.02
    for (h, w) in [(28, 37), (5, 7), (37, 37), (4, 6)]:
        out = torch.empty(1 + h * w, 1024, device="cuda")
        g.check(g.lib().mab_posembed_resize(g.ptr(pos), 37, 1024, h, w, g.ptr(out), g.stream()))
        torch.cuda.synchronize()
        ref = orc.interpolate_pos_embed(pos.cpu(), h, w)[0]
        e = g.rel(out.cpu(), ref)
        assert e < 1e-5, (h, w, e)


def test_edge_mask_matches_numpy():
    """mask kernels vs the oracle's numpy (the reference's CPU code path) on identical inputs"""
    import numpy as np
    from oracle import mapanything_oracle as orc
    from tests import gpu_helpers as g

    torch.manual_seed(1)
    V, H, W = 2, 61, 83
    yy, xx = torch.meshgrid(torch.linspace(-1, 1, H), torch.linspace(-1, 1, W), indexing="ij")
    total_diff = 0
    for v in range(V):
        z = 2 + 0.5 * torch.sin(3 * xx + v) * torch.cos(2 * yy) + 0.004 * torch.randn(H, W)
        z[20:30, 30:50] += 1.5  # a depth discontinuity
        pts_cam = torch.stack([xx * z, yy * z, z], -1)
        Rm = torch.linalg.qr(torch.randn(3, 3))[0]
        pts = pts_cam @ Rm.t() + torch.tensor([0.1, -0.2, 0.3])
        fm = torch.rand(H, W) > 0.1
        n, nm = orc.points_to_normals(pts.numpy(), fm.numpy())
        ne = orc.normals_edge(n, 5.0, nm)
        de = orc.depth_edge(pts_cam[..., 2].numpy(), 0.03, fm.numpy())
        ref = fm.numpy() & ~(de & ne)
        ws = torch.empty(H * W * 17 + 256, dtype=torch.uint8, device="cuda")
        out = torch.empty(1, H, W, dtype=torch.bool, device="cuda")
        pc = pts_cam.cuda().contiguous().view(1, H, W, 3)
        g.check(g.lib().mab_edge_mask(g.ptr(pts.cuda().contiguous()), g.ptr(pc), g.ptr(fm.cuda().contiguous()),
                                      1, H, W, 5.0, 0.03, g.ptr(ws), g.ptr(out), g.stream()))
        torch.cuda.synchronize()
        diff = int((out[0].cpu().numpy() != ref).sum())
        total_diff += diff
        assert 0.05 < ref.mean() < 0.95, "test scene should exercise both mask values"
    assert total_diff <= 2, f"{total_diff} mask pixels differ from numpy"


@pytest.mark.parametrize("M,N,K,act,post", [(300, 768, 640, 0, 0), (2072, 1024, 1024, 0, 0),
                                             (1037, 1024, 768, 1, 1), (128, 256, 32, 1, 0)])
def test_gemm_tf32(M, N, K, act, post):
    """fp32-operand GEMM on the TF32 tensor-core path (geometric-input encoders' 1x1 convs) vs torch
    fp32: TF32 keeps 10 mantissa bits, so the error budget is ~1e-3 of bf16's."""
    from tests import gpu_helpers as g

    torch.manual_seed(M + N)
    A = torch.randn(M, K, device="cuda")
    B = torch.randn(N, K, device="cuda") / K ** 0.5
    bias = torch.randn(N, device="cuda") * 0.1
    res = torch.randn(M, N, device="cuda")
    out = torch.empty(M, N, device="cuda")
    g.check(g.lib().mab_gemm_tf32(g.ptr(A), g.ptr(B), M, N, K, g.ptr(bias), g.ptr(res), act, post, g.ptr(out),
                                  g.stream()))
    torch.cuda.synchronize()
    y = A.double() @ B.double().t() + bias.double()
    if post:
        ref = F.gelu(y + res.double()) if act else y + res.double()
    else:
        ref = (F.gelu(y) if act else y) + res.double()
    e = g.rel(out.double(), ref)
    assert e < 1.5e-3, f"tf32 gemm rel-L2 {e}"


@pytest.mark.parametrize("B,H,W,Cin,Cpad,Cout", [(1, 8, 16, 32, 32, 256), (2, 5, 7, 196, 256, 588),
                                                  (1, 28, 37, 588, 640, 768)])
def test_conv3x3_tf32(B, H, W, Cin, Cpad, Cout):
    """fp32 NHWC implicit-GEMM conv on the TF32 path with padded channels, GELU after the residual
    (ResidualBlock, dense_rep_encoder.py:31-52), in place on the residual buffer."""
    from tests import gpu_helpers as g

    torch.manual_seed(H * W + Cin)
    x = torch.zeros(B, H, W, Cpad, device="cuda")
    x[..., :Cin] = torch.randn(B, H, W, Cin, device="cuda")
    w = torch.randn(Cout, Cin, 3, 3, device="cuda") / (9 * Cin) ** 0.5
    b = torch.randn(Cout, device="cuda") * 0.1
    res = torch.randn(B, H, W, Cout, device="cuda")
    wp = torch.empty(Cout, 9, Cpad, device="cuda")
    g.check(g.lib().mab_pack_conv3_weight_f32(g.ptr(w), Cout, Cin, Cpad, g.ptr(wp), g.stream()))
    out = res.clone()
    g.check(g.lib().mab_conv3x3_tf32(g.ptr(x), B, H, W, Cpad, g.ptr(wp), Cout, g.ptr(b), g.ptr(out), 1, 1,
                                     g.ptr(out), g.stream()))
    torch.cuda.synchronize()
    ref = F.conv2d(x[..., :Cin].double().permute(0, 3, 1, 2), w.double(), b.double(), padding=1).permute(0, 2, 3, 1)
    ref = F.gelu(ref + res.double())
    e = g.rel(out.double(), ref)
    assert e < 1.5e-3, f"tf32 conv rel-L2 {e}"
