"""Error bounds of the two in-kernel function approximations, restated in numpy fp32 (CPU):

* the degree-3 exp2 that a share of the attention softmax evaluates on the FMA pipe
  (csrc/attention.cuh exp2_poly2), against 2^x;
* the sigmoid-form GELU of the GEMM / conv epilogues (csrc/gemm.cuh gelu_erf), against the exact-erf
  GELU the reference uses (nn.GELU default).

The constants are parsed from the kernel sources, so a changed coefficient is re-checked here before
it reaches a GPU.  The bounds are the ones DESIGN.md quotes; both sit below the bf16 rounding
(2^-9 relative) of the value the kernel stores next.
"""
import math
import os
import re

import numpy as np

CSRC = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "mapanything_b200", "csrc")
f32 = np.float32


def _fma(a, b, c):
    # fp32 fma: the product of two floats is exact in double; one rounding to fp32 at the end
    return (a.astype(np.float64) * b.astype(np.float64) + c.astype(np.float64)).astype(f32)


def _floats(text):
    return [f32(m) for m in re.findall(r"(-?\d+\.\d+(?:e[+-]?\d+)?)f", text)]


def test_exp2_polynomial_error_bound():
    src = open(os.path.join(CSRC, "attention.cuh")).read()
    pc3, _, pc2, _ = _floats(re.search(r"pc3 = pack_f2\(.*", src).group(0))[:4]
    pc1, _, pc0, _ = _floats(re.search(r"pc1 = pack_f2\(.*", src).group(0))[:4]
    magic = f32(12582912.0)
    assert "12582912.f" in src

    # the softmax argument is (s - max) * scale * log2e <= 0 on the fast path, up to +8 under the lazy
    # row-max threshold; below -125 it is clamped
    x = np.concatenate([np.linspace(-140, 8.5, 2_000_001), np.arange(-126, 9) + 0.5,
                        np.nextafter(np.arange(-126, 9) + 0.5, 100)]).astype(f32)
    xc = np.maximum(x, f32(-125.0))
    r = (xc + magic).astype(f32)
    fr = _fma((r - magic).astype(f32), np.full_like(xc, -1), xc)
    assert float(np.abs(fr).max()) <= 0.5
    p = _fma(np.full_like(fr, pc3), fr, np.full_like(fr, pc2))
    p = _fma(p, fr, np.full_like(fr, pc1))
    p = _fma(p, fr, np.full_like(fr, pc0))
    e = (p.view(np.int32) + (r.view(np.int32) << 23)).view(f32)   # exponent add, as the kernel does
    ref = np.exp2(xc.astype(np.float64))
    relerr = np.abs(e.astype(np.float64) - ref) / ref
    assert np.isfinite(e).all() and (e > 0).all()
    assert relerr.max() < 1.0e-4, relerr.max()          # 7.5e-5 minimax + fp32 rounding
    assert relerr.max() < 2.0 ** -9 / 10                  # an order below the bf16 rounding of P


def test_gelu_sigmoid_form_error_bound():
    src = open(os.path.join(CSRC, "gemm.cuh")).read()
    body = src[src.index("const float xc = fminf"):src.index("return x * fast_rcp")]
    a2, a1 = _floats(re.search(r"float pnl = fmaf\(.*", body).group(0))[:2]
    a0 = _floats(re.search(r"pnl = fmaf\(pnl, x2, .*", body).group(0))[0]

    x = np.concatenate([np.linspace(-12, 12, 1_000_001), np.linspace(-1e-3, 1e-3, 1001)]).astype(f32)
    xc = np.clip(x, f32(-5), f32(5))
    x2 = (xc * xc).astype(f32)
    pnl = _fma(x2, np.full_like(x2, a2), np.full_like(x2, a1))
    pnl = _fma(pnl, x2, np.full_like(x2, a0))
    e = np.exp2((xc * pnl).astype(f32).astype(np.float64))      # ex2.approx: 2 ulp, far below the fit error
    got = x.astype(np.float64) / (1.0 + e)
    erf = np.vectorize(math.erf)
    ref = 0.5 * x.astype(np.float64) * (1.0 + erf(x.astype(np.float64) / math.sqrt(2.0)))
    err = np.abs(got - ref)
    assert err.max() < 3.0e-5, err.max()                  # DESIGN.md / gemm.cuh: 2.6e-5 absolute
    # where the activation is large the error is relative to it: below half a bf16 ulp (2^-9)
    big = np.abs(ref) > 1e-2
    assert (err[big] / np.abs(ref[big])).max() < 2.0 ** -9
