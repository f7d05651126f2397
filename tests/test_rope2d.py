"""RoPE2D: the oracle against the live-reference golden vectors (CPU),
and mab_rope2d (csrc/elementwise.cuh rope2d_kernel) against the oracle and the golden vectors (GPU)."""
import ctypes as C
import glob
import os

import numpy as np
import pytest
import torch

from oracle import rope2d_oracle

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "rope2d_*.npz")))
# float tolerance: cos/sin of angles up to ~40 rad in fp32 (argument rounding 40 * 2^-24 = 2.4e-6);
# the reference's torch fallback rounds the angle as t * inv_freq, curope.cpp as (fwd * p) / powf(...)
TOL = 2e-5


def test_golden_files_exist():
    assert len(GOLDEN) == 3


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_oracle_matches_reference_golden(path):
    g = np.load(path)
    out = rope2d_oracle.rope2d(g["tokens"], g["positions"], float(g["base"]), 1.0)
    assert np.abs(out - g["out"]).max() < TOL
    # fwd = -1 is the inverse rotation (curope's backward pass): round trip
    back = rope2d_oracle.rope2d(out, g["positions"], float(g["base"]), -1.0)
    assert np.abs(back - g["tokens"]).max() < TOL


def _run(tokens, positions, base, fwd, pitch=None):
    from tests.gpu_helpers import check, lib, ptr, stream

    B, N, H, D = tokens.shape
    dtype = 0 if tokens.dtype == torch.float32 else 1
    check(lib().mab_rope2d(ptr(tokens), dtype, B * N, H, D, pitch or H * D, ptr(positions), base, fwd, stream()),
          "mab_rope2d")
    torch.cuda.synchronize()


@pytest.mark.gpu
@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_cuda_matches_golden_and_oracle(path):
    g = np.load(path)
    tok = torch.from_numpy(g["tokens"]).cuda().contiguous()
    pos = torch.from_numpy(g["positions"]).cuda().contiguous()
    _run(tok, pos, float(g["base"]), 1.0)
    assert float((tok.cpu() - torch.from_numpy(g["out"])).abs().max()) < TOL
    ora = rope2d_oracle.rope2d(g["tokens"], g["positions"], float(g["base"]), 1.0)
    assert float((tok.cpu() - torch.from_numpy(ora)).abs().max()) < TOL
    _run(tok, pos, float(g["base"]), -1.0)  # inverse rotation restores the input
    assert float((tok.cpu() - torch.from_numpy(g["tokens"])).abs().max()) < TOL


@pytest.mark.gpu
def test_cuda_bf16_in_place_on_packed_qkv():
    """q and k thirds of a packed [rows, 3*H*D] bf16 qkv buffer rotated where they lie; v untouched."""
    torch.manual_seed(3)
    N, H, D = 1036 * 3, 12, 64
    qkv = torch.randn(N, 3 * H * D, device="cuda").bfloat16()
    ys, xs = torch.meshgrid(torch.arange(28), torch.arange(37), indexing="ij")
    pos = torch.stack([ys.reshape(-1), xs.reshape(-1)], -1).repeat(3, 1).cuda().contiguous()
    ref = qkv.clone()
    for part in range(2):
        view = qkv[:, part * H * D:]
        from tests.gpu_helpers import check, lib, ptr, stream
        check(lib().mab_rope2d(C.c_void_p(view.data_ptr()), 1, N, H, D, 3 * H * D, ptr(pos), 100.0, 1.0, stream()),
              "mab_rope2d")
    torch.cuda.synchronize()
    assert torch.equal(qkv[:, 2 * H * D:], ref[:, 2 * H * D:])
    for part in range(2):
        src = ref[:, part * H * D:(part + 1) * H * D].float().cpu().numpy().reshape(1, N, H, D)
        ora = rope2d_oracle.rope2d(src, pos.cpu().numpy()[None], 100.0, 1.0).reshape(N, H * D)
        got = qkv[:, part * H * D:(part + 1) * H * D].float().cpu().numpy()
        assert np.abs(got - ora).max() <= 2.0 ** -8 * max(1.0, np.abs(ora).max())  # one bf16 rounding


@pytest.mark.gpu
def test_cuda_argument_errors_and_empty():
    from tests.gpu_helpers import lib, ptr, stream

    tok = torch.zeros(4, 2, 24, device="cuda")
    pos = torch.zeros(4, 2, dtype=torch.int64, device="cuda")
    assert lib().mab_rope2d(ptr(tok), 0, 4, 2, 24, 48, ptr(pos), 100.0, 1.0, stream()) < 0  # D % 16
    assert lib().mab_rope2d(ptr(tok), 0, 0, 2, 16, 32, ptr(pos), 100.0, 1.0, stream()) == 0  # empty
