"""mab_export_points_static: the allocation-free variant of mab_export_points (scratch in a file-scope
`static __device__` array, include/mab200.h) against the workspace entry point, through the C ABI.

No counterpart in the reference.  Collected last on
purpose (file name).  Integer / byte / index work and fp32 arithmetic
in the same operand order, so the tolerance is the one tests/test_export.py uses against the oracle:
bit-exact.  Two calls in a row with different inputs: the second call must not see the first call's
scratch contents."""
import ctypes as C

import pytest
import torch

from mapanything_b200 import _lib

pytestmark = pytest.mark.gpu


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def _scene(seed, V, H, W, keep):
    g = torch.Generator(device="cuda").manual_seed(seed)
    d = torch.rand(V, H, W, device="cuda", generator=g) * 4.0 - 0.4       # some depths <= 0 (dropped)
    K = torch.tensor([[440.0, 0, W / 2], [0, 440.0, H / 2], [0, 0, 1]], device="cuda")[None].repeat(V, 1, 1).contiguous()
    P = torch.eye(4, device="cuda")[None].repeat(V, 1, 1)
    P[:, :3, 3] = torch.rand(V, 3, device="cuda", generator=g)
    P = P.contiguous()
    rgb = torch.rand(V, H, W, 3, device="cuda", generator=g)
    m = (torch.rand(V, H, W, device="cuda", generator=g) < keep).view(torch.uint8)
    return d, K, P, m, rgb


def _outputs(V, H, W):
    n = V * H * W
    return (torch.full((V, H, W, 3), -7.0, device="cuda"), torch.full((n, 3), -7.0, device="cuda"),
            torch.full((n, 3), 9, dtype=torch.uint8, device="cuda"),
            torch.full((V + 1,), -1, dtype=torch.int64, device="cuda"))


def _existing(lib, st, d, K, P, m, rgb):
    V, H, W = d.shape
    dense, xyz, col, counts = _outputs(V, H, W)
    nbytes = lib.mab_export_workspace_bytes(V, H, W)
    ws = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    _lib.check(lib.mab_export_points(_ptr(d), _ptr(K), _ptr(P), _ptr(m), _ptr(rgb), V, H, W, _ptr(dense), _ptr(xyz),
                                     _ptr(col), V * H * W, _ptr(counts), _ptr(ws), nbytes, st), "mab_export_points")
    return dense, xyz, col, counts


def _static(lib, st, d, K, P, m, rgb):
    V, H, W = d.shape
    dense, xyz, col, counts = _outputs(V, H, W)
    _lib.check(lib.mab_export_points_static(_ptr(d), _ptr(K), _ptr(P), _ptr(m), _ptr(rgb), V, H, W, _ptr(dense),
                                            _ptr(xyz), _ptr(col), V * H * W, _ptr(counts), st),
               "mab_export_points_static")
    return dense, xyz, col, counts


def test_static_scratch_export_equals_workspace_export_twice_in_a_row():
    lib = _lib.load()
    torch.cuda.set_device(0)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    # call 1: 8 views 392x518 (the benched shape), call 2: fewer, ragged tiles, a sparser mask -- the
    # scratch still holds call 1's counts and offsets when call 2 starts
    for seed, V, H, W, keep in ((11, 8, 392, 518, 0.6), (12, 3, 70, 98, 0.2)):
        scene = _scene(seed, V, H, W, keep)
        want = _existing(lib, st, *scene)
        got = _static(lib, st, *scene)
        torch.cuda.synchronize()
        n = int(want[3][-1])
        assert 0 < n < V * H * W
        assert torch.equal(got[3], want[3])                      # per-view counts and the total
        assert torch.equal(got[0], want[0])                      # dense world points
        assert torch.equal(got[1][:n], want[1][:n]) and torch.equal(got[2][:n], want[2][:n])
        assert bool((got[1][n:] == -7.0).all()) and bool((got[2][n:] == 9).all())   # nothing stored beyond the count
        valid = (scene[3] != 0) & (scene[0] > 0)
        assert n == int(valid.sum()) and torch.equal(got[1][:n], got[0][valid])     # stable boolean-mask order


def test_static_scratch_export_rejects_sizes_beyond_the_maximum():
    lib = _lib.load()
    torch.cuda.set_device(0)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    t = torch.zeros(16, device="cuda")
    c = torch.zeros(16, dtype=torch.int64, device="cuda")
    # 330 views of 392x518 = 65 670 tiles of 1024 pixels > MAB_EXPORT_STATIC_MAX_TILES (65 536):
    # refused before any launch, so the tiny buffers are never touched
    rc = lib.mab_export_points_static(_ptr(t), _ptr(t), None, None, None, 330, 392, 518, None, None, None, 0,
                                      _ptr(c), st)
    assert rc == -1 and b"MAB_EXPORT_STATIC_MAX_TILES" in lib.mab_last_error()
    torch.cuda.synchronize()
