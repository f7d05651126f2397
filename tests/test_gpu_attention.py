"""tcgen05 flash attention vs fp32 softmax(QK^T/8)V on the same bf16 inputs."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(params=["short_kernel", "long_kernel"])
def kernel(request):
    """Both attention kernels at every test size: the one-Q-tile kernel (two CTAs per SM, production
    choice up to 6000 keys) and the two-Q-tile kernel."""
    from tests import gpu_helpers as g

    g.check(g.lib().mab_set_attention_short_max(1 << 30 if request.param == "short_kernel" else 0))
    yield request.param
    g.check(g.lib().mab_set_attention_short_max(-1))


def _ref(qkv, heads, num_seqs, q_len, kv_len, q_stride, kv_stride, kv=None):
    D = heads * 64
    kvt = qkv if kv is None else kv
    out = torch.zeros((qkv.shape[0], D), dtype=torch.float32, device=qkv.device)
    lse = torch.zeros((qkv.shape[0], heads), dtype=torch.float32, device=qkv.device)
    for s in range(num_seqs):
        q = qkv[s * q_stride:s * q_stride + q_len, :D].float().view(q_len, heads, 64).transpose(0, 1)
        k = kvt[s * kv_stride:s * kv_stride + kv_len, D:2 * D].float().view(kv_len, heads, 64).transpose(0, 1)
        v = kvt[s * kv_stride:s * kv_stride + kv_len, 2 * D:].float().view(kv_len, heads, 64).transpose(0, 1)
        sc = q @ k.transpose(1, 2) * 0.125
        p = torch.softmax(sc, dim=-1)
        out[s * q_stride:s * q_stride + q_len] = (p @ v).transpose(0, 1).reshape(q_len, D)
        lse[s * q_stride:s * q_stride + q_len] = torch.logsumexp(sc, dim=-1).t()
    return out, lse


@pytest.mark.parametrize("heads,num_seqs,length,peaky", [
    (1, 1, 128, False),     # exactly one tile
    (1, 1, 100, False),     # one partial tile (masking)
    (2, 1, 256, False),     # two kv tiles (online-softmax rescale)
    (2, 3, 333, True),      # several sequences, tails, peaky logits
    (16, 2, 1037, True),    # encoder shape
    (12, 2, 1036, False),   # frame attention shape
    (12, 1, 2073, True),    # global attention, 2 views + scale token
    (3, 1, 5000, True),     # long sequence: many kv tiles, ring wrap-around of every barrier
])
def test_attention_self(heads, num_seqs, length, peaky, kernel):
    from tests import gpu_helpers as g

    torch.manual_seed(heads * 100 + length)
    D = heads * 64
    rows = num_seqs * length
    qkv = torch.randn(rows, 3 * D, device="cuda")
    if peaky:
        qkv[:, :2 * D] *= 3.0
    qkv = qkv.bfloat16()
    out, lse = g.attention(qkv, heads, num_seqs, length, length, length, length, lse=True)
    ref, ref_lse = _ref(qkv, heads, num_seqs, length, length, length, length)
    e = g.rel(out.float(), ref)
    assert e < 8e-3, f"attention rel-L2 {e}"
    assert (lse - ref_lse).abs().max() < 2e-2


def test_attention_separate_kv(kernel):
    """global layer on a view shard: local queries against the gathered K/V of every rank"""
    from tests import gpu_helpers as g

    torch.manual_seed(3)
    heads, D = 12, 768
    q = torch.randn(1037, 3 * D, device="cuda").bfloat16()
    kv = torch.randn(3109, 3 * D, device="cuda").bfloat16()
    out = g.attention(q, heads, 1, 1037, 3109, 0, 0, kv=kv)
    ref, _ = _ref(q, heads, 1, 1037, 3109, 0, 0, kv=kv)
    assert g.rel(out.float(), ref) < 8e-3


def test_attention_ring_carry(kernel):
    """ring global attention on one GPU: the K/V of the scene arrive as uneven blocks (as they do
    from the other ranks of a view-sharded run) and the online-softmax state is carried between the
    per-block launches; the result must equal attention over the concatenated K/V."""
    import ctypes as C
    from tests import gpu_helpers as g

    torch.manual_seed(9)
    heads, D = 12, 768
    q = (torch.randn(2 * 1036 + 1, 3 * D, device="cuda") * 2).bfloat16()
    blocks = [q, (torch.randn(2072, 3 * D, device="cuda") * 2).bfloat16(),
              (torch.randn(1036, 3 * D, device="cuda") * 2).bfloat16()]
    rows = q.shape[0]
    out = torch.zeros((rows, D), dtype=torch.bfloat16, device="cuda")
    co = torch.empty((rows, D), dtype=torch.float32, device="cuda")
    cml = torch.empty((rows, heads, 2), dtype=torch.float32, device="cuda")
    for i, b in enumerate(blocks):
        g.check(g.lib().mab_attention_carry(g.ptr(q), rows, g.ptr(b), b.shape[0], heads, 1, rows, b.shape[0], 0, 0,
                                            0.125, g.ptr(out), g.ptr(co), g.ptr(cml), int(i > 0),
                                            int(i < len(blocks) - 1), g.stream()), "mab_attention_carry")
    torch.cuda.synchronize()
    kv = torch.cat(blocks, 0)
    ref, _ = _ref(q, heads, 1, rows, kv.shape[0], 0, 0, kv=kv)
    assert g.rel(out.float(), ref) < 8e-3
