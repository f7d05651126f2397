"""Host-side multi-rank logic on CPU: the view shard table and the ring schedule of global attention
(parallel.ring_pass) with world_size 2 and 3 over gloo.  `attend` is a torch restatement of what the
CUDA kernel does per ring step (online-softmax state carried between K/V blocks), so the test checks
that the schedule feeds every rank every block exactly once, with the right row counts, in an order
that reproduces full attention."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from mapanything_b200 import parallel


def test_shard_table():
    for n in (1, 2, 3, 8, 9, 13, 64, 511, 512, 2000):
        for w in (1, 2, 3, 4, 8):
            if n < w:
                with pytest.raises(ValueError):
                    parallel.shard_table(n, w)
                continue
            t = parallel.shard_table(n, w)
            assert t[0][0] == 0 and t[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(t, t[1:]))
            sizes = [hi - lo for lo, hi in t]
            assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)
            rows = parallel.rows_per_rank(n, w, 1036)
            assert sum(rows) == n * 1036 + 1 and rows[0] == sizes[0] * 1036 + 1
            assert parallel.shard_views(n, w, w - 1) == t[-1]


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_views, T, result, schedule="ring"):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        heads, hd = 2, 8
        D = heads * hd
        rows = parallel.rows_per_rank(n_views, world, T)
        g = torch.Generator().manual_seed(5)
        qkv_all = torch.randn(sum(rows), 3 * D, generator=g)  # every rank builds the same scene
        # global row order: rank blocks back to back (rank 0's block ends with the scale token row)
        off = [0]
        for r in rows:
            off.append(off[-1] + r)
        local = qkv_all[off[rank]:off[rank + 1]].clone()
        q = local[:, :D].view(-1, heads, hd).transpose(0, 1)
        state = {"m": torch.full((heads, rows[rank]), -float("inf")), "l": torch.zeros(heads, rows[rank]),
                 "o": torch.zeros(heads, rows[rank], hd), "seen": []}

        def attend(block, n, step, src):
            assert block.shape[0] >= n and n == rows[src]
            assert (step == 0) == (src == rank)
            k = block[:n, D:2 * D].view(n, heads, hd).transpose(0, 1)
            v = block[:n, 2 * D:].view(n, heads, hd).transpose(0, 1)
            s = q @ k.transpose(1, 2) * hd ** -0.5
            m_new = torch.maximum(state["m"], s.max(-1).values)
            alpha = torch.exp(state["m"] - m_new)
            p = torch.exp(s - m_new[..., None])
            state["l"] = state["l"] * alpha + p.sum(-1)
            state["o"] = state["o"] * alpha[..., None] + p @ v
            state["m"] = m_new
            state["seen"].append(src)

        if schedule == "ring":
            bufs = [torch.empty(max(rows), 3 * D) for _ in range(2)]
            parallel.ring_pass(local, rows, bufs, attend)
            assert sorted(state["seen"]) == list(range(world))
        else:  # all-to-all gather of the K|V columns, local block first, then <= 2 remote spans
            kv_all = torch.full((sum(rows), 2 * D), float("nan"))
            calls = []

            def attend_span(kv, n, first, last):
                calls.append((n, first, last))
                blk = torch.cat([torch.zeros(n, D), kv[:n]], dim=1)  # attend() reads columns D:3D
                state["seen"].append(-1)
                k = blk[:n, D:2 * D].view(n, heads, hd).transpose(0, 1)
                v = blk[:n, 2 * D:].view(n, heads, hd).transpose(0, 1)
                sc = q @ k.transpose(1, 2) * hd ** -0.5
                m_new = torch.maximum(state["m"], sc.max(-1).values)
                alpha = torch.exp(state["m"] - m_new)
                pr = torch.exp(sc - m_new[..., None])
                state["l"] = state["l"] * alpha + pr.sum(-1)
                state["o"] = state["o"] * alpha[..., None] + pr @ v
                state["m"] = m_new

            parallel.gather_pass(rows, kv_all, lambda dst: dst.copy_(local[:, D:]),
                                 lambda first, last: attend_span(local[:, D:], rows[rank], first, last),
                                 attend_span)
            assert not torch.isnan(kv_all).any()  # every block arrived
            assert torch.equal(kv_all, qkv_all[:, D:])
            assert sum(n for n, _, _ in calls) == sum(rows) and len(calls) <= 3
            assert calls[0][1] and calls[-1][2] and not any(c[1] for c in calls[1:])
            assert not any(c[2] for c in calls[:-1])
        out = (state["o"] / state["l"][..., None]).transpose(0, 1).reshape(rows[rank], D)
        K = qkv_all[:, D:2 * D].view(-1, heads, hd).transpose(0, 1)
        Vv = qkv_all[:, 2 * D:].view(-1, heads, hd).transpose(0, 1)
        ref = (torch.softmax(q @ K.transpose(1, 2) * hd ** -0.5, -1) @ Vv).transpose(0, 1).reshape(rows[rank], D)
        result[rank] = float((out - ref).abs().max())
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n_views", [(2, 2), (2, 5), (3, 7)])
def test_ring_pass_gloo(world, n_views):
    ctx = mp.get_context("spawn")
    with ctx.Manager() as man:
        result = man.dict()
        port = _free_port()
        procs = [ctx.Process(target=_worker, args=(r, world, port, n_views, 11, result)) for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(120)
            assert p.exitcode == 0, f"rank exited with {p.exitcode}"
        assert len(result) == world
        assert max(result.values()) < 1e-5, dict(result)


@pytest.mark.parametrize("world,n_views", [(2, 3), (3, 7)])
def test_gather_pass_gloo(world, n_views):
    ctx = mp.get_context("spawn")
    with ctx.Manager() as man:
        result = man.dict()
        port = _free_port()
        procs = [ctx.Process(target=_worker, args=(r, world, port, n_views, 11, result, "gather"))
                 for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(120)
            assert p.exitcode == 0, f"rank exited with {p.exitcode}"
        assert len(result) == world
        assert max(result.values()) < 1e-5, dict(result)


def _view_pe_worker(rank, world, port, total, result):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from mapanything_b200 import _lib
        from mapanything_b200.model import MapAnything

        class FakeLib:
            def mab_set_view_pe_rows(self, handle, rows, n):
                self.rows = [rows[i] for i in range(n)]
                return 0

        m = MapAnything.__new__(MapAnything)   # host logic only: no device, no library
        m._cfg = _lib.MabConfig(aat_nonref_view_pe=1)
        m._lib, m._handle, m._pinned_view_pe_rows = FakeLib(), None, None
        m._max_views_for_pe, m._rand_view_pe, m._check_view_count = 1000, True, True
        m._device = torch.device("cpu")
        torch.manual_seed(100 + rank)          # the ranks' generators disagree: rank 0's draw must win
        lo, hi = parallel.shard_views(total, world, rank)
        m._set_view_pe_rows(lo, hi - lo, total)
        sharded_rows = m._lib.rows
        # the same ranks running INDEPENDENT scenes (each holds its whole scene): no collective, own draw
        torch.manual_seed(200 + rank)
        if rank == 1:
            m._set_view_pe_rows(0, 4, 4)       # rank 1 alone: would hang if a broadcast were issued
        result[rank] = (lo, sharded_rows, m._lib.rows)
    finally:
        dist.destroy_process_group()


def test_view_pe_rows_agree_across_ranks_gloo():
    """use_pe_for_non_reference_views on a view-sharded scene: every rank draws from its own generator,
    rank 0's rows are broadcast, and each rank hands the library the rows of its own views."""
    world, total = 2, 7
    ctx = mp.get_context("spawn")
    with ctx.Manager() as man:
        result = man.dict()
        port = _free_port()
        procs = [ctx.Process(target=_view_pe_worker, args=(r, world, port, total, result)) for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(120)
            assert p.exitcode == 0, f"rank exited with {p.exitcode}"
        torch.manual_seed(100)
        want = [0] + torch.randint(low=1, high=1000, size=(total - 1,)).tolist()
        got = []
        for r in range(world):
            lo, rows, _ = result[r]
            assert lo == len(got)
            got += rows
        assert got == want
        torch.manual_seed(201)
        assert result[1][2] == [0] + torch.randint(low=1, high=1000, size=(3,)).tolist()
