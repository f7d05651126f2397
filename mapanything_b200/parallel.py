"""View sharding and the ring (context-parallel) schedule of global attention.

The reference has no multi-GPU inference path; this is the B200-native extension of the hot path
named by BASELINE.json: encoder, frame attention and heads are per-view independent, the one
exchange step is the global attention of the even info-sharing blocks
(alternating_attention_transformer.py:657-661), where every query must see every key/value.

Rank r keeps its queries resident and, for `world` steps, attends to the K/V block it currently
holds while the next block travels round the ring (`torch.distributed` P2P: NCCL over NVLink on the
GPU box, gloo in the CPU tests).  The attention kernel carries the online-softmax state between
steps, so no partial outputs are materialised.  Nothing here touches CUDA directly: `attend` is a
callback, which is what lets the schedule be tested with gloo on CPU.
"""
from typing import Callable, List, Sequence, Tuple

import torch
import torch.distributed as dist


def shard_table(num_views: int, world: int) -> List[Tuple[int, int]]:
    """Contiguous view blocks, sizes differing by at most one, earlier ranks take the extra view.
    View 0 (the reference view) always lands on rank 0."""
    if num_views < world:
        raise ValueError(f"cannot shard {num_views} views over {world} ranks: every rank needs a view")
    base, extra = divmod(num_views, world)
    table, lo = [], 0
    for r in range(world):
        hi = lo + base + (1 if r < extra else 0)
        table.append((lo, hi))
        lo = hi
    return table


def shard_views(num_views: int, world: int, rank: int) -> Tuple[int, int]:
    return shard_table(num_views, world)[rank]


def rows_per_rank(num_views: int, world: int, tokens_per_view: int) -> List[int]:
    """Info-sharing rows held by each rank: its views' tokens, plus the scale token on rank 0."""
    return [(hi - lo) * tokens_per_view + (1 if r == 0 else 0)
            for r, (lo, hi) in enumerate(shard_table(num_views, world))]


def ring_pass(local_block: torch.Tensor, rows: Sequence[int], recv_bufs: Sequence[torch.Tensor],
              attend: Callable[[torch.Tensor, int, int, int], None], group=None) -> None:
    """One global-attention layer over the ring.

    local_block : this rank's [rows[rank], C] block of packed qkv rows
    rows        : rows held by every rank
    recv_bufs   : two [max(rows), C] scratch tensors (double buffer)
    attend(block, n_rows, step, src_rank): run attention of the local queries against `block`
        (n_rows valid rows, produced by src_rank); step 0 is the first, step world-1 the last.

    The transfer of block s+1 is issued BEFORE the attention on block s is launched, so on NCCL it
    runs on the communicator's stream concurrently with the attention kernel."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    nxt_rank, prv_rank = (rank + 1) % world, (rank - 1) % world
    if group is not None:
        nxt_rank, prv_rank = dist.get_global_rank(group, nxt_rank), dist.get_global_rank(group, prv_rank)
    cur = local_block
    for step in range(world):
        src = (rank - step) % world
        reqs = []
        if step < world - 1:
            incoming = recv_bufs[step % 2][: rows[(rank - step - 1) % world]]
            reqs = dist.batch_isend_irecv([
                dist.P2POp(dist.isend, cur[: rows[src]], nxt_rank, group),
                dist.P2POp(dist.irecv, incoming, prv_rank, group)])
        attend(cur, rows[src], step, src)
        for r in reqs:
            r.wait()
        if step < world - 1:
            cur = incoming
