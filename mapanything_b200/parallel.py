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


def gather_pass(rows: Sequence[int], kv_all: torch.Tensor, export_local: Callable[[torch.Tensor], None],
                attend_local: Callable[[bool, bool], None],
                attend_kv: Callable[[torch.Tensor, int, bool, bool], None], group=None) -> None:
    """One global-attention layer with an all-to-all gather of the K|V blocks.

    rows         : rows held by every rank
    kv_all       : [sum(rows), C] buffer, the same layout on every rank: rank r's K|V block occupies
                   rows offs[r]:offs[r+1]
    export_local(dst)                  : write this rank's K|V block into dst (its slice of kv_all)
    attend_local(first, last)          : attention of the local queries against the local block
    attend_kv(block, n, first, last)   : ... against n gathered rows

    Every rank posts its sends to and receives from all peers in one group (NVSwitch gives each pair
    full bandwidth) and runs the local-block attention while the blocks travel; the remote blocks
    are then attended in at most two contiguous spans (before / after the own block), the
    online-softmax state being carried between the calls.  Traffic per rank and layer: the other
    ranks' K|V rows once -- the same bytes as the ring, in one step instead of world-1."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    offs = [0]
    for r in rows:
        offs.append(offs[-1] + r)
    own = kv_all[offs[rank]: offs[rank + 1]]
    export_local(own)
    ops = []
    for d in range(1, world):
        dst, src = (rank + d) % world, (rank - d) % world
        gdst = dist.get_global_rank(group, dst) if group is not None else dst
        gsrc = dist.get_global_rank(group, src) if group is not None else src
        ops.append(dist.P2POp(dist.isend, own, gdst, group))
        ops.append(dist.P2POp(dist.irecv, kv_all[offs[src]: offs[src + 1]], gsrc, group))
    reqs = dist.batch_isend_irecv(ops) if ops else []
    spans = [(a, b) for a, b in ((0, offs[rank]), (offs[rank + 1], offs[world])) if b > a]
    attend_local(True, not spans)
    for r in reqs:
        r.wait()
    for i, (a, b) in enumerate(spans):
        attend_kv(kv_all[a:b], b - a, False, i == len(spans) - 1)


SELF_CHECK_KEYS = ("pts3d", "pts3d_cam", "ray_directions", "depth_along_ray", "conf", "non_ambiguous_mask_logits",
                   "cam_trans", "cam_quats", "metric_scaling_factor", "intrinsics")


def sharded_vs_single(model, views, exchange=("gather", "gather_host", "ring"), infer_kwargs=None, group=None):
    """Self-check of the view-sharded path on the GPUs of this job: every rank runs `views` (the WHOLE
    scene, a list of view dicts, identical on every rank) once on its own GPU and once view-sharded over
    the process group, per exchange schedule, and compares its shard of the sharded result with the
    same views of the single-GPU result.  Returns {"worst_rel_l2", "per_key", "mask_agree", "schedules"}
    reduced over ranks (max of the errors, min of the agreement).  Used by bench.py --gpus N (N > 1)
    before its timed region and by tests/dist_sharded_check.py.

    The two runs execute the same kernels on the same data except for the order in which the K/V blocks
    of the global layers are presented, which re-draws the bf16 rounding noise downstream: agreement is
    to the bf16 tolerance of the path (1e-2), not to fp32 round-off."""
    import os

    world, rank = dist.get_world_size(group), dist.get_rank(group)
    kw = dict(infer_kwargs or {})
    n = len(views)
    lo, hi = shard_views(n, world, rank)
    full = model.infer([dict(v) for v in views], **kw)
    per_key, agree = {}, 1.0
    saved = os.environ.get("MAB_EXCHANGE")
    try:
        saved_host = os.environ.get("MAB_EXCHANGE_HOST")
        for mode in exchange:
            # "gather": the library's one-call sharded forward (mab_forward_sharded, NCCL issued in C);
            # "gather_host" / "ring": the per-layer exchange sequenced from Python through torch.distributed
            os.environ["MAB_EXCHANGE"] = "gather" if mode == "gather_host" else mode
            os.environ["MAB_EXCHANGE_HOST"] = "1" if mode == "gather_host" else "0"
            mine = model.infer([dict(v) for v in views[lo:hi]], view_range=(lo, hi), total_views=n, **kw)
            for p, r in zip(mine, full[lo:hi]):
                both = (p["mask"] & r["mask"]).double() if "mask" in p else None
                for k in SELF_CHECK_KEYS:
                    a, b = p[k].double(), r[k].double()
                    if both is not None and k in ("pts3d", "pts3d_cam", "depth_along_ray"):
                        a, b = a * both, b * both  # infer() zeroes masked pixels
                    e = float((a - b).norm() / (b.norm() + 1e-30))
                    per_key[k] = max(per_key.get(k, 0.0), e)
                if "mask" in p:
                    agree = min(agree, float((p["mask"] == r["mask"]).float().mean()))
    finally:
        for name, val in (("MAB_EXCHANGE", saved), ("MAB_EXCHANGE_HOST", saved_host)):
            if val is None:
                os.environ.pop(name, None)
            else:
                os.environ[name] = val
    dev = full[0]["pts3d"].device
    t = torch.tensor([per_key[k] for k in SELF_CHECK_KEYS] + [-agree], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    vals = t.tolist()
    per_key = {k: vals[i] for i, k in enumerate(SELF_CHECK_KEYS)}
    return {"worst_rel_l2": max(per_key.values()), "per_key": per_key, "mask_agree": -vals[-1],
            "schedules": list(exchange), "views": n, "world": world}
