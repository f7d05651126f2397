"""TEST INFRASTRUCTURE -- run under torch.distributed.run on N GPUs (by hand: `gpurun --gpus 8`):
many-view scenes through the view-sharded path (BASELINE.json configs[3]/[4]).

    python -m torch.distributed.run --nproc-per-node 8 ... tests/dist_manyview_check.py VIEWS [host] [ring]

The scene is 64 distinct views tiled VIEWS/64 times, so no oracle is needed: a view and its copies on the
SAME rank see every key/value block in the same order and must agree bit for bit (32-bit index or
byte-offset arithmetic anywhere in the path breaks this at these sizes); copies held by DIFFERENT ranks
see the blocks in another order (own block first), which re-draws the bf16 rounding noise: they agree to
the bf16 tolerance of the path (1e-2).  `host`: through infer(host_outputs=True), the predictions of every
view streamed to pinned host memory under the dense head (27 GB at 2000 views).  Prints one JSON line with
the step time, the peak device memory per rank and the two agreement figures."""
import json
import os
import sys
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
H, W, DISTINCT = 392, 518, 64


def main():
    n_views = int(sys.argv[1])
    host = "host" in sys.argv[2:]
    if "ring" in sys.argv[2:]:
        os.environ["MAB_EXCHANGE"] = "ring"
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    from mapanything_b200 import MapAnything, parallel, synthetic

    model = MapAnything(device=dev, head_chunk=8)
    model.load_state_dict(synthetic.synthetic_state_dict())
    lo, hi = parallel.shard_views(n_views, world, rank)
    base = synthetic.synthetic_views(DISTINCT, H, W, seed=1)
    torch.cuda.reset_peak_memory_stats(dev)
    if host:
        views = [{"img": base[i % DISTINCT]["img"].pin_memory(), "data_norm_type": ["dinov2"]} for i in range(lo, hi)]
        run = lambda: model.infer([dict(v) for v in views], view_range=(lo, hi), total_views=n_views,
                                  host_outputs=True, memory_efficient_inference=True)
    else:
        imgs = torch.cat([base[i % DISTINCT]["img"] for i in range(lo, hi)], 0).to(dev)
        run = lambda: model.run_resident(imgs, H, W, view_range=(lo, hi), total_views=n_views)
    run()  # warm-up (allocations, NCCL channels)
    dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    res = run()
    dist.barrier(); torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    depth = (torch.cat([p["depth_along_ray"] for p in res], 0) if host else res[0]["depth_along_ray"]).reshape(hi - lo, -1)
    conf = (torch.cat([p["conf"] for p in res], 0) if host else res[0]["conf"]).reshape(hi - lo, -1)
    assert bool(torch.isfinite(depth.float()).all()) and bool(torch.isfinite(conf.float()).all())
    # same-rank copies: views g and g + 64 (neither is the scene's view 0) must be identical
    same_rank_max = 0.0
    pairs = 0
    for g in range(lo, hi - DISTINCT):
        if g == 0:
            continue
        d = float((depth[g - lo].float() - depth[g + DISTINCT - lo].float()).abs().max())
        same_rank_max = max(same_rank_max, d)
        pairs += 1
    # cross-rank copies: every rank contributes its first copy of distinct view 1: a digest of 16384 pixels of
    # the confidence map (NOT of the depth: infer() zeroes the masked pixels of the depth, so the ~0.5 % of
    # pixels whose mask bit flips between two bf16 noise draws would count with their full magnitude)
    first = next(g for g in range(lo, hi) if g % DISTINCT == 1) if any(g % DISTINCT == 1 for g in range(lo, hi)) else None
    dig = conf[first - lo, :16384].float().to(dev) if first is not None else torch.full((16384,), float("nan"), device=dev)
    allv = [torch.empty_like(dig) for _ in range(world)]
    dist.all_gather(allv, dig)
    have = [t for t in allv if bool(torch.isfinite(t).all())]
    cross = max(float((t - have[0]).norm() / have[0].norm()) for t in have) if len(have) > 1 else 0.0
    peak = torch.tensor([torch.cuda.max_memory_allocated(dev) / 2**30], device=dev)
    dist.all_reduce(peak, op=dist.ReduceOp.MAX)
    stats = torch.tensor([same_rank_max, float(pairs)], device=dev, dtype=torch.float64)
    dist.all_reduce(stats[:1], op=dist.ReduceOp.MAX)
    dist.all_reduce(stats[1:], op=dist.ReduceOp.SUM)
    if rank == 0:
        flops = (1369.1 * n_views + 39.57 * n_views * n_views) * 1e9
        print(json.dumps({"views": n_views, "gpus": world, "api": "infer(host_outputs=True)" if host else "run_resident",
                          "exchange": os.environ.get("MAB_EXCHANGE", "gather"), "s_per_step": dt,
                          "views_per_s": n_views / dt, "tflops_per_gpu": flops / dt / 1e12 / world,
                          "peak_device_gib_per_rank": float(peak.item()),
                          "same_rank_copies_max_abs_diff": float(stats[0].item()), "same_rank_pairs": int(stats[1].item()),
                          "cross_rank_copies_rel_l2": cross, "timing": "wall clock around one step, barrier + synchronize on both sides"}),
              flush=True)
    assert float(stats[0].item()) == 0.0, float(stats[0].item())
    assert cross < 1e-2, cross
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
