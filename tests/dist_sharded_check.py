"""TEST INFRASTRUCTURE -- run under torch.distributed.run on >= 2 GPUs (tests/test_gpu_sharded.py, and by
hand `gpurun --gpus 2`): view-sharded infer over NCCL, with BOTH exchange schedules of the global layers
(all-to-all K|V gather, ring), must reproduce the single-GPU result.  bench.py --gpus N runs the same
comparison (parallel.sharded_vs_single) before its timed region, which is where the driver sees it."""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

# the two runs differ only in the order the K/V blocks of the global layers are presented in, which
# re-draws the bf16 rounding noise downstream (each path is ~4e-3 from fp32): bf16 tolerance, not round-off
TOL = 1e-2
# measured on 2 B200s (profiles/r2_sharded_parity_n2.json): every key, the near-cancelling mask logits and the
# small-image intrinsics fit included, agrees to <= 4.2e-3; one tolerance for all

def main():
    n_views, H, W = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
    mode = sys.argv[4] if len(sys.argv) > 4 else "img"
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=dev)
    from mapanything_b200 import MapAnything, parallel, synthetic

    kwargs = {}
    sd = synthetic.synthetic_state_dict()
    if mode == "viewpe":  # use_pe_for_non_reference_views: rank 0's draw of the rows is broadcast
        margs = {"use_pe_for_non_reference_views": True, "max_num_views_for_pe": 16,
                 "use_rand_idx_pe_for_non_reference_views": True}
        kwargs["info_sharing_config"] = {"model_type": "alternating_attention", "module_args": margs}
        sd = synthetic.synthetic_state_dict(view_pe_rows=16)
    model = MapAnything(device=dev, head_chunk=3, **kwargs)
    model.load_state_dict(sd)
    views = synthetic.synthetic_views(n_views, H, W, seed=3)
    if mode == "mm":  # rays + sparse depth + poses: pose algebra needs view 0 and a scene-wide norm
        views = synthetic.add_multimodal_inputs(views, seed=4)
    if mode == "viewpe":
        # the single-GPU run and the sharded runs must use the SAME rows: pin one draw for all of them
        rows = torch.cat([torch.zeros(1, dtype=torch.int64), torch.randperm(15)[: n_views - 1] + 1]).to(dev)
        dist.broadcast(rows, src=0)
        model._pinned_view_pe_rows = rows.cpu()
    rep = parallel.sharded_vs_single(model, views)
    if rank == 0:
        print("sharded-vs-single " + json.dumps(rep), flush=True)
    for k, e in rep["per_key"].items():
        assert e < TOL, (k, e, rep)
    assert rep["mask_agree"] > 0.97, rep
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
