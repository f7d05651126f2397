#!/usr/bin/env python
"""bench.py -- views/sec of the MapAnything inference hot path on B200 (contract: task prompt / DESIGN.md).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--views V] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A step = one whole `infer()` of ONE scene (encoder -> 24 alternating-attention blocks -> DPT / pose /
scale heads -> fused pointmap + post-processing) on synthetic views (518x392, DINOv2-normalised) with
random-init weights of the production architecture.  N=1: BASELINE.json configs[1], 8 views on one
B200.  N>1: one scene of 8*N views, view-sharded over N ranks (global attention: all-to-all K|V gather
over NCCL by default, ring above MAB_GATHER_MAX_BYTES), after a sharded-vs-single parity check on the
same GPUs (`sharded_parity`).  Every line also carries `fixed_scene` (the same 64-view scene at every N)
and, at N = 8, `c4_512` (BASELINE.json configs[3]).

JSON keys: see DESIGN.md "Measurement".  `value` times the C-ABI forward+postprocess with inputs
resident in HBM (CUDA events, max over ranks); `e2e` times `MapAnything.infer` from pinned HOST images
with every prediction tensor copied back to pinned host memory inside the timed region.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

H_IMG, W_IMG = 392, 518  # "518x392" is WxH (image.py:40-52)
VIEWS_PER_GPU = 8
METRIC = "views/sec (518px, bf16)"
UNIT = "views/s"


print_line = print  # replaced in main(): the JSON line goes to the ORIGINAL stdout, everything else to stderr


def flops_per_scene(n):
    """image-only algorithmic FLOPs of one scene of n views."""
    return (1369.1 * n + 39.57 * n * n) * 1e9


class ClockSampler:
    """nvidia-smi sampler running during the timed region (B200_PROFILING.md clocks line)."""

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, power = [], [], set(), []
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); power.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


def _reference_runner():
    """(kind, build_s, fn) with fn(views, device, amp) -> predictions: the UNMODIFIED reference from
    oracle/_ref (copied there from /root/reference by oracle/make_ref.py at build time; kind "reference"),
    else the oracle port (kind "port").  Test/bench infrastructure: never on the product path."""
    from mapanything_b200 import synthetic
    from oracle import make_ref

    sd = synthetic.synthetic_state_dict()
    root = make_ref.ref_root()
    t0 = time.perf_counter()
    if root is not None:
        os.environ["MAB_REFERENCE_ROOT"] = root
        import importlib

        from oracle import ref_shim
        importlib.reload(ref_shim)
        import contextlib
        import io
        with contextlib.redirect_stdout(io.StringIO()):  # the reference prints while it builds
            model = ref_shim.build_reference_model(sd)
        state = {"dev": "cpu"}

        def run(views, device="cpu", amp=False):
            if state["dev"] != device:
                model.to(device)
                state["dev"] = device
            vs = [{k: (t.to(device) if torch.is_tensor(t) else t) for k, t in v.items()} for v in views]
            with contextlib.redirect_stdout(io.StringIO()):
                return model.infer(vs, use_amp=amp, amp_dtype="bf16")

        def forward_only(views, device):
            """The reference's forward() alone, the way its infer() calls it (model.py:2304-2338): views on the
            device, validated and pre-processed once outside the timed call, bf16 autocast."""
            from mapanything.utils.inference import (preprocess_input_views_for_inference,
                                                     validate_input_views_for_inference)
            if state["dev"] != device:
                model.to(device)
                state["dev"] = device
            vs = [{k: (t.to(device) if torch.is_tensor(t) else t) for k, t in v.items()} for v in views]
            pv = preprocess_input_views_for_inference(validate_input_views_for_inference(vs))
            model._configure_geometric_input_config(use_calibration=True, use_depth=True, use_pose=True,
                                                    use_depth_scale=True, use_pose_scale=True)

            def step():
                with torch.inference_mode(), torch.autocast("cuda", enabled=True, dtype=torch.bfloat16), \
                        contextlib.redirect_stdout(io.StringIO()):
                    return model.forward(pv, memory_efficient_inference=False)

            return step, model._restore_original_geometric_input_config

        run.forward_only = forward_only
        return "reference", time.perf_counter() - t0, run
    from oracle import mapanything_oracle as orc

    def run_port(views, device="cpu", amp=False):
        sdd = sd if device == "cpu" else {k: v.to(device) for k, v in sd.items()}
        vs = [{k: (t.to(device) if torch.is_tensor(t) else t) for k, t in v.items()} for v in views]
        return orc.infer(sdd, vs, amp=amp)

    return "port", time.perf_counter() - t0, run_port


def _kind_text(kind):
    return ("the unmodified reference MapAnything.infer(use_amp=False) imported from oracle/_ref" if kind == "reference"
            else "oracle/mapanything_oracle.py (restatement pinned to the live reference by tests/golden)")


def reference_arm(args, rank, world):
    """The reference's own CPU path for this hot path (fp32, all host cores), timed on the box's host
    cores: the UNMODIFIED reference when oracle/_ref travelled with the snapshot, else the oracle port.
    N=1: every step is the whole BASELINE configs[1] scene (8 views 518x392) -- the same config as the
    GPU arm.  N>1 (a scene of 8*N views): each step is a bounded 8-view sample of that scene."""
    if rank != 0:
        return
    from mapanything_b200 import synthetic

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    n_views = args.views or VIEWS_PER_GPU * world
    sample_views = min(n_views, VIEWS_PER_GPU)
    kind, build_s, run = _reference_runner()
    views = synthetic.synthetic_views(sample_views, H_IMG, W_IMG, seed=1)
    warm = max(args.warmup, 1)
    # bounded sample: the first warm-up step is clocked; if K + W steps of the whole scene would not end
    # within the budget (a few minutes), every step takes a sub-sample of 4 or 2 of its views instead
    # (CPU views/s barely depends on the scene size at <= 8 views: the quadratic term is 2 % of the work)
    budget = float(os.environ.get("MAB_REF_BUDGET_S", "420"))
    t0 = time.perf_counter()
    run(views)
    first = time.perf_counter() - t0
    while sample_views > 2 and first * (sample_views / len(views)) * (args.steps + warm - 1) > budget:
        sample_views //= 2
    views = views[:sample_views]
    for _ in range(warm - 1):
        run(views)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        run(views)
    dt = (time.perf_counter() - t0) / args.steps
    v = sample_views / dt
    sample = (f"the whole scene: {sample_views} views 518x392 per step" if sample_views == n_views else
              f"{sample_views} of the scene's {n_views} views 518x392 per step")
    print_line(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": warm, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"MapAnything image-only infer, {n_views} views 518x392 (WxH), one scene, random-init weights of the production architecture",
                   "views": n_views, "sample": sample + " (fp32, CPU)", "same_config": sample_views == n_views},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": kind,
                         "sample": sample + ", " + _kind_text(kind) + f", fp32 torch CPU, {cores} threads"},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def cpu_and_gpu_baseline_legs(dev, budget_s=25.0):
    """Rank 0, N=1: (cpu_baseline, reference_gpu).  cpu_baseline = the reference's CPU path on a bounded
    sample (whole 8-view scenes for ~budget_s).  reference_gpu = the SAME reference code moved to this B200
    and run the way its README does (infer(use_amp=True, amp_dtype="bf16"): eager PyTorch, library SDPA /
    cuBLASLt / cuDNN), the bar SURVEY.md 8(d) names; forward() alone is timed too because the reference's
    post-processing runs in numpy on the host."""
    from mapanything_b200 import synthetic

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    kind, build_s, run = _reference_runner()
    views = synthetic.synthetic_views(VIEWS_PER_GPU, H_IMG, W_IMG, seed=1)
    t0 = time.perf_counter()
    reps = 0
    while True:
        run(views)
        reps += 1
        if time.perf_counter() - t0 > budget_s * 0.6 or reps >= 4:
            break
    dt = (time.perf_counter() - t0) / reps
    cpu = {"value": VIEWS_PER_GPU / dt, "unit": UNIT, "cores": cores, "kind": kind,
           "sample": f"{reps} x infer of the whole 8-view scene 518x392 (fp32, {_kind_text(kind)}, {cores} threads)"}
    gpu = None
    try:
        dname = str(dev)
        for _ in range(2):
            run(views, dname, True)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        n = 3
        for _ in range(n):
            run(views, dname, True)
        torch.cuda.synchronize()
        dt_inf = (time.perf_counter() - t0) / n
        gpu = {"value": VIEWS_PER_GPU / dt_inf, "unit": UNIT, "ms_per_step": dt_inf * 1e3, "kind": kind,
               "what": "infer(use_amp=True, amp_dtype='bf16') of the 8-view scene on the same B200, inputs on the "
                       "device, wall clock incl. the reference's host-side numpy post-processing"}
        if hasattr(run, "forward_only"):
            # the reference's forward() alone (no post-processing): its eager library kernels against this
            # repo's forward, CUDA events, 10 calls after 3 warm-ups
            step, restore = run.forward_only(views, dname)
            try:
                for _ in range(3):
                    step()
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(10):
                    step()
                e1.record()
                torch.cuda.synchronize()
                ms_f = e0.elapsed_time(e1) / 10
                gpu["forward_only"] = {"ms_per_step": ms_f, "views_per_s": VIEWS_PER_GPU / (ms_f * 1e-3),
                                       "what": "the reference's forward() alone under bf16 autocast (eager PyTorch: "
                                               "library SDPA / cuBLASLt / cuDNN), no post-processing, CUDA events"}
            finally:
                restore()
    except Exception as e:  # the baseline leg must never take the bench line down
        gpu = {"unavailable": f"{type(e).__name__}: {e}"[:200]}
    return cpu, gpu


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--views", type=int, default=0, help="views in the scene (default 8 per GPU)")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-roofline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-to-host leg (many-view sweeps)")
    ap.add_argument("--no-fixed-scene", action="store_true", help="skip the fixed 64-view (and, at N=8, 512-view) legs")
    ap.add_argument("--multimodal", action="store_true",
                    help="BASELINE.json configs[2]: images + intrinsics + poses + sparse depth (use with --views 32)")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    # stdout carries ONE JSON line.  Libraries write to file descriptor 1 behind Python's back (NCCL prints its
    # version banner there whatever NCCL_DEBUG_FILE says): fd 1 is pointed at stderr for the whole run and
    # the line goes out through a private duplicate of the original stdout.
    sys.stdout.flush()
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    global print_line
    print_line = lambda text: (real_stdout.write(text + "\n"), real_stdout.flush())
    if args.impl == "reference":
        reference_arm(args, rank, world)
        return
    assert args.warmup >= 3 or os.environ.get("MAB_BENCH_ALLOW_SHORT"), "timing rules: at least 3 warm-up steps"
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: libmab200 has no CPU path")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # stdout carries ONE JSON line: whatever NCCL logs (its version banner at NCCL_DEBUG >= VERSION,
        # INFO lines when the caller asks for them) goes to stderr
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=dev)
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run"

    from mapanything_b200 import MapAnything, _lib, synthetic
    from mapanything_b200.parallel import shard_views

    n_views = args.views or VIEWS_PER_GPU * world
    lo, hi = shard_views(n_views, world, rank)
    model = MapAnything(device=dev, head_chunk=8)
    model.load_state_dict(synthetic.synthetic_state_dict())
    lib = model._lib

    # this rank's views of the scene: pinned host copies (e2e) and a device-resident stack (value)
    all_views = synthetic.synthetic_views(n_views, H_IMG, W_IMG, seed=1) if n_views <= 64 else None
    if all_views is None:  # many-view scenes: tile 64 distinct views (content does not change the work)
        base = synthetic.synthetic_views(64, H_IMG, W_IMG, seed=1)
        all_views = [base[i % 64] for i in range(n_views)]
    host_views = [{"img": v["img"].pin_memory(), "data_norm_type": ["dinov2"]} for v in all_views[lo:hi]]
    dev_imgs = torch.cat([v["img"] for v in host_views], 0).to(dev)
    V = hi - lo

    mm_dev_views = None
    if args.multimodal:
        mm_all = synthetic.add_multimodal_inputs(all_views, seed=2)[lo:hi]  # poses follow the global view index
        mm_dev_views = [{k: (t.to(dev) if torch.is_tensor(t) else t) for k, t in v.items()} for v in mm_all]
        host_views = [{k: (t.pin_memory() if torch.is_tensor(t) else t) for k, t in v.items()} for v in mm_all]

    def step_resident():
        if mm_dev_views is not None:  # geometric inputs resident on the device, public infer()
            return model.infer([dict(v) for v in mm_dev_views], view_range=(lo, hi), total_views=n_views)
        outs, post = model.run_resident(dev_imgs, H_IMG, W_IMG, view_range=(lo, hi), total_views=n_views)
        return outs, post

    host_out = {}

    def step_e2e():
        # the public call with host tensors on both sides: pinned host images in, every prediction
        # tensor back in pinned host memory (copied out range by range under the dense head)
        preds = model.infer([dict(v) for v in host_views], view_range=(lo, hi), total_views=n_views,
                            host_outputs=True)
        nbytes = 0
        for p in preds:
            for k, t in p.items():
                assert t.device.type == "cpu"
                nbytes += t.numel() * t.element_size()
        return nbytes

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()) / steps

    # N > 1: before anything is timed, the view-sharded path must reproduce the single-GPU result on
    # this box (both exchange schedules; an uneven multi-modal scene and a full-resolution one).  The
    # result goes into the JSON line; a failure ends the run non-zero after the line is printed.
    shard_par = None
    if world > 1:
        from mapanything_b200 import parallel
        small = synthetic.add_multimodal_inputs(synthetic.synthetic_views(2 * world + 1, 70, 98, seed=3), seed=4)
        rep_a = parallel.sharded_vs_single(model, small)
        rep_b = parallel.sharded_vs_single(model, synthetic.synthetic_views(2 * world, H_IMG, W_IMG, seed=3))
        # tolerance of tests/dist_sharded_check.py: the two runs differ in the K/V block order of the global
        # layers only, which re-draws the bf16 rounding noise downstream -> bf16 tolerance of the path, every key
        def _ok(rep, small_img):
            return all(e < 1e-2 for e in rep["per_key"].values()) and rep["mask_agree"] > 0.985
        shard_par = {"worst_rel_l2": max(rep_a["worst_rel_l2"], rep_b["worst_rel_l2"]),
                     "mask_agree": min(rep_a["mask_agree"], rep_b["mask_agree"]),
                     "ok": bool(_ok(rep_a, True) and _ok(rep_b, False)),
                     "what": "shard of the view-sharded result vs the same views of the single-GPU result, every rank, "
                             "both exchange schedules (gather, ring), max over ranks",
                     "cases": {f"{2 * world + 1} views 70x98 multi-modal (uneven shards)": rep_a,
                               f"{2 * world} views 518x392 image-only": rep_b}}

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    l0 = lib.mab_launch_count()
    ms_step = timed(step_resident, args.steps, args.warmup)
    launches = (lib.mab_launch_count() - l0) // (args.steps + args.warmup)
    clocks = sampler.stop() if rank == 0 else None

    d2h = [0]

    def e2e_fn():
        d2h[0] = step_e2e()

    ms_e2e = timed(e2e_fn, max(3, args.steps // 2), 3) if not args.no_e2e else float("nan")
    h2d_bytes = V * 3 * H_IMG * W_IMG * 4

    roof = None
    if not args.no_roofline:
        # per-launch CUDA events (recorded by the library on the launch stream) over K more steps
        lib.mab_prof_enable(1)
        barrier()
        for _ in range(args.steps):
            step_resident()
        cats = {}
        for cat, name in enumerate(["gemm", "conv3", "conv3_regress", "attention", "layernorm", "elementwise", "postprocess",
                                    "gemm_tf32", "conv3_tf32"]):
            n, ms, fl, by = C.c_longlong(), C.c_double(), C.c_double(), C.c_double()
            _lib.check(lib.mab_prof_query(cat, C.byref(n), C.byref(ms), C.byref(fl), C.byref(by)))
            cats[name] = dict(launches=n.value // args.steps, ms=ms.value / args.steps, flops=fl.value / args.steps,
                              bytes=by.value / args.steps)
        lib.mab_prof_enable(0)
        peaks = {}
        try:
            with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
                peaks = json.load(f)
        except Exception:
            pass
        peak_tf = peaks.get("bf16_tflops_sustained", 1400.0)
        peak_src = "MEASURED_PEAKS.json bf16_tflops_sustained (kernel timed inside a long step)" if peaks else "fallback 1.4 PFLOP/s sustained (B200_PROFILING.md)"
        total_ms = sum(c["ms"] for c in cats.values())
        dom = max(("gemm", "attention", "conv3"), key=lambda k: cats[k]["ms"])
        d = cats[dom]
        ach = d["flops"] / (d["ms"] * 1e-3) / 1e12 if d["ms"] > 0 else 0.0
        kernel_name = {"gemm": "gemm_kernel<MODE_GEMM> (tcgen05 linear layers)",
                       "attention": "attention_kernel (tcgen05 flash attention)",
                       "conv3": "gemm_kernel<MODE_CONV3> (tcgen05 implicit-GEMM 3x3 conv)"}[dom]
        # DRAM traffic per launch of that kernel class from the committed `ncu --set full` capture of the
        # 8-view step of THIS round (profiles/r2_traffic.json, derived from the capture's summary CSV with the
        # git head recorded); only quoted for the workload it was captured on
        traffic, traffic_src = None, None
        try:
            with open(os.path.join(ROOT, "profiles", "r2_traffic.json")) as f:
                tj = json.load(f)
            if n_views == VIEWS_PER_GPU and world == 1 and dom in tj:
                traffic, traffic_src = tj[dom], tj["source"]
        except Exception:
            pass
        roof = {"bound": "tensor", "kernel": kernel_name, "achieved": ach, "peak": peak_tf, "unit": "TFLOP/s",
                "frac": ach / peak_tf, "traffic": traffic, "traffic_unit": "bytes per launch (ncu dram read + write)",
                "traffic_source": traffic_src, "peak_source": peak_src,
                "launches_per_step": d["launches"], "avg_launch_ms": d["ms"] / max(d["launches"], 1),
                "share_of_step": d["ms"] / total_ms if total_ms else None,
                "how": "library-recorded CUDA events around every launch on the launch stream, K extra steps after the timed region",
                "per_kernel": {k: {"ms_per_step": round(c["ms"], 4), "launches": c["launches"],
                                   "tflops": round(c["flops"] / (c["ms"] * 1e-3) / 1e12, 1) if c["ms"] > 0 and c["flops"] else None,
                                   "gbs": round(c["bytes"] / (c["ms"] * 1e-3) / 1e9, 1) if c["ms"] > 0 and c["bytes"] else None}
                               for k, c in cats.items()}}

    # N > 1: the same GPUs used scene-parallel (every rank runs its own 8-view scene, no exchange):
    # the serving-style throughput next to the one-scene-sharded number above
    scene_par = None
    if world > 1 and not args.views:
        own = dev_imgs[:VIEWS_PER_GPU] if V >= VIEWS_PER_GPU else None
        if own is not None:
            ms_sp = timed(lambda: model.run_resident(own, H_IMG, W_IMG), args.steps, 3)
            scene_par = {"value": VIEWS_PER_GPU * world / (ms_sp * 1e-3), "unit": UNIT, "ms_per_step": ms_sp,
                         "what": f"{world} independent {VIEWS_PER_GPU}-view scenes, one per GPU, no communication"}

    # fixed-scene legs (strong scaling, driver-clocked): the SAME 64-view scene at every N, and at N = 8
    # BASELINE.json configs[3] (512 views, one scene)
    def fixed_leg(views_total, steps, warmup):
        flo, fhi = shard_views(views_total, world, rank)
        base = synthetic.synthetic_views(64, H_IMG, W_IMG, seed=1)
        imgs = torch.cat([base[i % 64]["img"] for i in range(flo, fhi)], 0).to(dev)
        ms = timed(lambda: model.run_resident(imgs, H_IMG, W_IMG, view_range=(flo, fhi), total_views=views_total),
                   steps, warmup)
        del imgs
        return {"views": views_total, "ms_per_step": ms, "views_per_s": views_total / (ms * 1e-3), "steps": steps,
                "warmup": warmup, "tflops_per_gpu": flops_per_scene(views_total) / (ms * 1e-3) / 1e12 / world}

    fixed = c4 = None
    if not args.views and not args.multimodal and not args.no_fixed_scene:
        fixed = fixed_leg(64, 3, 1)
        if world == 8:
            c4 = fixed_leg(512, 2, 1)

    # single GPU: the same step replayed from a CUDA graph (the forward is one fixed kernel sequence on one
    # stream); reported next to `value`, not instead of it
    graph_leg = None
    if world == 1 and not args.views and not args.multimodal:
        try:
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                model.run_resident(dev_imgs, H_IMG, W_IMG)
            ms_g = timed(g.replay, args.steps, 3)
            graph_leg = {"ms_per_step": ms_g, "views_per_s": n_views / (ms_g * 1e-3),
                         "what": "run_resident captured once with torch.cuda.graph and replayed"}
            del g
        except Exception as e:
            graph_leg = {"unavailable": f"{type(e).__name__}: {e}"[:200]}

    # the bf16 dense head (opt-in, head_precision="bf16"): bf16 operands where the reference runs fp32 / TF32
    bf16_leg = None
    if world == 1 and not args.views and not args.multimodal:
        m16 = MapAnything(device=dev, head_chunk=8, head_precision="bf16")
        m16.load_state_dict(synthetic.synthetic_state_dict())
        ms16 = timed(lambda: m16.run_resident(dev_imgs, H_IMG, W_IMG), args.steps, 3)
        bf16_leg = {"ms_per_step": ms16, "views_per_s": n_views / (ms16 * 1e-3),
                    "what": "the same step with MapAnything(head_precision='bf16'); parity of that head: DESIGN.md section 5"}
        del m16

    cpu = ref_gpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu, ref_gpu = cpu_and_gpu_baseline_legs(dev)

    if rank == 0:
        value = n_views / (ms_step * 1e-3)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": f"MapAnything {'multi-modal (images + intrinsics + poses + sparse depth)' if args.multimodal else 'image-only'} infer, {n_views} views 518x392 (WxH), one scene, random-init weights of the production architecture (DINOv2 ViT-L/14 + 24-layer alternating attention + DPT/pose/scale heads)",
                       "views": n_views, "views_per_gpu": V, "parallelism": "single GPU" if world == 1 else f"one scene view-sharded x{world}; global attention: all-to-all K|V gather over NCCL/NVLink overlapped with the local block, carried online softmax",
                       "algorithmic_tflop_per_gpu_per_step": flops_per_scene(n_views) / 1e12 / world,
                       "note": "views per GPU are fixed but global attention is quadratic in the scene: algorithmic work per GPU grows as 10.95 + 2.53*N TFLOP with N GPUs (13.5 at N=1, 31.2 at N=8)",
                       "l2": "working set per step (activations of all views, > 1 GB) exceeds the 126 MB L2; no explicit flush"},
            "tflops_per_gpu": (flops_per_scene(n_views) + (125.5e9 * n_views if args.multimodal else 0.0)) / (ms_step * 1e-3) / 1e12 / world,
            "e2e": {"value": n_views / (ms_e2e * 1e-3), "unit": UNIT, "ms_per_step": ms_e2e,
                    "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": d2h[0],
                    "api": "MapAnything.infer(views, host_outputs=True): pinned host images in, all prediction tensors returned in pinned host memory"},
            "gpu_launches": int(launches) * args.steps,
            "gpu_launches_per_step": int(launches),
            "clocks": clocks,
        }
        if scene_par is not None:
            line["scene_parallel"] = scene_par
        if shard_par is not None:
            line["sharded_parity"] = shard_par
        if fixed is not None:
            line["fixed_scene"] = fixed
        if c4 is not None:
            line["c4_512"] = c4
        if ref_gpu is not None:
            line["reference_gpu"] = ref_gpu
        if graph_leg is not None:
            line["cuda_graph_replay"] = graph_leg
        if bf16_leg is not None:
            line["head_bf16"] = bf16_leg
        line["config"]["head_precision"] = "tf32 (fp32 activations, TF32 contractions: the reference's precision for the dense head; model.py:1774)"
        if roof is not None:
            line["roofline"] = roof
        if cpu is not None:
            line["cpu_baseline"] = cpu
        print_line(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    if shard_par is not None and not shard_par["ok"]:
        sys.exit("sharded_parity failed: " + json.dumps(shard_par["cases"]))


if __name__ == "__main__":
    main()
