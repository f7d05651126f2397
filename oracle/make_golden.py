"""TEST INFRASTRUCTURE ONLY -- generate tests/golden/*.npz from the LIVE reference.

Run in the build container (where /root/reference exists):

    python -m oracle.make_golden            # writes tests/golden/<case>.npz and prints the
                                            # reference-vs-restatement agreement for every key

Each case runs the unmodified reference `MapAnything.infer(views, use_amp=False)` on CPU/fp32 with
the deterministic synthetic checkpoint of oracle/weights.py and stores the prediction dicts
(sub-sampled for the full-resolution case so the fixture stays small; masks are bit-packed in
full).  tests/test_oracle_golden.py re-runs the restatement (oracle/mapanything_oracle.py) on the
same inputs anywhere (no /root/reference needed) and compares against these files; the `-m gpu`
tests compare the CUDA path against the same files.
"""
import os
import sys
import time

import numpy as np
import torch

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)

from oracle import weights  # noqa: E402

GOLDEN_DIR = os.path.join(_ROOT, "tests", "golden")

DENSE_KEYS = ["pts3d", "pts3d_cam", "ray_directions", "depth_along_ray", "depth_z", "conf",
              "non_ambiguous_mask_logits", "img_no_norm"]
SMALL_KEYS = ["cam_trans", "cam_quats", "camera_poses", "intrinsics", "metric_scaling_factor"]
MASK_KEYS = ["non_ambiguous_mask", "mask"]

# name -> (num_views, H, W, multimodal, subsample stride, infer kwargs)
CASES = {
    "c1_n2_392x518": (2, 392, 518, False, 7, {}),
    "c1_loose_n2_392x518": (2, 392, 518, False, 7, dict(edge_depth_threshold=0.1, edge_normal_threshold=30.0)),
    "small_n3_70x98": (3, 70, 98, False, 1, {}),
    "small_mm_n3_70x98": (3, 70, 98, True, 1, {}),
    "small_conf_n2_56x84": (2, 56, 84, False, 1, dict(apply_confidence_mask=True, confidence_percentile=25)),
}


# info_sharing module_args of the variant cases (configs/model/info_sharing/aat_ifr_24_layers_escaling.yaml,
# aat_ifr_24_layers_no_ref_view.yaml); every other case runs the production config
CASE_MODULE_ARGS = {
    "small_escaling_n3_70x98": dict(use_entropy_scaling=True),
    "small_noref_n3_70x98": dict(distinguish_ref_and_non_ref_views=False),
}
CASES["small_escaling_n3_70x98"] = (3, 70, 98, False, 1, {})
CASES["small_noref_n3_70x98"] = (3, 70, 98, False, 1, {})
# aat_ifr_48_layers.yaml: 48 blocks of width 1024 / 16 heads, three intermediate features (the DPT head
# then takes no encoder features, model.py:324-327) and no proj_embed (aat.py:121-124)
CASE_MODULE_ARGS["small_48l_n2_56x84"] = dict(name="aat_48_layers_ifr", indices=[11, 23, 35], size="48_layers",
                                              depth=48, dim=1024, num_heads=16)
CASES["small_48l_n2_56x84"] = (2, 56, 84, False, 1, {})
# use_pe_for_non_reference_views (aat.py:30,628-645) with the rows drawn at random from a seeded global
# generator (the default) and in sequence; a small table keeps the fixture's checkpoint cheap
CASE_MODULE_ARGS["small_viewpe_rand_n3_70x98"] = dict(use_pe_for_non_reference_views=True, max_num_views_for_pe=64)
CASE_MODULE_ARGS["small_viewpe_seq_n3_70x98"] = dict(use_pe_for_non_reference_views=True, max_num_views_for_pe=64,
                                                     use_rand_idx_pe_for_non_reference_views=False)
CASES["small_viewpe_rand_n3_70x98"] = (3, 70, 98, False, 1, {})
CASES["small_viewpe_seq_n3_70x98"] = (3, 70, 98, False, 1, {})
# gat_ifr_24_layers.yaml: model_type "global_attention" (every block global, every view encoded)
CASE_MODEL_TYPE = {"small_gat_n3_70x98": "global_attention"}
CASE_MODULE_ARGS["small_gat_n3_70x98"] = dict(name="gat_24_layers_ifr", max_num_views=64,
                                              use_rand_idx_pe_for_non_reference_views=True, indices=[11, 17],
                                              norm_intermediate=True, size="24_layers", depth=24,
                                              gradient_checkpointing=False)
CASES["small_gat_n3_70x98"] = (3, 70, 98, False, 1, {})
# a single view: global attention over one view's tokens + the scale token, identity pose
CASES["small_n1_56x84"] = (1, 56, 84, False, 1, {})
CASE_SEED = 1234  # torch.manual_seed before every run of a case: the random view-PE rows are part of the case


def case_state_dict(name):
    """The synthetic checkpoint of a case: same generator, shapes of that case's trunk."""
    margs = CASE_MODULE_ARGS.get(name, {})
    rows = margs.get("max_num_views_for_pe", 1000) if margs.get("use_pe_for_non_reference_views") else 1
    rows = margs.get("max_num_views", rows)  # global_attention: always a [max_num_views, dim] table
    return weights.synthetic_state_dict(aat_dim=margs.get("dim", 768), aat_depth=margs.get("depth", 24),
                                        view_pe_rows=rows)

def case_info_sharing_config(name):
    """info_sharing_config overrides of a case, as MapAnything(info_sharing_config=...) takes them."""
    return {"model_type": CASE_MODEL_TYPE.get(name, "alternating_attention"), "custom_positional_encoding": None,
            "module_args": dict(CASE_MODULE_ARGS.get(name, {}))}


def case_views(name):
    n, H, W, mm, _, _ = CASES[name]
    views = weights.synthetic_views(n, H, W, seed=1)
    if mm:
        views = weights.add_multimodal_inputs(views, seed=2)
    return views


def pack_outputs(preds, stride):
    out = {}
    for i, p in enumerate(preds):
        for k in DENSE_KEYS:
            t = p[k].detach().cpu().numpy()
            out[f"v{i}.{k}"] = np.ascontiguousarray(t[:, ::stride, ::stride]).astype(np.float32)
        for k in SMALL_KEYS:
            out[f"v{i}.{k}"] = p[k].detach().cpu().numpy().astype(np.float32)
        for k in MASK_KEYS:
            m = p[k].detach().cpu().numpy().astype(bool)
            out[f"v{i}.{k}.shape"] = np.array(m.shape, np.int64)
            out[f"v{i}.{k}.bits"] = np.packbits(m.reshape(-1))
    return out


def main(only=None):
    from oracle import mapanything_oracle as orc
    from oracle import ref_shim

    os.makedirs(GOLDEN_DIR, exist_ok=True)
    models = {}
    for name, (n, H, W, mm, stride, kw) in CASES.items():
        if only and name not in only:
            continue
        margs = CASE_MODULE_ARGS.get(name, {})
        sd = case_state_dict(name)
        mkey = repr(sorted(margs.items())) + CASE_MODEL_TYPE.get(name, "")
        oopts = orc.aat_opts_for(case_info_sharing_config(name))
        if mkey not in models:
            models[mkey] = ref_shim.build_reference_model(sd, margs, CASE_MODEL_TYPE.get(name))
            # the derived manifest is the live reference's key space and shapes for this variant
            live = {k: list(v.shape) for k, v in models[mkey].state_dict().items()}
            mine = {k: list(v.shape) for k, v in sd.items()}
            if not margs.get("distinguish_ref_and_non_ref_views", True):
                mine.pop("info_sharing.view_pos_table")
            assert live == mine, sorted(set(live.items()) ^ set(mine.items()))[:8]
        model = models[mkey]
        views = case_views(name)
        t0 = time.time()
        # the view-PE rows the reference draws (torch.randint from the global CPU generator, after a
        # device- and input-dependent number of other draws: model.py:999,1056,1223,1313-1362) are part
        # of the fixture; the tests hand the same rows to the restatement and to the CUDA path
        drawn, randint = [], torch.randint

        def spy(*a, **k):
            drawn.append(randint(*a, **k))
            return drawn[-1]

        torch.manual_seed(CASE_SEED)
        torch.randint = spy
        try:
            ref = model.infer([dict(v) for v in views], use_amp=False, amp_dtype="fp32", **kw)
        finally:
            torch.randint = randint
        t_ref = time.time() - t0
        gold = pack_outputs(ref, stride)
        if drawn:
            assert len(drawn) == 1
            gold["view_pe_rows"] = drawn[0].numpy().astype(np.int64)
        np.savez_compressed(os.path.join(GOLDEN_DIR, name + ".npz"), **gold)
        t0 = time.time()
        torch.randint = (lambda *a, **k: drawn[0].clone()) if drawn else randint
        try:
            mine = pack_outputs(orc.infer(sd, [dict(v) for v in views], aat_opts=oopts, **kw), stride)
        finally:
            torch.randint = randint
        if drawn:
            mine["view_pe_rows"] = gold["view_pe_rows"]
        t_orc = time.time() - t0
        print(f"== {name}: reference {t_ref:.1f}s, restatement {t_orc:.1f}s")
        worst = 0.0
        for k in sorted(gold):
            a, b = gold[k], mine[k]
            if k.endswith(".bits"):
                nd = int(np.unpackbits(a ^ b).sum())
                print(f"   {k:44s} differing bits {nd}")
            elif k.endswith(".shape"):
                assert (a == b).all(), k
            else:
                rel = float(np.linalg.norm(a - b) / (np.linalg.norm(a) + 1e-30))
                worst = max(worst, rel)
                if rel > 1e-5:
                    print(f"   {k:44s} rel-L2 {rel:.3e}")
        print(f"   worst rel-L2 over float keys: {worst:.3e}")


if __name__ == "__main__":
    main(sys.argv[1:] or None)
