"""CPU-side checks: the C-ABI library builds, loads and exports every symbol include/mab200.h
declares (no compute call), and the host-side validation logic mirrors the reference."""
import os
import re
import subprocess

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    import __graft_entry__ as ge
    from mapanything_b200 import _lib

    ge.build()
    lib = _lib.load(allow_build=False)
    header = open(os.path.join(ROOT, "include", "mab200.h")).read()
    declared = set(re.findall(r"\b(mab_[a-z0-9_]+)\s*\(", header))
    assert declared, "no declarations found"
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in mab200.h but not exported"
        assert name in _lib.SYMBOLS, f"{name} has no ctypes signature"
    assert lib.mab_version() == 100


def test_validation_mirrors_reference():
    from mapanything_b200.model import validate_input_views_for_inference as val

    v = {"img": torch.zeros(1, 3, 28, 28), "data_norm_type": ["dinov2"]}
    assert val([v]) == [v]
    for bad in ([], [{"img": v["img"]}],
                [dict(v, intrinsics=torch.eye(3)[None], ray_directions=torch.zeros(1, 28, 28, 3))],
                [dict(v, depth_z=torch.zeros(1, 28, 28))],
                [v, dict(v, camera_poses=torch.eye(4)[None])],
                [v, {"data_norm_type": ["dinov2"], "camera_poses": torch.eye(4)[None]}]):
        with pytest.raises(ValueError):
            val(bad)


def test_product_fails_loudly_without_gpu():
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from mapanything_b200 import MapAnything

    with pytest.raises(RuntimeError):
        MapAnything()


def test_head_ranges_cover_all_views_in_order():
    from mapanything_b200.model import head_ranges

    assert head_ranges(8, 4, 4) == [(0, 4), (4, 8)]
    assert head_ranges(8, 0, 4) == [(0, 4), (4, 8)]          # 0: the library's head_chunk
    assert head_ranges(7, 3, 4) == [(0, 3), (3, 6), (6, 7)]  # ragged last range
    assert head_ranges(9, (4, 2, 1), 4) == [(0, 4), (4, 6), (6, 7), (7, 8), (8, 9)]  # last size repeats
    assert head_ranges(3, (), 2) == [(0, 2), (2, 3)]
    assert head_ranges(0, 4, 4) == []
    for V in (1, 5, 32, 33):
        for chunk in (1, 4, 64, (8, 4), (3,), (0, -2)):
            r = head_ranges(V, chunk, 4)
            assert r[0][0] == 0 and r[-1][1] == V
            assert all(a[1] == b[0] for a, b in zip(r, r[1:])) and all(lo < hi for lo, hi in r)


def test_constructor_rejects_unbuilt_trunk_configs_before_touching_the_gpu():
    """The configs/model/info_sharing/*.yaml variants: the unbuilt ones are refused by name, the built
    ones get as far as the device check (RuntimeError here: this container has no GPU)."""
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from mapanything_b200 import MapAnything

    def cfg(**margs):
        return {"model_type": "alternating_attention", "custom_positional_encoding": None,
                "module_args": dict(depth=24, indices=[11, 17], **margs)}

    for bad in (cfg(dim=512, num_heads=8), cfg(dim=1024, num_heads=12),
                dict(cfg(), model_type="cross_attention"), dict(cfg(), custom_positional_encoding=lambda x, p: x)):
        with pytest.raises(NotImplementedError):
            MapAnything(info_sharing_config=bad)
    with pytest.raises(NotImplementedError):
        MapAnything(pred_head_config={"type": "dpt"})
    with pytest.raises(TypeError, match="max_num_views"):      # gat_ifr_*.yaml: required by the GAT constructor
        MapAnything(info_sharing_config=dict(cfg(), model_type="global_attention"))
    with pytest.raises(ValueError, match="None implemented"):  # model.py:259-267
        MapAnything(info_sharing_config=dict(cfg(), custom_positional_encoding="RoPE100"))
    with pytest.raises(ValueError, match="2 or 3 indices"):   # model.py:328-331
        MapAnything(info_sharing_config=dict(cfg(), module_args=dict(depth=24, indices=[5, 11, 17, 23])))
    for ok in (cfg(), cfg(use_entropy_scaling=True), cfg(distinguish_ref_and_non_ref_views=False),
               cfg(use_pe_for_non_reference_views=True, max_num_views_for_pe=64),
               dict(cfg(max_num_views=1000, use_rand_idx_pe_for_non_reference_views=True),
                    model_type="global_attention"),                  # gat_ifr_24_layers.yaml
               cfg(distinguish_ref_and_non_ref_views=True, max_num_views_for_pe=1000,
                   use_rand_idx_pe_for_non_reference_views=True),    # aat_ifr_24_layers_w_view_pe.yaml
               dict(cfg(), module_args=dict(depth=48, indices=[11, 23, 35], dim=1024, num_heads=16))):  # aat_ifr_48_layers.yaml
        with pytest.raises(RuntimeError, match="CUDA device"):
            MapAnything(info_sharing_config=ok)


def test_entry_points_refuse_bad_arguments_before_any_cuda_call():
    """Argument errors come back as status -1 with a message from mab_last_error, before the first
    CUDA call (so this runs without a GPU and computes nothing)."""
    import ctypes as C

    from mapanything_b200 import _lib

    lib = _lib.load()
    one = C.c_void_p(16)  # a non-null, 16-byte aligned address that is never dereferenced
    calls = {
        "mab_gemm": (None, one, 8, 8, 8, None, None, None, 0, None, one, None),
        "mab_gemm_tf32": (one, one, 8, 8, 8, None, None, 0, 0, None, None),
        "mab_layernorm": (None, 4, 768, one, one, 1e-6, one, None, None),
        "mab_attention": (one, 10, one, 10, 12, 2, 8, 8, 8, 8, 0.125, one, None, None),  # 2 x 8 rows > 10
        "mab_attention_carry": (one, 8, None, 8, 12, 1, 8, 8, 0, 0, 0.125, one, None, None, 0, 0, None),
        "mab_conv3x3": (one, 1, 8, 8, 64, one, 64, None, None, None, 0, None, None, None),
        "mab_conv3x3_tf32": (one, 1, 8, 8, 48, one, 64, None, None, 0, 0, one, None),
        "mab_pack_conv3_weight": (None, 64, 64, one, None),
        "mab_pack_conv3_weight_f32": (one, 64, 40, 32, one, None),
        "mab_resize_bilinear": (one, 1, 8, 8, 12, 16, 16, 16, 16, one, None, None),
        "mab_posembed_resize": (None, 37, 1024, 28, 37, one, None),
        "mab_rope2d": (one, 7, 4, 2, 64, 128, one, 100.0, 1.0, None),
        "mab_rays_from_intrinsics": (one, 0, 8, 8, one, None),
        "mab_normalize_rays": (one, 0, one, None),
        "mab_recover_intrinsics": (None, 1, 8, 8, one, None),
        "mab_resize_nearest_crop": (None, 8, 8, 4, 4, 0, 0, 4, 4, one, None),
        "mab_edge_mask": (None, one, one, 1, 8, 8, 5.0, 0.03, one, one, None),
    }
    for name, args in calls.items():
        st = getattr(lib, name)(*args)
        msg = lib.mab_last_error().decode()
        assert st == -1 and msg.startswith("argument error"), (name, st, msg)
    with pytest.raises(RuntimeError, match="argument error"):
        _lib.check(lib.mab_normalize_rays(one, 0, one, None), "mab_normalize_rays")


def test_create_refuses_inconsistent_configs():
    import ctypes as C

    from mapanything_b200 import _lib
    from mapanything_b200.model import default_config

    lib = _lib.load()
    for change in (dict(aat_tap1=24), dict(aat_tap0=17), dict(aat_dim=1024), dict(aat_dim=512, aat_heads=8),
                   dict(aat_tap2=17), dict(aat_tap2=24), dict(enc_heads=12),
                   dict(aat_nonref_view_pe=1, aat_ref_view_pe=0),
                   dict(head_chunk=0), dict(patch=16), dict(enc_depth=0)):
        cfg = _lib.MabConfig(**dict(default_config(), **change))
        handle = C.c_void_p()
        assert lib.mab_create(C.byref(cfg), 0, C.byref(handle)) == -1, change
        assert lib.mab_last_error().decode().startswith("argument error") and not handle.value


def test_state_dict_bookkeeping_mirrors_torch_strict_loading():
    from mapanything_b200 import synthetic
    from mapanything_b200.model import check_state_dict

    sd = synthetic.synthetic_state_dict()
    assert check_state_dict(sd) == ([], [], [])
    # either name of an aliased tensor satisfies the model (dense_head.0.* == dpt_feature_head.*)
    canonical = {k: v for k, v in sd.items() if synthetic.canonical_key(k) == k}
    aliases = {k: v for k, v in sd.items() if synthetic.canonical_key(k) != k or not k.startswith("dpt_")}
    assert len(canonical) < len(sd) and check_state_dict(canonical) == ([], [], [])
    assert check_state_dict(aliases) == ([], [], [])

    bad = dict(sd)
    del bad["encoder.model.blocks.3.attn.qkv.weight"]
    bad["encoder.model.blocks.3.attn.qkv.wieght"] = sd["encoder.model.blocks.3.attn.qkv.weight"]
    bad["pose_head.fc_rot.weight"] = torch.zeros(3, 784)
    missing, unexpected, mismatch = check_state_dict(bad)
    assert missing == ["encoder.model.blocks.3.attn.qkv.weight"]
    assert unexpected == ["encoder.model.blocks.3.attn.qkv.wieght"]
    assert len(mismatch) == 1 and "pose_head.fc_rot.weight" in mismatch[0] and "(4, 784)" in mismatch[0]

    # aat_ifr_24_layers_no_ref_view.yaml: such a checkpoint has no view_pos_table buffer
    noref = {k: v for k, v in sd.items() if k != "info_sharing.view_pos_table"}
    assert check_state_dict(noref, ref_view_pe=False) == ([], [], [])
    assert check_state_dict(noref)[0] == ["info_sharing.view_pos_table"]
    assert check_state_dict(sd, ref_view_pe=False)[1] == ["info_sharing.view_pos_table"]
    # a shallower trunk: the extra blocks of a 24-layer checkpoint are unexpected, none is missing
    missing, unexpected, _ = check_state_dict(sd, aat_depth=20)
    assert not missing and unexpected and all(k.startswith("info_sharing.self_attention_blocks.2") for k in unexpected)


def test_state_dict_bookkeeping_follows_the_trunk_width_and_depth():
    """aat_ifr_48_layers.yaml: 48 blocks of width 1024, no proj_embed (aat.py:121-124), the consumers of
    trunk tokens follow the width.  oracle/make_golden.py pins manifest_for against the live reference."""
    from mapanything_b200 import synthetic
    from mapanything_b200.model import check_state_dict

    man = synthetic.manifest_for(1024, 48)
    assert len(man) == 1166 and "info_sharing.proj_embed.weight" not in man
    assert man["info_sharing.self_attention_blocks.47.attn.qkv.weight"] == [3072, 1024]
    assert man["pose_head.proj.weight"] == [784, 1024, 1, 1] and man["scale_head.proj.weight"] == [196, 1024]
    assert man["dpt_feature_head.input_process.0.0.0.weight"] == [96, 1024, 1, 1]   # tap0 instead of the encoder
    assert man["dense_head.0.input_process.3.0.0.weight"] == [768, 1024, 1, 1]
    assert synthetic.manifest_for() == synthetic.load_manifest()
    shapes = {k: torch.empty(v, device="meta") for k, v in man.items()}
    assert check_state_dict(shapes, aat_depth=48, aat_dim=1024) == ([], [], [])
    missing, unexpected, mismatch = check_state_dict(shapes)      # into the production configuration
    assert "info_sharing.proj_embed.weight" in missing and unexpected and mismatch


def test_view_pe_table_follows_max_num_views():
    """use_pe_for_non_reference_views: the view_pos_table buffer has max_num_views_for_pe rows of the
    sinusoid of aat.py:190-200 (row 0 = the reference-view row of every other configuration)."""
    import numpy as np

    from mapanything_b200 import synthetic
    from mapanything_b200.model import check_state_dict

    man = synthetic.manifest_for(view_pe_rows=64)
    assert man["info_sharing.view_pos_table"] == [64, 768]
    tab = synthetic._gen("info_sharing.view_pos_table", (64, 768))
    pos, j = 37, 501
    want = np.cos(pos / np.power(10000, 2 * (j // 2) / 768))
    assert abs(tab[pos, j] - want) < 1e-6 and tab[0, 0] == 0 and tab[0, 1] == 1
    assert (tab[:1] == synthetic._gen("info_sharing.view_pos_table", (1, 768))).all()
    shapes = {k: torch.empty(v, device="meta") for k, v in man.items()}
    assert check_state_dict(shapes, view_pe_rows=64) == ([], [], [])
    assert len(check_state_dict(shapes)[2]) == 1   # into the production configuration: size mismatch


def test_view_pe_rows_are_drawn_like_the_reference(monkeypatch):
    """aat.py:628-636 / global_attention_transformer.py:552-556: row 0 for the reference view, then
    torch.randint(1, max_num_views_for_pe, (V-1,)) from the global CPU generator, or 1..V-1; the rows of
    a rank's own views reach the library, one per view."""
    from mapanything_b200 import _lib
    from mapanything_b200.model import MapAnything

    class FakeLib:
        def mab_set_view_pe_rows(self, handle, rows, n):
            self.rows = [rows[i] for i in range(n)]
            return 0

    m = MapAnything.__new__(MapAnything)   # host logic only: no device, no library
    m._cfg = _lib.MabConfig(aat_nonref_view_pe=1)
    m._lib, m._handle, m._pinned_view_pe_rows = FakeLib(), None, None
    m._max_views_for_pe, m._rand_view_pe, m._check_view_count = 64, True, True

    torch.manual_seed(7)
    want = torch.randint(low=1, high=64, size=(4,))
    torch.manual_seed(7)
    rows = m._draw_view_pe_rows(5)
    assert rows.tolist() == [0] + want.tolist()
    torch.manual_seed(7)
    m._set_view_pe_rows(2, 2, 5)            # a shard holding views 2 and 3 of 5
    assert m._lib.rows == want.tolist()[1:3]
    m._rand_view_pe = False
    m._set_view_pe_rows(0, 5, 5)
    assert m._lib.rows == [0, 1, 2, 3, 4]
    m._pinned_view_pe_rows = torch.tensor([0, 9, 8, 7, 6])   # one draw for all scenes of a batched call
    m._set_view_pe_rows(1, 3, 5)
    assert m._lib.rows == [9, 8, 7]
    with pytest.raises(AssertionError, match="Expected less than 64 views"):
        m._set_view_pe_rows(0, 65, 65)
    m._cfg = _lib.MabConfig(aat_nonref_view_pe=0)             # other configurations: nothing is set
    m._lib.rows, m._check_view_count = None, False
    m._set_view_pe_rows(0, 5, 5)
    assert m._lib.rows is None


def test_production_config_is_accepted_and_other_heads_are_refused():
    """The constructor computes exactly one head / adaptor configuration (pred_head/dpt_pose_scale.yaml +
    adaptor_config/raydirs_depth_pose_confidence_mask_scale.yaml): the production config, as config.json
    carries it, passes the checks; a deviating one is refused by name instead of being mis-computed."""
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import copy
    import json

    from mapanything_b200 import MapAnything
    from oracle import ref_shim

    cfg = json.loads(json.dumps(ref_shim.reference_config()))  # +-inf survive as Infinity, like huggingface_hub's file
    keys = ("name", "encoder_config", "info_sharing_config", "pred_head_config", "geometric_input_config")
    prod = {k: cfg[k] for k in keys if k in cfg}
    with pytest.raises(RuntimeError, match="CUDA device"):
        MapAnything(**copy.deepcopy(prod))

    def variant(path, value):
        c = copy.deepcopy(prod)
        d = c
        for k in path[:-1]:
            d = d[k]
        d[path[-1]] = value
        return c

    for path, value in ((("pred_head_config", "adaptor_type"), "pointmap+confidence"),
                        (("pred_head_config", "dpt_adaptor", "depth_mode"), "linear"),
                        (("pred_head_config", "dpt_adaptor", "confidence_vmin"), 0),
                        (("pred_head_config", "dpt_adaptor", "ray_directions_normalize_to_unit_sphere"), False),
                        (("pred_head_config", "pose_adaptor", "quaternions_normalize"), False),
                        (("pred_head_config", "scale_adaptor", "mode"), "linear"),
                        (("pred_head_config", "feature_head", "feature_dim"), 128),
                        (("pred_head_config", "feature_head", "hooks"), [0, 1, 2]),
                        (("pred_head_config", "pose_head", "num_resconv_block"), 3),
                        (("pred_head_config", "regressor_head", "output_dim"), 4),
                        (("encoder_config", "name"), "dinov2_giant"),
                        (("encoder_config", "data_norm_type"), "dust3r")):
        with pytest.raises(NotImplementedError):
            MapAnything(**variant(path, value))


def test_nccl_library_is_found_and_exports_the_p2p_api():
    """mab_set_comm resolves NCCL with dlopen from the path mapanything_b200.nccl_comm names: the library
    torch ships must be there and export the calls mab_forward_sharded issues (no communicator is made)."""
    import ctypes

    from mapanything_b200 import nccl_comm

    path = nccl_comm.nccl_library_path()
    assert os.path.exists(path), path
    lib = ctypes.CDLL(path)
    for name in ("ncclGetUniqueId", "ncclCommInitRank", "ncclGroupStart", "ncclGroupEnd", "ncclSend", "ncclRecv",
                 "ncclBroadcast", "ncclGetErrorString"):
        assert hasattr(lib, name), name
    assert ctypes.sizeof(nccl_comm._UniqueId) == 128


def test_reference_tree_for_the_bench_legs():
    """oracle/make_ref.py: where /root/reference is mounted the two packages are copied into the git-ignored
    oracle/_ref/ (what travels to the GPU box); ref_root() names a tree that holds both packages."""
    from oracle import make_ref

    if not os.path.isdir("/root/reference/mapanything"):
        pytest.skip("no reference mount here")
    dst = make_ref.make()
    assert dst and all(os.path.isdir(os.path.join(dst, p)) for p in make_ref.PACKAGES)
    root = make_ref.ref_root()
    assert root and os.path.isfile(os.path.join(root, "mapanything", "models", "mapanything", "model.py"))
    tracked = subprocess.run(["git", "ls-files", "oracle/_ref"], capture_output=True, text=True, cwd=ROOT).stdout
    assert tracked.strip() == "", "oracle/_ref must stay out of history"
