"""CPU-side checks: the C-ABI library builds, loads and exports every symbol include/mab200.h
declares (no compute call), and the host-side validation logic mirrors the reference."""
import os
import re

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

    for bad in (cfg(dim=1024, num_heads=16), cfg(use_pe_for_non_reference_views=True),
                dict(cfg(), model_type="global_attention"), dict(cfg(), custom_positional_encoding="RoPE100")):
        with pytest.raises(NotImplementedError):
            MapAnything(info_sharing_config=bad)
    with pytest.raises(NotImplementedError):
        MapAnything(pred_head_config={"type": "dpt"})
    for ok in (cfg(), cfg(use_entropy_scaling=True), cfg(distinguish_ref_and_non_ref_views=False),
               cfg(distinguish_ref_and_non_ref_views=True, max_num_views_for_pe=1000,
                   use_rand_idx_pe_for_non_reference_views=True)):   # aat_ifr_24_layers_w_view_pe.yaml
        with pytest.raises(RuntimeError, match="CUDA device"):
            MapAnything(info_sharing_config=ok)
