"""Image input path.

CPU: the oracle (oracle/image_oracle.py, a restatement of Pillow's 8-bit resampling + the reference's
resize/crop/normalise logic) against (a) the LIVE Pillow of this image and (b) golden vectors produced by
the LIVE reference `load_images`.  GPU: mab_preprocess_image / mapanything_b200.image against oracle and
golden.  uint8 stage: bit-exact; fp32 normalisation: bit-exact (same operations in the same order)."""
import glob
import os

import numpy as np
import pytest
import torch

from oracle import image_oracle as io
from oracle.make_golden_image import CASES, synthetic_photo

GOLDEN_DIR = os.path.join(os.path.dirname(__file__), "golden")


def _golden(name):
    return np.load(os.path.join(GOLDEN_DIR, name + ".npz"))


def test_golden_files_exist():
    assert len(glob.glob(os.path.join(GOLDEN_DIR, "image_*.npz"))) == len(CASES)


@pytest.mark.parametrize("shape", [((480, 640), (321, 200), "lanczos"), ((100, 150), (518, 392), "bicubic"),
                                   ((37, 53), (53, 20), "lanczos"), ((300, 300), (300, 150), "lanczos"),
                                   ((64, 64), (7, 5), "lanczos"), ((5, 9), (40, 33), "bicubic")])
def test_oracle_resize_equals_live_pillow(shape):
    PIL = pytest.importorskip("PIL.Image")
    (h, w), (ow, oh), f = shape
    img = np.random.default_rng(h * w).integers(0, 256, (h, w, 3)).astype(np.uint8)
    res = PIL.Resampling.LANCZOS if f == "lanczos" else PIL.Resampling.BICUBIC
    ref = np.asarray(PIL.fromarray(img).resize((ow, oh), resample=res))
    assert np.array_equal(io.pil_resize(img, ow, oh, f), ref)


@pytest.mark.parametrize("name", list(CASES))
def test_oracle_matches_reference_golden(name):
    g = _golden(name)
    sizes, kw = CASES[name]
    tw, th = io.target_size([(w, h) for h, w in sizes], kw.get("resize_mode", "fixed_mapping"), kw.get("size"))
    for i in range(len(sizes)):
        _, out = io.preprocess_image(g[f"src{i}"], tw, th)
        assert out.shape == g[f"out{i}"].shape
        assert np.array_equal(out, g[f"out{i}"]), (name, i, float(np.abs(out - g[f"out{i}"]).max()))
        assert tuple(g["true_shape"][i]) == (th, tw)


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(CASES))
def test_cuda_matches_golden_and_oracle(name):
    from mapanything_b200 import image

    g = _golden(name)
    sizes, kw = CASES[name]
    srcs = [g[f"src{i}"] for i in range(len(sizes))]
    views = image.preprocess_decoded(srcs, return_u8=True, **kw)
    for i, v in enumerate(views):
        got = v["img"][0].cpu().numpy()
        assert got.shape == g[f"out{i}"].shape
        assert np.array_equal(got, g[f"out{i}"]), (name, i, float(np.abs(got - g[f"out{i}"]).max()))
        u8, _ = io.preprocess_image(srcs[i], got.shape[2], got.shape[1])
        assert np.array_equal(v["img_u8"].cpu().numpy(), u8)      # byte stage against the oracle
        assert tuple(v["true_shape"][0]) == tuple(g["true_shape"][i])
        assert v["data_norm_type"] == ["dinov2"] and v["idx"] == i and v["instance"] == str(i)


@pytest.mark.gpu
def test_cuda_full_hd_and_edge_cases():
    """1080p -> 518x294 (the 16:9 mapping), portrait, no-resize identity, tiny images; vs the oracle."""
    from mapanything_b200 import image

    for (h, w), kw in [((1080, 1920), {}), ((1600, 900), {}), ((392, 518), {}), ((14, 14), dict(resize_mode="square", size=14)),
                       ((29, 1000), dict(resize_mode="fixed_size", size=(140, 28)))]:
        src = synthetic_photo(h, w, h + w)
        v = image.preprocess_decoded([src], return_u8=True, **kw)[0]
        tw, th = io.target_size([(w, h)], kw.get("resize_mode", "fixed_mapping"), kw.get("size"))
        u8, out = io.preprocess_image(src, tw, th)
        assert np.array_equal(v["img_u8"].cpu().numpy(), u8), (h, w)
        assert np.array_equal(v["img"][0].cpu().numpy(), out), (h, w)


def test_argument_errors_mirror_reference():
    from mapanything_b200 import image

    for kw in (dict(resize_mode="nope"), dict(resize_mode="square"), dict(resize_mode="square", size=(1, 2)),
               dict(resize_mode="fixed_size", size=5), dict(resize_mode="fixed_size", size=(14.0, 14))):
        with pytest.raises(ValueError):
            image.preprocess_decoded([np.zeros((4, 4, 3), np.uint8)], **kw)
    with pytest.raises(ValueError):
        image.preprocess_decoded([np.zeros((4, 4, 3), np.uint8)], norm_type="unknown")


# ---- preprocess_inputs: image + depth + intrinsics ------------------------------------------------
from oracle.make_golden_inputs import CASES as IN_CASES, case_views  # noqa: E402


@pytest.mark.parametrize("shape", [((480, 640), (518, 388)), ((100, 150), (518, 777)), ((37, 53), (20, 53)),
                                   ((7, 9), (100, 50))])
def test_oracle_nearest_equals_live_cv2(shape):
    cv2 = pytest.importorskip("cv2")
    (h, w), (ow, oh) = shape
    src = np.random.default_rng(h).random((h, w)).astype(np.float32)
    ref = cv2.resize(src, (ow, oh), fx=ow / w, fy=oh / h, interpolation=cv2.INTER_NEAREST)
    assert np.array_equal(io.cv2_resize_nearest(src, ow, oh), ref)


@pytest.mark.parametrize("name", list(IN_CASES))
def test_oracle_inputs_match_reference_golden(name):
    g = _golden(name)
    specs, kw = IN_CASES[name]
    tw, th = io.target_size([(s[1], s[0]) for s in specs], kw["resize_mode"], kw["size"])
    views = case_views(name)
    for i in range(len(specs)):
        K = g[f"K{i}"] if f"K{i}" in g else None
        if "ray_directions" in views[i]:  # image.py:572-590: intrinsics recovered from the rays
            from oracle import mapanything_oracle as orc
            K = orc.recover_pinhole_intrinsics(views[i]["ray_directions"][None])[0].numpy()
        img, d, Ko = io.preprocess_view(g[f"img{i}"], tw, th, g[f"depth{i}"], K)
        assert np.array_equal(img, g[f"out_img{i}"]) and np.array_equal(d, g[f"out_depth{i}"])
        if f"K{i}" in g:
            assert np.array_equal(Ko, g[f"out_K{i}"])
        elif K is not None:  # recovered by least squares: float tolerance
            assert np.abs(Ko - g[f"out_K{i}"]).max() < 1e-3 * np.abs(g[f"out_K{i}"]).max()


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(IN_CASES))
def test_cuda_inputs_match_golden(name):
    from mapanything_b200 import image

    g = _golden(name)
    specs, kw = IN_CASES[name]
    views = case_views(name)
    assert all(np.array_equal(v["img"].numpy(), g[f"img{i}"]) for i, v in enumerate(views))
    out = image.preprocess_inputs(views, **kw)
    for i, o in enumerate(out):
        assert np.array_equal(o["img"][0].cpu().numpy(), g[f"out_img{i}"]), (name, i)
        assert np.array_equal(o["depth_z"][0].cpu().numpy(), g[f"out_depth{i}"]), (name, i)
        if f"K{i}" in g:
            assert np.array_equal(o["intrinsics"][0].cpu().numpy(), g[f"out_K{i}"])
        elif f"out_K{i}" in g:  # from ray_directions: least squares in double here, fp32 in the reference
            assert np.abs(o["intrinsics"][0].cpu().numpy() - g[f"out_K{i}"]).max() < 1e-3 * np.abs(g[f"out_K{i}"]).max()
        assert o["data_norm_type"] == ["dinov2"]


def test_oracle_resize_equals_live_pillow_random_sizes():
    """Seeded sweep over odd sizes and scale factors (shrink -> LANCZOS, enlarge -> BICUBIC), both passes,
    single-pass cases (one dimension unchanged) and extreme aspect ratios."""
    PIL = pytest.importorskip("PIL.Image")
    rng = np.random.default_rng(2024)
    for _ in range(40):
        h, w = int(rng.integers(3, 120)), int(rng.integers(3, 160))
        if rng.random() < 0.5:   # shrink
            oh, ow, f = int(rng.integers(1, h + 1)), int(rng.integers(1, w + 1)), "lanczos"
        else:                    # enlarge
            oh, ow, f = int(rng.integers(h, 3 * h + 1)), int(rng.integers(w, 3 * w + 1)), "bicubic"
        img = rng.integers(0, 256, (h, w, 3)).astype(np.uint8)
        res = PIL.Resampling.LANCZOS if f == "lanczos" else PIL.Resampling.BICUBIC
        ref = np.asarray(PIL.fromarray(img).resize((ow, oh), resample=res))
        assert np.array_equal(io.pil_resize(img, ow, oh, f), ref), (h, w, oh, ow, f)


def test_oracle_crop_plan_matches_resize_plan_without_intrinsics():
    for (w1, h1, tw, th) in [(640, 480, 518, 392), (1920, 1080, 518, 294), (90, 60, 126, 84), (333, 777, 294, 518)]:
        rw, rh, filt, left, top = io.resize_plan(w1, h1, tw, th)
        assert io.crop_plan(w1, h1, tw, th)[:5] == (rw, rh, filt, left, top)
        assert rw >= tw and rh >= th and 0 <= left <= rw - tw and 0 <= top <= rh - th
