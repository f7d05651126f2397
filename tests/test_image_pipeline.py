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
