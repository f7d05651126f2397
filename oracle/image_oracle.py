"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's image input path.

`load_images` (mapanything/utils/image.py:134-332) -> `crop_resize_if_necessary`
(mapanything/utils/cropping.py:385-467) -> `rescale_image_and_other_optional_info` (:188-283: target
`floor(size * (max(out / in) + 1e-8))`, PIL LANCZOS when shrinking, BICUBIC otherwise) -> centred crop
(:441-447) -> torchvision `ToTensor` + `Normalize` (image.py:297-300: u8 / 255, then (x - mean) / std
in fp32).

The resampling arithmetic is a THIRD-PARTY dependency that is not in /root/reference: Pillow
(`PIL.Image.resize`; 12.2.0 installed in this image, unpinned by the reference's pyproject).  Its
published algorithm (Pillow src/libImaging/Resample.c: `precompute_coeffs`, `normalize_coeffs_8bpc`,
`ImagingResampleHorizontal_8bpc` / `Vertical_8bpc`) is restated here: per output sample a window of
`support * max(scale, 1)` input samples around the centre `(xx + 0.5) * scale`, double-precision
filter weights normalised to sum 1, quantised to 22-bit fixed point (round half away from zero), integer
accumulation from 2^21, `>> 22`, clip to [0, 255]; horizontal pass first with an 8-bit intermediate,
then the vertical pass.  Pinned bit-for-bit against the LIVE Pillow and the LIVE reference `load_images`
by tests/test_image_pipeline.py (CPU, in any container that has Pillow) and by the golden vectors
tests/golden/image_*.npz (oracle/make_golden_image.py).  Only tests/ may import this file.
"""
import math

import numpy as np

PRECISION_BITS = 32 - 8 - 2

RESOLUTION_MAPPINGS = {  # image.py:40-66
    518: {1.000: (518, 518), 1.321: (518, 392), 1.542: (518, 336), 1.762: (518, 294), 2.056: (518, 252),
          3.083: (518, 168), 0.757: (392, 518), 0.649: (336, 518), 0.567: (294, 518), 0.486: (252, 518)},
    512: {1.000: (512, 512), 1.333: (512, 384), 1.524: (512, 336), 1.778: (512, 288), 2.000: (512, 256),
          3.200: (512, 160), 0.750: (384, 512), 0.656: (336, 512), 0.562: (288, 512), 0.500: (256, 512)},
}
DINOV2_MEAN, DINOV2_STD = (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)  # image_normalizations.py:21


def _sinc(x):
    if x == 0.0:
        return 1.0
    x = x * math.pi
    return math.sin(x) / x


def lanczos_filter(x):
    return _sinc(x) * _sinc(x / 3) if -3.0 <= x < 3.0 else 0.0


def bicubic_filter(x):
    a = -0.5
    x = abs(x)
    if x < 1.0:
        return ((a + 2.0) * x - (a + 3.0)) * x * x + 1
    if x < 2.0:
        return (((x - 5) * x + 8) * x - 4) * a
    return 0.0


FILTERS = {"lanczos": (lanczos_filter, 3.0), "bicubic": (bicubic_filter, 2.0)}


def precompute_coeffs(in_size, out_size, filt):
    """Resample.c precompute_coeffs + normalize_coeffs_8bpc for the full box (0, in_size).
    -> (ksize, bounds [out,2] int32 (xmin, count), kk [out, ksize] int32)."""
    f, fsupport = FILTERS[filt]
    scale = filterscale = float(np.float32(in_size) - np.float32(0)) / out_size
    if filterscale < 1.0:
        filterscale = 1.0
    support = fsupport * filterscale
    ksize = int(math.ceil(support)) * 2 + 1
    bounds = np.zeros((out_size, 2), np.int32)
    kk = np.zeros((out_size, ksize), np.int32)
    ss = 1.0 / filterscale
    for xx in range(out_size):
        center = 0.0 + (xx + 0.5) * scale
        xmin = max(int(center - support + 0.5), 0)
        xmax = min(int(center + support + 0.5), in_size) - xmin
        w = [f((x + xmin - center + 0.5) * ss) for x in range(xmax)]
        ww = 0.0
        for v in w:
            ww += v
        for x in range(xmax):
            v = w[x] / ww if ww != 0.0 else w[x]
            kk[xx, x] = int(-0.5 + v * (1 << PRECISION_BITS)) if v < 0 else int(0.5 + v * (1 << PRECISION_BITS))
        bounds[xx] = (xmin, xmax)
    return ksize, bounds, kk


def _pass(img, out_size, filt, axis):
    """one separable pass over `axis` of a uint8 [H, W, C] image"""
    _, bounds, kk = precompute_coeffs(img.shape[axis], out_size, filt)
    src = np.moveaxis(img, axis, 0).astype(np.int64)
    out = np.empty((out_size,) + src.shape[1:], np.uint8)
    for xx in range(out_size):
        x0, n = bounds[xx]
        acc = np.tensordot(kk[xx, :n].astype(np.int64), src[x0:x0 + n], axes=(0, 0)) + (1 << (PRECISION_BITS - 1))
        out[xx] = np.clip(acc >> PRECISION_BITS, 0, 255).astype(np.uint8)
    return np.moveaxis(out, 0, axis)


def pil_resize(img, out_w, out_h, filt):
    """PIL.Image.resize((out_w, out_h), resample=filt) for an RGB uint8 [H, W, 3] array."""
    H, W = img.shape[:2]
    if out_w != W:
        img = _pass(img, out_w, filt, 1)   # ImagingResample: horizontal pass first
    if out_h != H:
        img = _pass(img, out_h, filt, 0)
    return img


def resize_plan(w1, h1, target_w, target_h):
    """cropping.py:230-244 and :441-447 -> (resize_w, resize_h, filter, crop_left, crop_top)."""
    inp = np.array([w1, h1])
    scale_final = max(np.array([target_w, target_h]) / inp) + 1e-8
    rw, rh = (int(v) for v in np.floor(inp * scale_final).astype(int))
    return rw, rh, ("lanczos" if scale_final < 1 else "bicubic"), (rw - target_w) // 2, (rh - target_h) // 2


def target_size(sizes, resize_mode="fixed_mapping", size=None, patch_size=14, resolution_set=518):
    """image.py:248-284: common (W, H) for a list of (w, h) image sizes."""
    ar = [w / h for w, h in sizes]
    avg = sum(ar) / len(ar)
    if resize_mode == "fixed_mapping":
        keys = sorted(RESOLUTION_MAPPINGS[resolution_set].keys())
        return RESOLUTION_MAPPINGS[resolution_set][min(keys, key=lambda x: abs(x - avg))]
    if resize_mode == "square":
        return (round(size // patch_size) * patch_size,) * 2
    if resize_mode == "longest_side":
        if avg >= 1:
            return (size, round((size // patch_size) / avg) * patch_size)
        return (round((size // patch_size) * avg) * patch_size, size)
    if resize_mode == "fixed_size":
        return ((size[0] // patch_size) * patch_size, (size[1] // patch_size) * patch_size)
    raise ValueError(resize_mode)


def preprocess_image(img, target_w, target_h, mean=DINOV2_MEAN, std=DINOV2_STD):
    """uint8 [H1, W1, 3] -> (u8 crop [target_h, target_w, 3], normalised fp32 [3, target_h, target_w])."""
    h1, w1 = img.shape[:2]
    rw, rh, filt, left, top = resize_plan(w1, h1, target_w, target_h)
    res = pil_resize(img, rw, rh, filt)
    crop = res[top:top + target_h, left:left + target_w]
    x = crop.astype(np.float32).transpose(2, 0, 1) / np.float32(255)          # ToTensor
    m = np.asarray(mean, np.float32).reshape(3, 1, 1)
    s = np.asarray(std, np.float32).reshape(3, 1, 1)
    return crop, ((x - m) / s).astype(np.float32)                            # Normalize


# ---- multi-modal inputs: preprocess_inputs (image.py:335-675) --------------------------------------
def cv2_resize_nearest(src, out_w, out_h):
    """cv2.resize(src, (out_w, out_h), interpolation=cv2.INTER_NEAREST) -- a THIRD-PARTY dependency
    (opencv-python, unpinned by the reference): OpenCV imgproc resize.cpp `resizeNN`: source index
    min(floor(dst * (1 / (dsize / ssize))), ssize - 1) per axis, in double.  Pinned against the live cv2."""
    h, w = src.shape[:2]
    ifx, ify = 1.0 / (out_w / w), 1.0 / (out_h / h)
    xs = np.minimum(np.floor(np.arange(out_w) * ifx).astype(np.int64), w - 1)
    ys = np.minimum(np.floor(np.arange(out_h) * ify).astype(np.int64), h - 1)
    return src[ys][:, xs]


def crop_plan(w1, h1, target_w, target_h, intrinsics=None):
    """cropping.py:230-283 + :429-447 -> (resize_w, resize_h, filter, left, top, new_intrinsics or None)."""
    inp = np.array([w1, h1])
    scale_final = max(np.array([target_w, target_h]) / inp) + 1e-8
    out_res = np.floor(inp * scale_final).astype(int)
    rw, rh = int(out_res[0]), int(out_res[1])
    filt = "lanczos" if scale_final < 1 else "bicubic"
    if intrinsics is None:
        return rw, rh, filt, (rw - target_w) // 2, (rh - target_h) // 2, None

    def matrix_of_crop(K, in_res, o_res, scaling=1, offset_factor=0.5):  # cropping.py:283-318
        margins = np.asarray(in_res) * scaling - o_res
        assert np.all(margins >= 0.0)
        offset = offset_factor * margins
        Kc = K.copy(); Kc[0, 2] += 0.5; Kc[1, 2] += 0.5         # opencv_to_colmap_intrinsics
        Kc[:2, :] *= scaling
        Kc[:2, 2] -= offset
        Kc[0, 2] -= 0.5; Kc[1, 2] -= 0.5                         # colmap_to_opencv_intrinsics
        return Kc

    K = matrix_of_crop(np.asarray(intrinsics), inp, out_res, scaling=scale_final)      # rescale, :268-272
    Knew = matrix_of_crop(K, (rw, rh), (target_w, target_h), offset_factor=0.5)         # :430-435
    left, top = (int(v) for v in np.int32(np.round(K[:2, 2] - Knew[:2, 2])))            # :375-378
    Kout = K.copy(); Kout[0, 2] -= left; Kout[1, 2] -= top                              # :360-363
    return rw, rh, filt, left, top, Kout


def preprocess_view(img, target_w, target_h, depth=None, intrinsics=None, mean=DINOV2_MEAN, std=DINOV2_STD):
    """one view of preprocess_inputs: -> (normalised img [3,th,tw], depth [th,tw] or None, K or None)"""
    h1, w1 = img.shape[:2]
    rw, rh, filt, left, top, K = crop_plan(w1, h1, target_w, target_h, intrinsics)
    crop = pil_resize(img, rw, rh, filt)[top:top + target_h, left:left + target_w]
    x = crop.astype(np.float32).transpose(2, 0, 1) / np.float32(255)
    out = ((x - np.asarray(mean, np.float32).reshape(3, 1, 1)) / np.asarray(std, np.float32).reshape(3, 1, 1))
    d = None
    if depth is not None:
        d = cv2_resize_nearest(np.asarray(depth), rw, rh)[top:top + target_h, left:left + target_w]
    return out.astype(np.float32), d, K
