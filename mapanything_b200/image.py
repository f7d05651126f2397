"""Image input path on the GPU.

Host-side mirror of `mapanything.utils.image.load_images` (image.py:134-332): same arguments, same
target-size selection (`find_closest_aspect_ratio` + `RESOLUTION_MAPPINGS`, image.py:40-91,248-284),
same view dicts.  Files are decoded on the host with PIL exactly as the reference does (image.py:226-233);
everything after decoding -- PIL LANCZOS/BICUBIC resize, centred crop, ToTensor, Normalize
(cropping.py:230-244,441-447, image.py:297-300) -- runs on the device through libmab200's
`mab_preprocess_image`, bit-exact with Pillow's 8-bit resampling.  No CPU fallback for that part.
"""
import ctypes as C
import os
from typing import List, Optional, Sequence, Tuple, Union

import numpy as np
import torch

from . import _lib

# image.py:40-66
RESOLUTION_MAPPINGS = {
    518: {1.000: (518, 518), 1.321: (518, 392), 1.542: (518, 336), 1.762: (518, 294), 2.056: (518, 252),
          3.083: (518, 168), 0.757: (392, 518), 0.649: (336, 518), 0.567: (294, 518), 0.486: (252, 518)},
    512: {1.000: (512, 512), 1.333: (512, 384), 1.524: (512, 336), 1.778: (512, 288), 2.000: (512, 256),
          3.200: (512, 160), 0.750: (384, 512), 0.656: (336, 512), 0.562: (288, 512), 0.500: (256, 512)},
}
# uniception/models/encoders/image_normalizations.py:17-40
_IMAGENET = ((0.485, 0.456, 0.406), (0.229, 0.224, 0.225))
IMAGE_NORMALIZATION_DICT = {
    "dummy": ((0.0, 0.0, 0.0), (1.0, 1.0, 1.0)),
    "croco": _IMAGENET,
    "dust3r": ((0.5, 0.5, 0.5), (0.5, 0.5, 0.5)),
    "dinov2": _IMAGENET,
    "identity": ((0.0, 0.0, 0.0), (1.0, 1.0, 1.0)),
    "patch_embedder": _IMAGENET,
    "radio": ((0.0, 0.0, 0.0), (1.0, 1.0, 1.0)),
    "sea_raft": ((0.0, 0.0, 0.0), (1.0 / 255, 1.0 / 255, 1.0 / 255)),
    "unimatch": ((0.0, 0.0, 0.0), (1.0 / 255, 1.0 / 255, 1.0 / 255)),
    "roma": _IMAGENET,
    "cosmos": ((0.0, 0.0, 0.0), (0.5, 0.5, 0.5)),
}


def find_closest_aspect_ratio(aspect_ratio: float, resolution_set: int) -> Tuple[int, int]:
    """image.py:74-90"""
    keys = sorted(RESOLUTION_MAPPINGS[resolution_set].keys())
    return RESOLUTION_MAPPINGS[resolution_set][min(keys, key=lambda x: abs(x - aspect_ratio))]


def _target_size(sizes, resize_mode, size, patch_size, resolution_set):
    """image.py:240-284: one (W, H) for all images from their average aspect ratio."""
    aspect = [w / h for w, h in sizes]
    avg = sum(aspect) / len(aspect)
    if resize_mode == "fixed_mapping":
        return find_closest_aspect_ratio(avg, resolution_set)
    if resize_mode == "square":
        return (round((size // patch_size)) * patch_size,) * 2
    if resize_mode == "longest_side":
        if avg >= 1:
            return (size, round((size // patch_size) / avg) * patch_size)
        return (round((size // patch_size) * avg) * patch_size, size)
    return ((size[0] // patch_size) * patch_size, (size[1] // patch_size) * patch_size)  # fixed_size


def _validate(resize_mode, size):
    """image.py:164-190 (same messages' conditions, same exception type)"""
    valid = ["fixed_mapping", "longest_side", "square", "fixed_size"]
    if resize_mode not in valid:
        raise ValueError(f"Resize_mode must be one of {valid}, got '{resize_mode}'")
    if resize_mode in ["longest_side", "square", "fixed_size"] and size is None:
        raise ValueError(f"Size parameter is required for resize_mode='{resize_mode}'")
    if resize_mode in ["longest_side", "square"] and not isinstance(size, int):
        raise ValueError(f"Size must be an int for resize_mode='{resize_mode}', got {type(size)}")
    if resize_mode == "fixed_size":
        if not isinstance(size, (tuple, list)) or len(size) != 2:
            raise ValueError(f"Size must be a tuple/list of (width, height) for resize_mode='fixed_size', got {size}")
        if not all(isinstance(x, int) for x in size):
            raise ValueError(f"Size values must be integers for resize_mode='fixed_size', got {size}")


def preprocess_decoded(images: Sequence[Union[np.ndarray, torch.Tensor]], resize_mode="fixed_mapping", size=None,
                       norm_type="dinov2", patch_size=14, resolution_set=518, device="cuda",
                       return_u8=False) -> List[dict]:
    """Second pass of `load_images` (image.py:286-323) for already-decoded RGB uint8 [H, W, 3] images
    (numpy or torch, host or device): resize + crop + normalise on the device, one view dict per image."""
    _validate(resize_mode, size)
    if norm_type not in IMAGE_NORMALIZATION_DICT:
        raise ValueError(f"Unknown image normalization type: {norm_type}. "
                         f"Available options: {list(IMAGE_NORMALIZATION_DICT.keys())}")
    if len(images) == 0:
        raise ValueError("No valid images found")
    lib = _lib.load()
    dev = torch.device(device)
    if dev.type != "cuda":
        raise RuntimeError("mapanything_b200.image runs on a CUDA device only (no CPU path)")
    sizes = [(int(im.shape[1]), int(im.shape[0])) for im in images]
    tw, th = _target_size(sizes, resize_mode, size, patch_size, resolution_set)
    mean, std = IMAGE_NORMALIZATION_DICT[norm_type]
    c_mean, c_std = (C.c_float * 3)(*mean), (C.c_float * 3)(*std)
    st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    out = []
    for im in images:
        t = torch.as_tensor(im)
        assert t.dtype == torch.uint8 and t.dim() == 3 and t.shape[2] == 3, "RGB uint8 [H, W, 3] expected"
        t = t.to(dev, non_blocking=True).contiguous()
        plan = _lib.MabImagePlan()
        _lib.check(lib.mab_image_plan_make(int(t.shape[1]), int(t.shape[0]), tw, th, C.byref(plan)), "mab_image_plan_make")
        ws = torch.empty(lib.mab_preprocess_workspace_bytes(C.byref(plan)), dtype=torch.uint8, device=dev)
        img = torch.empty((1, 3, th, tw), dtype=torch.float32, device=dev)
        u8 = torch.empty((th, tw, 3), dtype=torch.uint8, device=dev) if return_u8 else None
        _lib.check(lib.mab_preprocess_image(C.c_void_p(t.data_ptr()), C.byref(plan), c_mean, c_std,
                                            C.c_void_p(img.data_ptr()), C.c_void_p(u8.data_ptr()) if return_u8 else None,
                                            C.c_void_p(ws.data_ptr()), ws.numel(), st), "mab_preprocess_image")
        view = dict(img=img, true_shape=np.int32([[th, tw]]), idx=len(out), instance=str(len(out)),
                    data_norm_type=[norm_type])
        if return_u8:
            view["img_u8"] = u8
        out.append(view)
    return out


def load_images(folder_or_list, resize_mode="fixed_mapping", size=None, norm_type="dinov2", patch_size=14,
                verbose=False, bayer_format=False, resolution_set=518, stride=1, device="cuda") -> List[dict]:
    """`mapanything.utils.image.load_images` (image.py:134-332) with the per-image work on the device.
    Same arguments and return value; `img` tensors live on `device`."""
    import PIL.Image
    from PIL.ImageOps import exif_transpose

    _validate(resize_mode, size)
    if isinstance(folder_or_list, str):
        root, content = folder_or_list, sorted(os.listdir(folder_or_list))
    elif isinstance(folder_or_list, list):
        root, content = "", folder_or_list
    else:
        raise ValueError(f"Bad {folder_or_list=} ({type(folder_or_list)})")
    exts = (".jpg", ".jpeg", ".png")
    decoded = []
    for i, path in enumerate(content):
        if i % stride != 0 or not path.lower().endswith(exts):
            continue
        try:
            if bayer_format:
                import cv2
                color = cv2.cvtColor(cv2.imread(os.path.join(root, path), cv2.IMREAD_UNCHANGED), cv2.COLOR_BAYER_RG2BGR)
                img = exif_transpose(PIL.Image.fromarray(color)).convert("RGB")
            else:
                img = exif_transpose(PIL.Image.open(os.path.join(root, path))).convert("RGB")
            decoded.append(np.asarray(img))
        except Exception as e:  # image.py:239-242
            if verbose:
                print(f"Warning: Could not load {path}: {e}")
    if not decoded:
        raise ValueError("No valid images found")
    return preprocess_decoded(decoded, resize_mode, size, norm_type, patch_size, resolution_set, device)


# ---- preprocess_inputs (image.py:335-675): images + depth + intrinsics, cropped consistently ---------
def _crop_plan(w1, h1, tw, th, intrinsics):
    """cropping.py:230-283 (rescale) + :429-447 (crop that follows the principal point when intrinsics are
    given, centred otherwise) + :360-363 (intrinsics of the crop).  3x3 host arithmetic in numpy, as the
    reference does it.  -> (resize_w, resize_h, left, top, K_out or None)"""
    inp = np.array([w1, h1])
    scale_final = max(np.array([tw, th]) / inp) + 1e-8
    out_res = np.floor(inp * scale_final).astype(int)
    rw, rh = int(out_res[0]), int(out_res[1])
    if intrinsics is None:
        return rw, rh, (rw - tw) // 2, (rh - th) // 2, None

    def matrix_of_crop(K, in_res, o_res, scaling=1, offset_factor=0.5):
        margins = np.asarray(in_res) * scaling - o_res
        assert np.all(margins >= 0.0)
        Kc = K.copy()
        Kc[0, 2] += 0.5; Kc[1, 2] += 0.5
        Kc[:2, :] *= scaling
        Kc[:2, 2] -= offset_factor * margins
        Kc[0, 2] -= 0.5; Kc[1, 2] -= 0.5
        return Kc

    K = matrix_of_crop(np.asarray(intrinsics), inp, out_res, scaling=scale_final)
    Knew = matrix_of_crop(K, (rw, rh), (tw, th))
    left, top = (int(v) for v in np.int32(np.round(K[:2, 2] - Knew[:2, 2])))
    Kout = K.copy()
    Kout[0, 2] -= left; Kout[1, 2] -= top
    return rw, rh, left, top, Kout


def _to_u8(img, view_idx, dev):
    """image.py:493-520 `to_pil_image`: tensor / array (H, W, 3) in [0, 1] or [0, 255], or a PIL image."""
    import PIL.Image

    if isinstance(img, PIL.Image.Image):
        return torch.from_numpy(np.asarray(img.convert("RGB") if img.mode != "RGB" else img)).to(dev)
    if isinstance(img, torch.Tensor):
        if img.ndim != 3 or img.shape[2] != 3:
            raise ValueError(f"Expected tensor shape (H, W, 3) for img in view {view_idx}, got {img.shape}")
        img = img.to(dev)
        return ((img * 255) if img.max() <= 1.0 else img).clamp(0, 255).byte()
    if isinstance(img, np.ndarray):
        if img.ndim != 3 or img.shape[2] != 3:
            raise ValueError(f"Expected array shape (H, W, 3) for img in view {view_idx}, got {img.shape}")
        if img.dtype != np.uint8:
            img = (img * 255).clip(0, 255).astype(np.uint8)
        return torch.from_numpy(img).to(dev)
    raise ValueError(f"Unsupported image type in view {view_idx}: {type(img)}")


def preprocess_inputs(input_views, resize_mode="fixed_mapping", size=None, norm_type="dinov2", patch_size=14,
                      resolution_set=518, verbose=False, device="cuda") -> List[dict]:
    """`mapanything.utils.image.preprocess_inputs` (image.py:335-675): same arguments, same view dicts;
    images (Pillow-exact resize + crop + normalise) and depth maps (cv2.INTER_NEAREST + crop) on the device;
    views that carry `ray_directions` get their intrinsics recovered by least squares (image.py:572-590,
    `mab_recover_intrinsics`)."""
    _validate(resize_mode, size)
    if not input_views:
        raise ValueError("input_views cannot be empty")
    if norm_type not in IMAGE_NORMALIZATION_DICT:
        raise ValueError(f"Unknown image normalization type: {norm_type}. "
                         f"Available options: {list(IMAGE_NORMALIZATION_DICT.keys())}")
    dev = torch.device(device)
    if dev.type != "cuda":
        raise RuntimeError("mapanything_b200.image runs on a CUDA device only (no CPU path)")
    for i, v in enumerate(input_views):
        if "img" not in v:
            raise ValueError(f"View {i} missing required 'img' key")
        if "intrinsics" in v and "ray_directions" in v:
            raise ValueError(f"View {i} cannot have both 'intrinsics' and 'ray_directions'. Please provide only one "
                             "as they are redundant (ray_directions can be used to recover intrinsics).")
    lib = _lib.load()
    u8s = [_to_u8(v["img"], i, dev).contiguous() for i, v in enumerate(input_views)]
    tw, th = _target_size([(int(t.shape[1]), int(t.shape[0])) for t in u8s], resize_mode, size, patch_size,
                          resolution_set)
    mean, std = IMAGE_NORMALIZATION_DICT[norm_type]
    c_mean, c_std = (C.c_float * 3)(*mean), (C.c_float * 3)(*std)
    st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    out = []
    for i, (view, u8) in enumerate(zip(input_views, u8s)):
        h1, w1 = int(u8.shape[0]), int(u8.shape[1])
        K = view.get("intrinsics")
        if "ray_directions" in view:  # image.py:572-590: intrinsics recovered from the rays by least squares
            rays = torch.as_tensor(view["ray_directions"])
            if rays.ndim != 3 or rays.shape[2] != 3:
                raise ValueError(f"Expected shape (H, W, 3) for ray_directions in view {i}, got {tuple(rays.shape)}")
            rays = rays.to(dev, torch.float32).contiguous()
            Kd = torch.empty((1, 3, 3), dtype=torch.float32, device=dev)
            _lib.check(lib.mab_recover_intrinsics(C.c_void_p(rays.data_ptr()), 1, int(rays.shape[0]), int(rays.shape[1]),
                                                  C.c_void_p(Kd.data_ptr()), st), "mab_recover_intrinsics")
            K = Kd[0].cpu().numpy()
        if K is not None:
            K = K.cpu().numpy() if isinstance(K, torch.Tensor) else K
            if not isinstance(K, np.ndarray) or K.shape != (3, 3):
                raise ValueError(f"Expected shape (3, 3) for intrinsics in view {i}, got {getattr(K, 'shape', type(K))}")
        rw, rh, left, top, Kout = _crop_plan(w1, h1, tw, th, K)
        plan = _lib.MabImagePlan()
        _lib.check(lib.mab_image_plan_make(w1, h1, tw, th, C.byref(plan)), "mab_image_plan_make")
        assert (plan.resize_w, plan.resize_h) == (rw, rh)
        plan.crop_left, plan.crop_top = left, top   # the crop follows the principal point (cropping.py:429-440)
        ws = torch.empty(lib.mab_preprocess_workspace_bytes(C.byref(plan)), dtype=torch.uint8, device=dev)
        img = torch.empty((1, 3, th, tw), dtype=torch.float32, device=dev)
        _lib.check(lib.mab_preprocess_image(C.c_void_p(u8.data_ptr()), C.byref(plan), c_mean, c_std,
                                            C.c_void_p(img.data_ptr()), None, C.c_void_p(ws.data_ptr()), ws.numel(), st),
                   "mab_preprocess_image")
        pv = {"img": img, "data_norm_type": [norm_type]}
        if "depth_z" in view:
            d = view["depth_z"]
            d = torch.as_tensor(d)
            if d.ndim != 2:
                raise ValueError(f"Expected shape (H, W) for depth_z in view {i}, got {tuple(d.shape)}")
            assert tuple(d.shape) == (h1, w1)  # cropping.py:221
            d = d.to(dev, torch.float32).contiguous()
            dz = torch.empty((1, th, tw), dtype=torch.float32, device=dev)
            _lib.check(lib.mab_resize_nearest_crop(C.c_void_p(d.data_ptr()), h1, w1, rw, rh, left, top, tw, th,
                                                   C.c_void_p(dz.data_ptr()), st), "mab_resize_nearest_crop")
            pv["depth_z"] = dz
        if Kout is not None:
            pv["intrinsics"] = torch.from_numpy(Kout)[None].to(dev)
        if "camera_poses" in view:  # image.py:621-652: batch dimension only
            cp = view["camera_poses"]
            if isinstance(cp, tuple):
                pv["camera_poses"] = tuple(torch.as_tensor(c)[None] for c in cp)
            elif isinstance(cp, (torch.Tensor, np.ndarray)):
                pv["camera_poses"] = torch.as_tensor(cp)[None]
            else:
                raise ValueError(f"Unsupported camera_poses format: {type(cp)}. Expected tuple (quats, trans) or "
                                 "matrix (tensor/array).")
        for k, val in view.items():
            if k not in ("img", "depth_z", "intrinsics", "ray_directions", "camera_poses"):
                pv[k] = val
        out.append(pv)
    return out
