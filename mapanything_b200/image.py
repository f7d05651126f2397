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
