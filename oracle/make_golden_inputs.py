"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/inputs_*.npz from the LIVE reference.

    python -m oracle.make_golden_inputs       (needs /root/reference; run in the build container)

Runs mapanything.utils.image.preprocess_inputs (image.py:335-675) on seeded synthetic views that
carry an image, a depth map and pinhole intrinsics (off-centre principal point: the crop then follows
the intrinsics, cropping.py:429-440), and stores inputs and outputs.
"""
import os

import numpy as np
import torch

from oracle import ref_shim
from oracle.make_golden_image import synthetic_photo

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CASES = {  # name -> (list of (h, w, fx, fy, cx, cy)), kwargs
    "inputs_shrink_intrinsics": ([(240, 330, 270.0, 268.0, 171.3, 113.9), (240, 330, 265.0, 265.0, 164.5, 119.5)],
                                 dict(resize_mode="fixed_size", size=(196, 140))),
    "inputs_up_intrinsics": ([(60, 96, 80.0, 82.0, 50.2, 27.7)], dict(resize_mode="longest_side", size=126)),
    "inputs_depth_no_intrinsics": ([(200, 260, 0, 0, 0, 0)], dict(resize_mode="square", size=112)),
    # ray_directions instead of intrinsics (image.py:572-590); the margin of the crop is even, so the
    # crop offset does not sit on a rounding tie of the recovered principal point
    "inputs_rays": ([(120, 200, 150.0, 148.0, 101.3, 58.6)], dict(resize_mode="fixed_size", size=(98, 70))),
}
RAY_CASES = {"inputs_rays"}


def case_views(name):
    specs, _ = CASES[name]
    views = []
    for i, (h, w, fx, fy, cx, cy) in enumerate(specs):
        rng = np.random.default_rng(500 + i)
        v = {"img": torch.from_numpy(synthetic_photo(h, w, 40 + i)),
             "depth_z": torch.from_numpy((rng.random((h, w)) * 5 + 0.5).astype(np.float32))}
        if fx:
            K = torch.tensor([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], dtype=torch.float32)
            if name in RAY_CASES:  # unit rays of that pinhole camera (geometry.py:186-241), restated
                ys, xs = torch.meshgrid(torch.arange(h).float(), torch.arange(w).float(), indexing="ij")
                d = torch.stack([(xs - cx) / fx, (ys - cy) / fy, torch.ones_like(xs)], -1)
                v["ray_directions"] = d / d.norm(dim=-1, keepdim=True)
            else:
                v["intrinsics"] = K
        views.append(v)
    return views


def main():
    ref_shim.install()
    from mapanything.utils.image import preprocess_inputs

    for name, (specs, kw) in CASES.items():
        views = case_views(name)
        out = preprocess_inputs([dict(v) for v in views], **kw)
        store = {}
        for i, (v, o) in enumerate(zip(views, out)):
            store[f"img{i}"] = v["img"].numpy(); store[f"depth{i}"] = v["depth_z"].numpy()
            if "intrinsics" in v:
                store[f"K{i}"] = v["intrinsics"].numpy(); store[f"out_K{i}"] = o["intrinsics"][0].numpy()
            if "ray_directions" in v:
                store[f"out_K{i}"] = o["intrinsics"][0].numpy()
            store[f"out_img{i}"] = o["img"][0].numpy(); store[f"out_depth{i}"] = o["depth_z"][0].numpy()
        path = os.path.join(ROOT, "tests", "golden", name + ".npz")
        np.savez_compressed(path, **store)
        print("wrote", path, os.path.getsize(path) // 1024, "KiB", tuple(out[0]["img"].shape), tuple(out[0]["depth_z"].shape))


if __name__ == "__main__":
    main()
