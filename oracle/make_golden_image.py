"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/image_*.npz from the LIVE reference + LIVE Pillow.

    python -m oracle.make_golden_image        (needs /root/reference; run in the build container)

Seeded synthetic photographs (smooth gradients + texture + noise) are written as PNG files into a
temporary folder and loaded by the reference's own `mapanything.utils.image.load_images`
(image.py:134-332); the decoded uint8 inputs and the normalised outputs are stored.
"""
import os
import tempfile

import numpy as np
import PIL.Image

from oracle import ref_shim

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def synthetic_photo(h, w, seed):
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    img = np.stack([127 + 120 * np.sin(xx / (7.0 + c) + yy / 13.0) * np.cos(yy / (5.0 + 2 * c)) for c in range(3)], -1)
    img += rng.normal(0, 25, img.shape)
    img[h // 4: h // 3, w // 5: w // 2] = 255   # saturated patches: exercise the clip
    img[h // 2: h // 2 + 9, :] = 0
    return np.clip(img, 0, 255).astype(np.uint8)


CASES = {  # name -> (list of (h, w)), load_images kwargs
    "image_fixedmap_lanczos": ([(480, 640)], dict()),                             # 4:3 -> 518x392, shrinking
    "image_lanczos_hd": ([(432, 768)], dict(resize_mode="fixed_size", size=(294, 168))),
    "image_bicubic_up": ([(60, 90), (64, 88)], dict(resize_mode="longest_side", size=126)),
    "image_square_crop": ([(210, 330)], dict(resize_mode="square", size=112)),
}


def main():
    ref_shim.install()
    from mapanything.utils.image import load_images

    for name, (sizes, kw) in CASES.items():
        with tempfile.TemporaryDirectory() as d:
            srcs = []
            for i, (h, w) in enumerate(sizes):
                im = synthetic_photo(h, w, 100 + i)
                PIL.Image.fromarray(im).save(os.path.join(d, f"{i:03d}.png"))
                srcs.append(im)
            views = load_images(d, **kw)
        store = {f"src{i}": s for i, s in enumerate(srcs)}
        store.update({f"out{i}": v["img"][0].numpy() for i, v in enumerate(views)})
        store["true_shape"] = np.stack([v["true_shape"][0] for v in views])
        path = os.path.join(ROOT, "tests", "golden", name + ".npz")
        np.savez_compressed(path, **store)
        print("wrote", path, os.path.getsize(path) // 1024, "KiB", [tuple(v["img"].shape) for v in views])


if __name__ == "__main__":
    main()
