"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/rope2d_*.npz from the LIVE reference.

    python -m oracle.make_golden_rope        (needs /root/reference; run in the build container)

The reference's pure-PyTorch RoPE2D (uniception/models/libs/croco/pos_embed.py:109-155, the class the
reference falls back to when cuRoPE is not compiled) is run on seeded inputs; tokens there are
(B, heads, N, D), stored here in cuRoPE's (B, N, heads, D) layout (curope.cpp:49-52).
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REFERENCE_ROOT = os.environ.get("MAB_REFERENCE_ROOT", "/root/reference")


def main():
    sys.path.insert(0, REFERENCE_ROOT)
    from uniception.models.libs.croco.pos_embed import RoPE2D  # the torch fallback (cuRoPE is not built)

    assert RoPE2D.__module__.endswith("pos_embed"), "expected the pure-PyTorch RoPE2D"
    cases = {"rope2d_b2_n37_h12_d64": (2, 37, 12, 64, 100.0), "rope2d_b1_n1036_h2_d64": (1, 1036, 2, 64, 100.0),
             "rope2d_b3_n5_h2_d32": (3, 5, 2, 32, 10000.0)}
    for name, (B, N, H, D, base) in cases.items():
        g = torch.Generator().manual_seed(7)
        tokens = torch.randn(B, N, H, D, generator=g)
        if N == 1036:  # the 28 x 37 patch grid of a 392 x 518 view (PositionGetter order: y-major)
            ys, xs = torch.meshgrid(torch.arange(28), torch.arange(37), indexing="ij")
            pos = torch.stack([ys.reshape(-1), xs.reshape(-1)], -1)[None].repeat(B, 1, 1)
        else:
            pos = torch.randint(0, 40, (B, N, 2), generator=g)
        rope = RoPE2D(freq=base)
        out = rope(tokens.permute(0, 2, 1, 3).contiguous(), pos).permute(0, 2, 1, 3).contiguous()
        path = os.path.join(ROOT, "tests", "golden", name + ".npz")
        np.savez_compressed(path, tokens=tokens.numpy(), positions=pos.numpy(), out=out.numpy(),
                            base=np.float32(base))
        print("wrote", path, os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main()
