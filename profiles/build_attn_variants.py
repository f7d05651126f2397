"""Builds attention-kernel variants of libmab200 into build_variants/ (kernel experiments; loaded with
MAB_LIB=<path>, see profiles/attn_variants.sh).  python profiles/build_attn_variants.py [name ...]"""
import os
import sys
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mapanything_b200 import build  # noqa: E402

VARIANTS = {
    "split1_turns": ["ATT_SPLIT=1", "ATT_NO_TURNS=0"],   # the round-1 kernel
    "split2_turns": ["ATT_NO_TURNS=0"],
    "split2_noturn_p0": ["ATT_POLY_MASK=0", "ATT_POLY_MASK_SHORT=0"],
    "split2_noturn_p3": ["ATT_POLY_MASK=0x0421"],
    "split2_noturn_p4": ["ATT_POLY_MASK=0x1111"],
    "split2_noturn_p6": ["ATT_POLY_MASK=0x2949"],
    "split2_noturn_p8": ["ATT_POLY_MASK=0x5555"],
}


def one(name):
    out = os.path.join(ROOT, "build_variants", f"libmab200_{name}.so")
    build.build(out=out, defines=VARIANTS[name])
    return out


if __name__ == "__main__":
    os.makedirs(os.path.join(ROOT, "build_variants"), exist_ok=True)
    with ThreadPoolExecutor(max_workers=6) as ex:
        for path in ex.map(one, sys.argv[1:] or list(VARIANTS)):
            print("built", path)
