"""Builds GEMM-epilogue variants of libmab200 into build_variants/ (kernel experiments, loaded with
MAB_LIB=<path>; profiles/gemm_variants.sh runs the parity tests and the shape microbench on each).
    python profiles/build_gemm_variants.py [name ...]"""
import os
import sys
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mapanything_b200 import build  # noqa: E402

VARIANTS = {
    "default": [],                                # lean EPI_PLAIN / EPI_RESID kernels (MAB_NO_EPI_RED=1 /
                                                  # MAB_NO_EPI_PLAIN=1 in the environment route back to the generic one)
    "tanhgelu": ["MAB_GELU_TANH"],                # GELU through one tanh.approx (faster, less accurate: not used)
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
