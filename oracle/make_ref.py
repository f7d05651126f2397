"""TEST / BENCH INFRASTRUCTURE -- makes the UNMODIFIED reference importable on the GPU box.

/root/reference exists only in the build container.  This script copies the two Python packages the
hot path imports (`mapanything/`, `uniception/`, ~3 MB of .py files) from where they lie under
/root/reference into the git-ignored `oracle/_ref/` (listed in .gitignore, not in .gpurunignore: it
travels to the GPU box with the snapshot, it never enters history).  `oracle/ref_shim.py` then builds
the reference `MapAnything` from that tree (MAB_REFERENCE_ROOT), which is what `bench.py --impl reference`
(CPU arm, kind "reference") and bench.py's `reference_gpu` leg (the reference's own eager bf16-autocast
GPU path: library SDPA / cuBLASLt / cuDNN) time.  Called by `__graft_entry__.build()` whenever
/root/reference is present; a no-op elsewhere.

    python -m oracle.make_ref
"""
import os
import shutil

SRC = "/root/reference"
DST = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")
PACKAGES = ("mapanything", "uniception")


def make(force=False):
    if not os.path.isdir(os.path.join(SRC, "mapanything")):
        return None
    for pkg in PACKAGES:
        dst = os.path.join(DST, pkg)
        if os.path.isdir(dst):
            if not force:
                continue
            shutil.rmtree(dst)
        shutil.copytree(os.path.join(SRC, pkg), dst,
                        ignore=shutil.ignore_patterns("__pycache__", "*.pyc", "*.so", "*.o", "build", "*.egg-info"))
    return DST


def ref_root():
    """Where the reference can be imported from on this machine, or None."""
    for root in (SRC, DST):
        if os.path.isdir(os.path.join(root, "mapanything")) and os.path.isdir(os.path.join(root, "uniception")):
            return root
    return None


if __name__ == "__main__":
    print(make(force=True))
