"""Recipe for oracle/_ref: a snapshot of the reference's OWN Python for the hot path, so that the reference arm of
bench.py (`--impl reference`, `cpu_baseline.kind = "reference"`) can time the UNMODIFIED reference on the GPU box's host
cores, where /root/reference does not exist.  TEST / MEASUREMENT INFRASTRUCTURE.

    python -m oracle.build_ref        (also run by __graft_entry__.build() when /root/reference is present)

The reference is pure Python (no build step): "building" it is copying the packages its `nets` registry imports
(nets/, data_utils/, trainer/, losses/, evaluation/, config/ -- 1.8 MB) from where they lie under /root/reference into
oracle/_ref/reference/.  oracle/_ref/ is git-ignored (reference sources never enter this repository's history) but not
gpurun-ignored, so it travels to the GPU box like the built .so files.  Nothing is edited: oracle/ref_harness.py applies
its shims at import time.
"""
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.environ.get("PROBTALK_REFERENCE", "/root/reference")
DST = os.path.join(ROOT, "oracle", "_ref", "reference")
PACKAGES = ("nets", "data_utils", "trainer", "losses", "evaluation", "config")


def snapshot(force=False):
    """Copy the packages; returns the destination or None when the reference tree is not present."""
    if not os.path.isdir(os.path.join(SRC, "nets")):
        return None
    stamp = os.path.join(DST, ".snapshot_of")
    if not force and os.path.exists(stamp) and open(stamp).read().strip() == SRC:
        return DST
    if os.path.isdir(DST):
        shutil.rmtree(DST)
    os.makedirs(DST)
    ignore = shutil.ignore_patterns("__pycache__", "*.pyc", "*.pth", "*.pkl", "*.wav", "*.mp4")
    for pkg in PACKAGES:
        if os.path.isdir(os.path.join(SRC, pkg)):
            shutil.copytree(os.path.join(SRC, pkg), os.path.join(DST, pkg), ignore=ignore)
    with open(stamp, "w") as f:
        f.write(SRC)
    return DST


if __name__ == "__main__":
    d = snapshot(force="--force" in sys.argv)
    print("snapshot:", d if d else "reference tree not found at %s" % SRC)
