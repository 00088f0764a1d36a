"""Build the checker: the C parts of the oracle (gcc) into oracle/_build/libsb_oracle.so, and - when the reference tree
is present (the build container) - a verbatim copy of the few reference files the CPU baseline and the dispatch tests
run, into the git-ignored ``oracle/_ref/`` so that they travel to the GPU box with the snapshot.

TEST INFRASTRUCTURE ONLY.  The reference is pure Python on this path: there is nothing to compile for ``oracle/_ref``;
its files are copied byte for byte (never edited, never committed - ``oracle/_ref/`` is in .gitignore) and imported
through ``oracle/ref_shim.py``.  ``bench.py --impl reference`` times ``CoreMahalanobis.predict`` from there
(``cpu_baseline.kind = "reference"``); ``tests/test_gpu_dispatch.py`` drives the drop-in classes through the
reference's own ``user_model.client_*`` functions from there.
"""
import filecmp
import os
import shutil
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
SOURCES = ["maha_lean.c", "routers.c"]
REFERENCE_ROOT = os.environ.get("SB200_REFERENCE_ROOT", "/root/reference")
REF_DIR = os.path.join(_HERE, "_ref")
# reference files on the hot path that run as they are under this image's Python
REF_FILES = [
    "components/outlier-detection/mahalanobis/CoreMahalanobis.py",
    "components/routers/epsilon-greedy/EpsilonGreedy.py",
    "components/routers/thompson-sampling/ThompsonSampling.py",
    "examples/combiners/mnist_combiner/MnistCombiner.py",
    "python/seldon_core/user_model.py",
    "python/seldon_core/metrics.py",
]


def build(force=False):
    out_dir = os.path.join(_HERE, "_build")
    os.makedirs(out_dir, exist_ok=True)
    out = os.path.join(out_dir, "libsb_oracle.so")
    srcs = [os.path.join(_HERE, s) for s in SOURCES if os.path.isfile(os.path.join(_HERE, s))]
    if force or not os.path.isfile(out) or any(os.path.getmtime(out) < os.path.getmtime(s) for s in srcs):
        cmd = ["gcc", "-O2", "-fPIC", "-shared", "-std=c11", "-ffp-contract=off", "-o", out] + srcs + ["-lm"]
        subprocess.check_call(cmd)
    build_ref()
    return out


def build_ref():
    """Copy REF_FILES from the reference tree into oracle/_ref/ (same relative paths).  A no-op where the tree is
    absent (the GPU box uses the copy that came with the snapshot).  Returns the number of files present."""
    if not os.path.isdir(REFERENCE_ROOT):
        return sum(os.path.isfile(os.path.join(REF_DIR, f)) for f in REF_FILES)
    n = 0
    for rel in REF_FILES:
        src, dst = os.path.join(REFERENCE_ROOT, rel), os.path.join(REF_DIR, rel)
        if not os.path.isfile(src):
            continue
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        if not (os.path.isfile(dst) and filecmp.cmp(src, dst, shallow=False)):
            shutil.copyfile(src, dst)
        n += 1
    return n


if __name__ == "__main__":
    print(build(force=True), build_ref(), "reference files in", REF_DIR)
