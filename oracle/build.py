"""Compile the C parts of the oracle (gcc) into oracle/_build/libsb_oracle.so.

TEST INFRASTRUCTURE ONLY.  The reference itself is pure Python, so there is no
``oracle/_ref`` binary to build: "reference" runs are done by importing
``/root/reference`` through ``oracle/ref_shim.py`` in the build container.
"""
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
SOURCES = ["maha_lean.c", "routers.c"]


def build(force=False):
    out_dir = os.path.join(_HERE, "_build")
    os.makedirs(out_dir, exist_ok=True)
    out = os.path.join(out_dir, "libsb_oracle.so")
    srcs = [os.path.join(_HERE, s) for s in SOURCES if os.path.isfile(os.path.join(_HERE, s))]
    if not force and os.path.isfile(out) and all(os.path.getmtime(out) >= os.path.getmtime(s) for s in srcs):
        return out
    cmd = ["gcc", "-O2", "-fPIC", "-shared", "-std=c11", "-ffp-contract=off", "-o", out] + srcs + ["-lm"]
    subprocess.check_call(cmd)
    return out


if __name__ == "__main__":
    print(build(force=True))
