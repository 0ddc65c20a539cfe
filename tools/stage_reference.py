#!/usr/bin/env python
"""Stage the UNMODIFIED reference's hot-path packages into baseline/_ref/ (git-ignored, NOT gpurun-ignored, so the copy
travels to the GPU box with the snapshot): bench.py then times the reference's own update_natural_param_batch as the
CPU baseline (cpu_baseline.kind = "reference") instead of the oracle port.  The reference has no setup.py /
pyproject.toml, so `pip install /root/reference` fails ("neither 'setup.py' nor 'pyproject.toml' found"); this copy is
the equivalent of that install.  Nothing under baseline/_ref is committed or imported by the product path.

    python tools/stage_reference.py [/root/reference]
"""
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PACKAGES = ["SSVI", "Probability", "Model", "Tensor"]


def stage(src="/root/reference", quiet=False):
    dst = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.isdir(os.path.join(src, "SSVI")):
        if not quiet:
            print("reference tree not found at %s: nothing staged" % src)
        return False
    os.makedirs(dst, exist_ok=True)
    for p in PACKAGES:
        t = os.path.join(dst, p)
        if os.path.isdir(t):
            shutil.rmtree(t)
        shutil.copytree(os.path.join(src, p), t, ignore=shutil.ignore_patterns("__pycache__", "*.pyc"))
    open(os.path.join(dst, "__init__.py"), "a").close()
    if not quiet:
        print("staged %s from %s into %s" % (", ".join(PACKAGES), src, dst))
    return True


if __name__ == "__main__":
    stage(sys.argv[1] if len(sys.argv) > 1 else "/root/reference")
