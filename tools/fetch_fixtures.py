#!/usr/bin/env python
"""Copies the reference's fixture images (dataset/dataset/*.jpg, the cards its acceptance test
walks, packages/lib/src/lib.rs:131-245) into tests/fixtures/dataset/ (git-ignored; see
tests/fixtures/README.md).  A no-op when the reference tree is absent (the GPU box)."""
import glob
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.environ.get("OICS_REFERENCE_DATASET", "/root/reference/dataset/dataset")
DST = os.path.join(ROOT, "tests", "fixtures", "dataset")


def fetch(verbose: bool = False) -> int:
    files = sorted(glob.glob(os.path.join(SRC, "*.jpg")))
    if not files:
        return 0
    os.makedirs(DST, exist_ok=True)
    n = 0
    for f in files:
        d = os.path.join(DST, os.path.basename(f))
        if not os.path.exists(d) or os.path.getsize(d) != os.path.getsize(f):
            shutil.copyfile(f, d)
            n += 1
    if verbose:
        print(f"{len(files)} fixture images in {DST} ({n} copied)")
    return len(files)


if __name__ == "__main__":
    sys.exit(0 if fetch(verbose=True) or True else 1)
