"""Builds liboics_cuda.so (sm_100a) in-tree with nvcc.  No torch involved.

    python -m omr_img_corrector_b200.build [--force] [--verbose]
"""
from __future__ import annotations

import concurrent.futures as cf
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
SO = os.path.join(HERE, "liboics_cuda.so")
HEADER = os.path.join(os.path.dirname(HERE), "include", "oics_cuda.h")

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
# --fmad=false: the coordinate tables, std-devs and area-resize taps must round every
# product and sum separately, exactly like the reference's un-fused x86 arithmetic.
NVCC_FLAGS = ["-O3", "-std=c++17", "-lineinfo", "--fmad=false", "-Xcompiler", "-fPIC,-ffp-contract=off,-fvisibility=hidden",
              "-Xptxas", "-v"]
# the FFT butterflies are the one place fused multiply-adds are wanted
FMAD_OK = {"fft.cu"}


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def sources():
    return sorted(f for f in os.listdir(CSRC) if f.endswith((".cu", ".cpp")))


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    nvcc = _nvcc()
    os.makedirs(OBJ, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".h")] + [HEADER, os.path.abspath(__file__)]
    jobs = []
    objs = []
    for f in sources():
        src = os.path.join(CSRC, f)
        obj = os.path.join(OBJ, f.rsplit(".", 1)[0] + ".o")
        objs.append(obj)
        if force or _stale(obj, [src] + headers):
            flags = [x for x in NVCC_FLAGS if not (f in FMAD_OK and x == "--fmad=false")]
            cmd = [nvcc, *ARCH, *flags, "-x", "cu", "-c", src, "-o", obj]
            jobs.append((f, cmd))

    def run(job):
        f, cmd = job
        p = subprocess.run(cmd, capture_output=True, text=True)
        return f, p.returncode, p.stdout + p.stderr

    if jobs:
        with cf.ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            for f, rc, out in ex.map(run, jobs):
                if verbose or rc != 0:
                    sys.stderr.write(f"--- {f}\n{out}\n")
                if rc != 0:
                    raise RuntimeError(f"nvcc failed on {f}")
                with open(os.path.join(OBJ, f + ".ptxas.log"), "w") as fh:
                    fh.write(out)
    if force or jobs or _stale(SO, objs):
        cmd = [nvcc, *ARCH, "-shared", "-cudart", "static", "-o", SO, *objs, "-ldl"]
        p = subprocess.run(cmd, capture_output=True, text=True)
        if p.returncode != 0:
            sys.stderr.write(p.stdout + p.stderr)
            raise RuntimeError("link failed")
    return SO


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
