"""Build libb200spai.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python -m learningsparsepreconditioner4gpu_b200.build [--force]

The shared library is written next to this file (it is git-ignored but travels to the GPU box with
the repo snapshot).  Objects are cached under csrc/_obj and rebuilt when a source or header is newer.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(CSRC, "_obj")
LIB = os.path.join(HERE, "libb200spai.so")
INCLUDE = os.path.join(os.path.dirname(HERE), "include")

SOURCES = ["csr.cu", "pcg.cu", "pcg_fused.cu", "assemble.cu", "gnn.cu", "dist.cu", "noalloc.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
    "-I", INCLUDE,
]
if os.environ.get("B200SPAI_PROBE"):  # development: cycle counters inside the streaming SpMV
    NVCC_FLAGS.append("-DB200SPAI_PROBE")
if os.environ.get("B200SPAI_S2_NO_PREFETCH"):  # development: A/B of the row-pointer prefetch in pcg_small2.cuh
    NVCC_FLAGS.append("-DB200SPAI_S2_NO_PREFETCH")
if os.environ.get("B200SPAI_PART_W"):  # development: row-block partition weight
    NVCC_FLAGS.append("-DB200SPAI_PART_W=" + os.environ["B200SPAI_PART_W"])
LIB = os.environ.get("B200SPAI_LIB_OUT", LIB)


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; libb200spai.so cannot be built")


def _newest_header_mtime() -> float:
    t = 0.0
    for d in (CSRC, INCLUDE):
        for f in os.listdir(d):
            if f.endswith((".h", ".cuh")):
                t = max(t, os.path.getmtime(os.path.join(d, f)))
    return t


def build(force: bool = False, verbose: bool = True) -> str:
    os.makedirs(OBJ, exist_ok=True)
    nvcc = _nvcc()
    hdr_t = _newest_header_mtime()
    sources = [s for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    jobs = []
    objs = []
    for src in sources:
        sp = os.path.join(CSRC, src)
        op = os.path.join(OBJ, src.replace(".cu", ".o"))
        objs.append(op)
        if force or not os.path.exists(op) or os.path.getmtime(op) < max(os.path.getmtime(sp), hdr_t):
            jobs.append([nvcc, *NVCC_FLAGS, "-c", sp, "-o", op])

    def run(cmd):
        if verbose:
            print("[build]", " ".join(cmd[-4:]), flush=True)
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + r.stdout + r.stderr)
        return r

    if jobs:
        with ThreadPoolExecutor(max_workers=min(4, len(jobs))) as ex:
            list(ex.map(run, jobs))
    if jobs or force or not os.path.exists(LIB) or any(os.path.getmtime(o) > os.path.getmtime(LIB) for o in objs):
        run([nvcc, "-shared", "-o", LIB, *objs, "-gencode", "arch=compute_100a,code=sm_100a"])
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
