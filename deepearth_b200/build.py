"""Build recipe for libde_grid4d.so (nvcc, sm_100a only, in-tree so the .so travels to the GPU box).

    python -m deepearth_b200.build [--force] [--verbose]

No torch headers are involved: the library is a plain C-ABI CUDA shared object
(include/de_grid4d.h) loaded through ctypes by deepearth_b200/_lib.py.
"""
from __future__ import annotations

import argparse
import concurrent.futures
import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
OBJ = os.path.join(PKG, "build")
LIB = os.path.join(PKG, "libde_grid4d.so")

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden",
          "-I", os.path.join(ROOT, "include")]
# per-file extra flags
EXTRA = {
    "geo2xyz.cu": ["-fmad=false"],   # one IEEE rounding per fp64 op, reference order
    "grid4d_field_kernels.cu": ["-fmad=false"],   # same prologue arithmetic, generated in place
}
SOURCES = ["c_abi.cu", "grid4d_kernels.cu", "grid4d_pair_kernels.cu", "grid4d_staged_kernels.cu", "grid4d_smem_kernels.cu", "geo2xyz.cu", "misc.cu", "mlp_head.cu", "mlp_head_bwd2.cu", "st_embed.cu", "st_embed_tc.cu", "grid4d_field_kernels.cu", "grid4d_small1d_kernels.cu", "proj_kernels.cu", "proj_bwd_kernels.cu", "grid4d_sorted_kernels.cu"]


def _nvcc() -> str:
    cand = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(cand):
        raise RuntimeError("nvcc not found: libde_grid4d.so cannot be built")
    return cand


def _stale(out: str, deps) -> bool:
    if not os.path.exists(out):
        return True
    t = os.path.getmtime(out)
    return any(os.path.getmtime(d) > t for d in deps)


def build_library(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OBJ, exist_ok=True)
    nvcc = _nvcc()
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    headers.append(os.path.join(ROOT, "include", "de_grid4d.h"))
    sources = [s for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    jobs = []
    for src in sources:
        obj = os.path.join(OBJ, src.replace(".cu", ".o"))
        if force or _stale(obj, [os.path.join(CSRC, src)] + headers):
            cmd = [nvcc, *ARCH, *COMMON, *EXTRA.get(src, []), "-c", os.path.join(CSRC, src), "-o", obj]
            if verbose:
                cmd.insert(1, "-Xptxas=-v")
            jobs.append(cmd)

    def run(cmd):
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + r.stdout + r.stderr)
        return r.stderr

    with concurrent.futures.ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        for log in ex.map(run, jobs):
            if verbose and log:
                print(log)
    objs = [os.path.join(OBJ, s.replace(".cu", ".o")) for s in sources]
    if force or jobs or _stale(LIB, objs):
        run([nvcc, *ARCH, "-shared", "-Xcompiler", "-fPIC", "-o", LIB, *objs, "-cudart", "static"])
    return LIB


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    a = ap.parse_args()
    print(build_library(a.force, a.verbose))
