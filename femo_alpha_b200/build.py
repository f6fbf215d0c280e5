"""Build libfemo_b200.so in-tree with nvcc for sm_100a (B200 only; no other arch is emitted)."""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libfemo_b200.so")
SOURCES = ["api.cu", "topology.cu", "assemble.cu", "assemble_tiles.cu", "solver.cu", "outputs.cu", "dist.cu", "mg.cu", "coarse.cu", "dynamic.cu", "hex.cu", "penalty_shape.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "-Xcompiler", "-fopenmp", "--use_fast_math=false"]


def _nvcc():
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    # every header of the library: a stale object file is exactly what would ship to the GPU box (*.so is git-ignored but pushed)
    headers = sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".hpp", ".h")))
    headers.append(os.path.join(HERE, "..", "include", "femo_b200.h"))
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    nvcc = _nvcc()
    flags = [f for f in NVCC_FLAGS if not f.startswith("--use_fast_math")]

    def compile_one(src):
        obj = os.path.join(objdir, src.replace(".cu", ".o"))
        path = os.path.join(CSRC, src)
        if force or _stale(obj, [path] + headers):
            cmd = [nvcc] + flags + (["-Xptxas", "-v"] if verbose else []) + ["-c", path, "-o", obj]
            r = subprocess.run(cmd, capture_output=True, text=True)
            if r.returncode != 0:
                raise RuntimeError("nvcc failed for %s:\n%s\n%s" % (src, r.stdout, r.stderr))
            if verbose:
                sys.stderr.write(r.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=len(SOURCES)) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    if force or _stale(LIB, objs):
        cmd = [nvcc, "-shared", "-o", LIB] + objs + ["-lcudart", "-ldl", "-lgomp"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n%s\n%s" % (r.stdout, r.stderr))
    build_metis()
    return LIB


METIS_LIB = os.path.join(HERE, "libfemo_metis.so")
METIS_ARCHIVES = ["/usr/local/cuda/targets/x86_64-linux/lib/libmetis_static.a", "/usr/local/cuda/lib64/libmetis_static.a"]


def build_metis(force: bool = False):
    """libfemo_metis.so: csrc/metis_shim.c + CUDA's libmetis_static.a (host only).  Optional: without the archive the
    graph partitioner falls back to RCB (femo_alpha_b200/partition.py)."""
    src = os.path.join(CSRC, "metis_shim.c")
    arch = next((a for a in METIS_ARCHIVES if os.path.exists(a)), None)
    if arch is None:
        return None
    if force or _stale(METIS_LIB, [src]):
        r = subprocess.run(["gcc", "-O2", "-fPIC", "-shared", "-o", METIS_LIB, src, arch, "-lm"], capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("libfemo_metis.so: link failed:\n%s\n%s" % (r.stdout, r.stderr))
    return METIS_LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
