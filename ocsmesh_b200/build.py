"""In-tree build of libocsb200.so (hand-written CUDA for sm_100a + the C ABI).

``python -m ocsmesh_b200.build`` or ``ocsmesh_b200.build.build()``.  nvcc
cross-compiles without a GPU; the resulting ``ocsmesh_b200/libocsb200.so`` is
git-ignored but travels to the GPU box with the working tree.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
INCLUDE = os.path.join(ROOT, "include")
OBJDIR = os.path.join(ROOT, "build", "obj")
LIB = os.path.join(HERE, "libocsb200.so")

SOURCES = [
    "ocsb_api.cu",
    "k_elementwise.cu",
    "k_limiter.cu",
    "k_interp.cu",
    "k_bucket.cu",
    "k_spline.cu",
    "k_feature.cu",
    "k_pass.cu",
    "k_nearest.cu",
    "k_constraint.cu",
    "k_warp.cu",
    "k_p2p.cu",
    "h_tiff.cu",
    "k_contour.cu",
    "k_stitch.cu",
    "k_region.cu",
    "k_sample.cu",
    "h_stitch.cu",
]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    # IEEE division / sqrt, no FMA contraction: results must match numpy/scipy
    "-prec-div=true", "-prec-sqrt=true", "-fmad=false",
    "-Xcompiler", "-fPIC,-O2,-Wall,-ffp-contract=off",
    "-I", INCLUDE, "-I", CSRC,
]


def _nvcc() -> str:
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found: libocsb200 cannot be built")
    return exe


def _digest(paths, extra=()) -> str:
    h = hashlib.sha256()
    h.update(" ".join(NVCC_FLAGS + list(extra)).encode())
    for p in sorted(paths):
        with open(p, "rb") as f:
            h.update(p.encode())
            h.update(f.read())
    return h.hexdigest()


def _headers():
    hs = [os.path.join(INCLUDE, f) for f in os.listdir(INCLUDE) if f.endswith(".h")]
    hs += [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cuh")]
    return hs


def _compile_one(src: str, verbose: bool, extra=(), objdir: str = OBJDIR) -> str:
    path = os.path.join(CSRC, src)
    obj = os.path.join(objdir, src.replace(".cu", ".o"))
    stamp = obj + ".sha"
    dig = _digest([path] + _headers(), extra)
    if os.path.exists(obj) and os.path.exists(stamp):
        with open(stamp) as f:
            if f.read().strip() == dig:
                return obj
    cmd = [_nvcc()] + NVCC_FLAGS + list(extra) + ["-c", path, "-o", obj]
    if verbose:
        print(" ".join(cmd), flush=True)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed on {src}:\n{r.stdout}\n{r.stderr}")
    if verbose and r.stderr.strip():
        print(r.stderr, file=sys.stderr)
    with open(stamp, "w") as f:
        f.write(dig)
    return obj


def build(verbose: bool = False, force: bool = False, extra=(), lib: str = LIB,
          objdir: str = OBJDIR) -> str:
    """Compile every CUDA source for sm_100a and link libocsb200.so.

    ``extra`` (more nvcc flags, e.g. ``-DOCSB_INTERP_MIN_CTAS=5``), ``lib`` and
    ``objdir`` build a tuning variant beside the product library for the
    tools/ sweeps (loaded through ``OCSB_LIB``)."""
    LIB, OBJDIR = lib, objdir  # noqa: N806 (shadow the module defaults)
    os.makedirs(OBJDIR, exist_ok=True)
    if force:
        for f in os.listdir(OBJDIR):
            os.remove(os.path.join(OBJDIR, f))
    srcs = [s for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    missing = set(SOURCES) - set(srcs)
    if missing:
        raise RuntimeError(f"missing CUDA sources: {sorted(missing)}")
    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(lambda s: _compile_one(s, verbose, extra, OBJDIR), srcs))
    newest = max(os.path.getmtime(o) for o in objs)
    if (not os.path.exists(LIB)) or os.path.getmtime(LIB) < newest or force:
        cmd = [_nvcc(), "-shared", "-o", LIB] + objs + [
            "-gencode", "arch=compute_100a,code=sm_100a",
            "-Xcompiler", "-fPIC"]
        if verbose:
            print(" ".join(cmd), flush=True)
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return LIB


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv, force="-f" in sys.argv))
