"""Build libeoc_sm100.so (the C-ABI CUDA library) in-tree with nvcc for sm_100a.

    python -m openeo_processes_dask_b200.build [--force] [--verbose]

Objects go to ``build/eoc/`` (git-ignored), the library to
``openeo_processes_dask_b200/_native/libeoc_sm100.so`` (git-ignored, shipped to the GPU box by
gpurun).  nvcc cross-compiles without a GPU.
"""

from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
OBJ_DIR = os.path.join(ROOT, "build", "eoc")
LIB_DIR = os.path.join(PKG, "_native")
LIB_PATH = os.path.join(LIB_DIR, "libeoc_sm100.so")

SOURCES = ["select.cu", "select_fused.cu", "select_fused2.cu", "select_tile.cu", "select_tile_fused.cu", "select_tile_masked.cu", "select_split.cu", "select_split_fused.cu", "select_split_masked.cu", "select_quad.cu", "select_quad_fused.cu", "select_quad_masked.cu", "select_quad_int.cu", "select_quad64.cu", "select_multi64.cu", "select_multi.cu", "select_split_wide64.cu", "select_split_wide128.cu", "select_wide.cu", "select_general.cu", "select_int_u8.cu", "select_int_i16.cu", "select_int_u16.cu", "tile.cu", "reduce.cu", "rows_static.cu", "masked.cu", "scan.cu", "order.cu", "zonal.cu", "interp.cu", "elementwise.cu", "tape.cu", "binary.cu", "api.cu", "synth.cu", "host_pipeline.cu"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
    "-Xptxas", "-v",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; cannot build libeoc_sm100.so")


def _newest_source_mtime() -> float:
    newest = 0.0
    for base, _, files in os.walk(CSRC):
        for f in files:
            newest = max(newest, os.path.getmtime(os.path.join(base, f)))
    newest = max(newest, os.path.getmtime(os.path.join(ROOT, "include", "eoc.h")))
    return newest


def _newest_header_mtime() -> float:
    newest = os.path.getmtime(os.path.join(ROOT, "include", "eoc.h"))
    for base, _, files in os.walk(CSRC):
        for f in files:
            if not f.endswith(".cu"):
                newest = max(newest, os.path.getmtime(os.path.join(base, f)))
    return newest


def _stale(src: str) -> bool:
    """A translation unit is recompiled when its object is missing or older than the .cu or any header."""
    obj = os.path.join(OBJ_DIR, src.replace(".cu", ".o"))
    if not os.path.exists(obj):
        return True
    return os.path.getmtime(obj) < max(os.path.getmtime(os.path.join(CSRC, src)), _newest_header_mtime())


def needs_build() -> bool:
    if not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < _newest_source_mtime():
        return True
    return any(_stale(s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s)))


def _compile(src: str, verbose: bool) -> str:
    obj = os.path.join(OBJ_DIR, src.replace(".cu", ".o"))
    cmd = [_nvcc(), *NVCC_FLAGS, "-c", os.path.join(CSRC, src), "-o", obj]
    res = subprocess.run(cmd, capture_output=True, text=True)
    log = os.path.join(OBJ_DIR, src + ".ptxas.log")
    with open(log, "w") as f:
        f.write(res.stderr)
    if res.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{res.stderr[-4000:]}")
    if verbose:
        print(res.stderr)
    return obj


def build_native(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return LIB_PATH
    os.makedirs(OBJ_DIR, exist_ok=True)
    os.makedirs(LIB_DIR, exist_ok=True)
    sources = [s for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    todo = [s for s in sources if force or _stale(s)]
    with ThreadPoolExecutor(max_workers=min(8, max(1, len(todo)))) as ex:
        list(ex.map(lambda s: _compile(s, verbose), todo))
    objs = [os.path.join(OBJ_DIR, s.replace(".cu", ".o")) for s in sources]
    cmd = [_nvcc(), "-shared", "-o", LIB_PATH, *objs, "-gencode", "arch=compute_100a,code=sm_100a",
           "-Xcompiler", "-fPIC", "-cudart", "static"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"link failed:\n{res.stderr[-4000:]}")
    return LIB_PATH


if __name__ == "__main__":
    path = build_native(force="--force" in sys.argv, verbose="--verbose" in sys.argv)
    print(path)
