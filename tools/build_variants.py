#!/usr/bin/env python3
"""Tuning builds: recompile ONE translation unit with extra -D flags and link it with the objects of the
regular build into openeo_processes_dask_b200/_native/variants/libeoc_<tag>.so (loaded through
EOC_LIBRARY_PATH).

    python tools/build_variants.py select_tile.cu mix0:-DEOC_MIX_NUM=0 mix25:-DEOC_MIX_NUM=2,-DEOC_MIX_DEN=5
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from openeo_processes_dask_b200 import build as B  # noqa: E402


def one(src, spec):
    tag, flags = spec.split(":", 1)
    flags = flags.split(",")
    obj = os.path.join(B.OBJ_DIR, f"{src[:-3]}.{tag}.o")
    subprocess.run([B._nvcc(), *B.NVCC_FLAGS, *flags, "-c", os.path.join(B.CSRC, src), "-o", obj],
                   check=True, capture_output=True)
    objs = [os.path.join(B.OBJ_DIR, s.replace(".cu", ".o")) for s in B.SOURCES if s != src] + [obj]
    out_dir = os.path.join(B.LIB_DIR, "variants")
    os.makedirs(out_dir, exist_ok=True)
    out = os.path.join(out_dir, f"libeoc_{tag}.so")
    subprocess.run([B._nvcc(), "-shared", "-o", out, *objs, "-gencode", "arch=compute_100a,code=sm_100a",
                    "-Xcompiler", "-fPIC", "-cudart", "static"], check=True, capture_output=True)
    return out


if __name__ == "__main__":
    src = sys.argv[1]
    with ThreadPoolExecutor(8) as ex:
        for path in ex.map(lambda s: one(src, s), sys.argv[2:]):
            print(path)
