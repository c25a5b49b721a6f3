#!/bin/bash
# quick GPU loop: gpu tests + per-reducer timings on a C2-shaped resident strip
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
for r in ${REDUCERS:-median mean sd max quantiles}; do timeout 120 python tools/profile_one.py $r --rows 512 --iters 5; done
