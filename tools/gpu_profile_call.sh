#!/bin/bash
# One gpurun call: plain runs first (must exit 0), then the ncu launch list of bench.py and one
# --set full capture per dominant kernel.  Usage: gpurun --timeout 1500 -- 'bash tools/gpu_profile_call.sh r1'
set -u
TAG=${1:-r1}
OUT=gpurun_out
mkdir -p $OUT
BENCH="python bench.py --steps 3 --warmup 3 --no-cpu"
$BENCH > $OUT/plain_bench_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file $OUT/launches_$TAG.csv $BENCH > $OUT/ncu_bench_$TAG.log 2>&1
echo "launch list rc=$?"
for r in ${REDUCERS:-median sd quantiles mean}; do
  CMD="python tools/profile_one.py $r --rows 512 --iters 3"
  $CMD > $OUT/plain_${r}_$TAG.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:'k_(select|columns|tile|scan|interp)' -s 1 -c 1 \
      -f -o $OUT/prof_${r}_$TAG $CMD > $OUT/ncu_${r}_$TAG.log 2>&1
  echo "$r full capture rc=$?"
done
ls -la $OUT
