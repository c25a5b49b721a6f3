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
# long columns (k_select_wide) and the binary elementwise kernel
CMD="python tools/profile_one.py median --rows 128 --iters 3 --t 365"
$CMD > $OUT/plain_median365_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'k_select_wide' -s 1 -c 1 \
    -f -o $OUT/prof_median365_$TAG $CMD > $OUT/ncu_median365_$TAG.log 2>&1
echo "median365 full capture rc=$?"
CMD="python tools/profile_one.py add --rows 512 --iters 3"
$CMD > $OUT/plain_add_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'k_binary' -s 1 -c 1 \
    -f -o $OUT/prof_add_$TAG $CMD > $OUT/ncu_add_$TAG.log 2>&1
echo "add full capture rc=$?"
ls -la $OUT
