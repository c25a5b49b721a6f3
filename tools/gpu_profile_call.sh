#!/bin/bash
# One gpurun call: plain runs first (must exit 0), then the ncu launch list of bench.py, one --set full capture per
# dominant kernel, and the compute-sanitizer passes.  Usage: gpurun --timeout 2400 -- 'bash tools/gpu_profile_call.sh r2'
set -u
TAG=${1:-r2}
OUT=gpurun_out
mkdir -p $OUT
BENCH="python bench.py --steps 3 --warmup 3 --no-cpu --windows 8"
$BENCH > $OUT/plain_bench_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv \
    --log-file $OUT/launches_$TAG.csv $BENCH > $OUT/ncu_bench_$TAG.log 2>&1
echo "launch list rc=$?"
cap() {  # name, kernel regex, command...
  local name=$1 rx=$2; shift 2
  "$@" > $OUT/plain_${name}_$TAG.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:"$rx" -s 1 -c 1 \
      -f -o $OUT/prof_${name}_$TAG "$@" > $OUT/ncu_${name}_$TAG.log 2>&1
  echo "$name full capture rc=$?"
  # gpurun brings back at most 64 MiB: keep the raw page (and the source page of the selection kernels) as CSV
  if [ -f $OUT/prof_${name}_$TAG.ncu-rep ]; then
    ncu -i $OUT/prof_${name}_$TAG.ncu-rep --page raw --csv > $OUT/raw_${name}_$TAG.csv 2>/dev/null
    case $name in median|quantiles|quant_sd|masked_median|median200|median365)
      ncu -i $OUT/prof_${name}_$TAG.ncu-rep --page source --csv > $OUT/source_${name}_$TAG.csv 2>/dev/null;; esac
    rm -f $OUT/prof_${name}_$TAG.ncu-rep
  fi
}
for r in ${REDUCERS:-median quantiles quant_sd masked_median sd mean ndvi_max}; do
  cap $r 'k_(select|columns|tile)' python tools/profile_one.py $r --rows 512 --iters 3
done
cap median200 'k_select' python tools/profile_one.py median --rows 256 --iters 3 --t 200
cap median365 'k_select' python tools/profile_one.py median --rows 128 --iters 3 --t 365
cap interp "k_interp" python tools/profile_one.py interp --rows 512 --iters 3
cap median_u16 "k_select" python tools/profile_dtype.py uint16 --rows 512
# compute-sanitizer is CLOSED on this pool (its runs left GPUs needing a reset); the small-case driver still runs and
# checks every hand-synchronised kernel against numpy.  SANITIZE=1 re-enables the sanitizer passes elsewhere.
python tools/sanitize_small.py > $OUT/plain_sanitize_$TAG.log 2>&1; echo "sanitize_small rc=$?"
if [ "${SANITIZE:-0}" = "1" ]; then
  python tools/sanitize_small.py > $OUT/plain_sanitize_$TAG.log 2>&1 &&
  compute-sanitizer --tool memcheck python tools/sanitize_small.py > $OUT/sanitizer_memcheck_$TAG.log 2>&1
  echo "memcheck rc=$?"
  timeout 900 compute-sanitizer --tool racecheck python tools/sanitize_small.py > $OUT/sanitizer_racecheck_$TAG.log 2>&1
  echo "racecheck rc=$?"
  tail -5 $OUT/plain_sanitize_$TAG.log; tail -12 $OUT/sanitizer_memcheck_$TAG.log; tail -12 $OUT/sanitizer_racecheck_$TAG.log
fi
ls -la $OUT | head -60
