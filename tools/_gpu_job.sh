set -u
for cfg in "EOC_HOST_NT=0 EOC_HOST_THREADS=8" "EOC_HOST_NT=1 EOC_HOST_THREADS=8" "EOC_HOST_NT=0 EOC_HOST_THREADS=16" "EOC_HOST_NT=1 EOC_HOST_THREADS=16" "EOC_HOST_NT=1 EOC_HOST_THREADS=12" "EOC_HOST_NT=1 EOC_HOST_THREADS=24" "EOC_HOST_NT=1 EOC_HOST_THREADS=16 EOC_STRIP_BYTES=33554432" "EOC_HOST_NT=1 EOC_HOST_THREADS=16 EOC_STRIP_BYTES=134217728"; do
  echo "== $cfg"; env $cfg python tools/pageable_e2e.py 2>&1 | tail -2
done
nproc; python -m pytest tests -m gpu -x -q -k "host_pipeline or dask" 2>&1 | tail -2
