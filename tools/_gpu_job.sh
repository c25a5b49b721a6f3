set -u
OUT=gpurun_out
mkdir -p $OUT
python bench.py --steps 10 --warmup 3 > $OUT/r2_bench_n1.json 2> $OUT/r2_bench_n1.err; echo "bench rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > $OUT/r2_bench_ref.json 2> $OUT/r2_bench_ref.err; echo "ref rc=$?"
tail -2 $OUT/r2_bench_n1.err
