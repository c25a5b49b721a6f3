set -u
OUT=gpurun_out
mkdir -p $OUT
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tests/multi_gpu_check.py > $OUT/r2_multi_gpu_check.log 2>&1; echo "multi rc=$?"; tail -3 $OUT/r2_multi_gpu_check.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 5 --warmup 3 > $OUT/r2_bench_n2.json 2> $OUT/r2_bench_n2.err; echo "bench2 rc=$?"; tail -3 $OUT/r2_bench_n2.err
