set -u
OUT=gpurun_out
mkdir -p $OUT
python -m pytest tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -15 > $OUT/r2d_zz_alone.log
python -m pytest tests -m gpu -x -q 2>&1 | tail -25 > $OUT/r2d_pytest.log
for cfg in "EOC_SELECT_SPLIT=0" "EOC_SELECT_SPLIT=1" "EOC_SELECT_SPLIT=1 EOC_SPLIT_W=256" "EOC_X=default"; do
  env $cfg python tools/ab_select.py --rows 1024 >> $OUT/r2d_ab.jsonl 2>> $OUT/r2d_ab.err
done
python bench.py --steps 5 --warmup 3 --cpu-seconds 6 > $OUT/r2d_bench.json 2> $OUT/r2d_bench.err
echo "bench rc=$?"
tail -3 $OUT/r2d_zz_alone.log; tail -4 $OUT/r2d_pytest.log; cat $OUT/r2d_ab.jsonl; tail -3 $OUT/r2d_bench.err
