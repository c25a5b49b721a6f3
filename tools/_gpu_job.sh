set -u
OUT=gpurun_out
mkdir -p $OUT
python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > $OUT/r2h_pytest.log
tail -4 $OUT/r2h_pytest.log
for t in 256 400 512; do
  rows=$((51200 / t))
  for cfg in "EOC_SELECT_SPLIT=0" "EOC_X=1"; do
  env $cfg python tools/ab_select.py --rows $rows --t $t --only median,quantiles >> $OUT/r2h_ab_wide.jsonl 2>> $OUT/r2h_ab.err
  done
done
for cfg in "EOC_SELECT_SPLIT=1" "EOC_SELECT_SPLIT=1 EOC_SPLIT_PAR=1"; do
  env $cfg python tools/ab_select.py --rows 1024 --only median,quantiles >> $OUT/r2h_ab_wide.jsonl 2>> $OUT/r2h_ab.err
done
cat $OUT/r2h_ab_wide.jsonl; tail -3 $OUT/r2h_ab.err
