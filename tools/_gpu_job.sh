set -u
OUT=gpurun_out
mkdir -p $OUT
python -m pytest tests -m gpu -x -q 2>&1 | tail -25 > $OUT/r2f_pytest.log
tail -4 $OUT/r2f_pytest.log
for t in 150 200 300 365; do
  rows=$((51200 / t))
  python tools/ab_select.py --rows $rows --t $t --only median,quantiles >> $OUT/r2f_ab_wide.jsonl 2>> $OUT/r2f_ab.err
done
cat $OUT/r2f_ab_wide.jsonl
