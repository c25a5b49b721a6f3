python tools/ab_select.py --rows 1024 --only masked_median,median
python tools/ab_select.py --rows 1024 --only masked_median,median --ballast-gb 90
python tools/ab_select.py --rows 1024 --only masked_median,median --slice-of 5490
python bench.py --steps 3 --warmup 3 --no-cpu --no-configs --no-e2e 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print({k:round(v.get('input_gbs',0)) for k,v in d['reducers'].items()})"
