GFORB_POISON_PYR=1 timeout -s KILL 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python bench.py --steps 60 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], {k:v['ms'] for k,v in d['stages'].items()})"
