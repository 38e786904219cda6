# scratch: the last quick A/B run of the round (kept for the pattern: variant libraries through GFORB_LIBRARY + one bench line each)
for v in "" ; do
  python bench.py --steps 60 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], {k:v['ms'] for k,v in d['stages'].items()})"
done
