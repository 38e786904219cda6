for v in "" rs10 rs12 rs16; do
  if [ -n "$v" ]; then export GFORB_LIBRARY=$PWD/gf-orb-slam2_b200/variants/libgforb_$v.so; fi
  echo "[$v]"; python bench.py --steps 60 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], {k:v['ms'] for k,v in d['stages'].items()})"
done
