for v in fw2 fw8; do
  export GFORB_LIBRARY=$PWD/gf-orb-slam2_b200/variants/libgforb_$v.so
  echo "[$v]"; python bench.py --steps 60 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], {k:v['ms'] for k,v in d['stages'].items()})"
done
