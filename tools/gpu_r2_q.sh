for e in 0 1 0 1; do
  if [ $e = 1 ]; then export GFORB_FORK_EARLY=1; else unset GFORB_FORK_EARLY; fi
  echo "[fork early=$e]"; python bench.py --steps 60 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'])"
done
