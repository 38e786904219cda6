for v in "" tc16; do
  if [ -n "$v" ]; then export GFORB_LIBRARY=$PWD/gf-orb-slam2_b200/variants/libgforb_$v.so; fi
  echo "[$v]"; timeout -s KILL 240 python -m pytest tests/test_gpu_match.py -x -q 2>&1 | tail -2
  timeout -s KILL 120 python tools/run_match_once.py 65536 4 | tail -2
  timeout -s KILL 120 python tools/run_match_once.py 8192 4 | tail -1
done
