python -m pytest tests/test_gpu_parity.py tests/test_gpu_vs_ref.py -x -q 2>&1 | tail -3
for v in "" d16 d4; do
  if [ -n "$v" ]; then export GFORB_LIBRARY=$PWD/gf-orb-slam2_b200/variants/libgforb_$v.so; fi
  python bench.py --steps 60 --warmup 3 --no-cpu-baseline > gpurun_out/r2k_$v.json 2> gpurun_out/r2k_$v.err || tail -5 gpurun_out/r2k_$v.err
  python -c "
import json;d=json.load(open('gpurun_out/r2k_$v.json'));print('[$v]', round(d['value']), round(d['e2e']['value']), d['stages']['describe']['ms'])"
done
