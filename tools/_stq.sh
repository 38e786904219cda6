python -m pytest tests/test_gpu_parity.py tests/test_gpu_vs_ref.py -x -q 2>&1 | tail -3 > gpurun_out/stq_pytest.log; cat gpurun_out/stq_pytest.log
for v in "" st128 st128s16 st128s4; do
  if [ -n "$v" ]; then export GFORB_LIBRARY=$PWD/gf-orb-slam2_b200/variants/libgforb_$v.so; fi
  python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/stq_$v.json 2> gpurun_out/stq_$v.err || tail -5 gpurun_out/stq_$v.err
  python -c "
import json,sys;d=json.load(open('gpurun_out/stq_$v.json'));print('$v', round(d['value']), round(d['e2e']['value']), d['stages']['stereo']['ms'])"
done
unset GFORB_LIBRARY
ncu --set full --clock-control none --import-source on --kernel-name stereo_kernel --launch-skip 3 --launch-count 1 -o gpurun_out/r2_stereo_v11 -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_st.log 2>&1
