# round 2, GPU call B: new bench (distinct frames per step, configs, reference arm = oracle/_ref)
python -m pytest tests/test_gpu_synthgen.py tests/test_gpu_parity.py -x -q 2>&1 | tail -5 > gpurun_out/r2b_pytest.log; cat gpurun_out/r2b_pytest.log
for c in 1 1euroc 2tum 2vga 2hd 4; do
  python bench.py --config $c > gpurun_out/r2b_bench_$c.json 2> gpurun_out/r2b_bench_$c.err || tail -5 gpurun_out/r2b_bench_$c.err
  tail -1 gpurun_out/r2b_bench_$c.err
  python -c "
import json;d=json.load(open('gpurun_out/r2b_bench_$c.json'));print('$c', round(d['value']), round(d['e2e']['value']), d['e2e'].get('link_frac'), d['steps'], d['clocks'], {k:v['ms'] for k,v in d['stages'].items()}, d.get('cpu_baseline',{}).get('value'))"
done
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2b_ref.json 2> gpurun_out/r2b_ref.err; cat gpurun_out/r2b_ref.json | cut -c 1-400
