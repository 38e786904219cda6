# round 2, GPU call E: SearchBySim3 + resize ping-pong
python -m pytest tests/test_gpu_sim3.py tests/test_gpu_parity.py tests/test_gpu_projsearch.py -x -q 2>&1 | tail -15 > gpurun_out/r2e_pytest.log; cat gpurun_out/r2e_pytest.log
python bench.py --no-cpu-baseline > gpurun_out/r2e_bench.json 2> gpurun_out/r2e_bench.err || tail -5 gpurun_out/r2e_bench.err
python -c "
import json;d=json.load(open('gpurun_out/r2e_bench.json'));print(round(d['value']), round(d['e2e']['value']), d['e2e'].get('link_frac'), {k:v['ms'] for k,v in d['stages'].items()})"
