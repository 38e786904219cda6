# round 2, GPU call C: resize kernel with running row pointers
python -m pytest tests/test_gpu_synthgen.py tests/test_gpu_parity.py tests/test_gpu_vs_ref.py -x -q 2>&1 | tail -5 > gpurun_out/r2c_pytest.log; cat gpurun_out/r2c_pytest.log
python bench.py --no-cpu-baseline > gpurun_out/r2c_bench.json 2> gpurun_out/r2c_bench.err || tail -5 gpurun_out/r2c_bench.err
python -c "
import json;d=json.load(open('gpurun_out/r2c_bench.json'));print(round(d['value']), round(d['e2e']['value']), d['e2e'].get('link_frac'), {k:v['ms'] for k,v in d['stages'].items()})"
