# round 2, GPU call D: full GPU suite + bench after the resize (64-row tiles), NVTX hooks, new tests
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/r2d_pytest.log; cat gpurun_out/r2d_pytest.log
python bench.py --no-cpu-baseline > gpurun_out/r2d_bench.json 2> gpurun_out/r2d_bench.err || tail -5 gpurun_out/r2d_bench.err
python -c "
import json;d=json.load(open('gpurun_out/r2d_bench.json'));print(round(d['value']), round(d['e2e']['value']), d['e2e'].get('link_frac'), {k:v['ms'] for k,v in d['stages'].items()})"
