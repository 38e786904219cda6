python -m pytest tests/test_gpu_parity.py tests/test_gpu_vs_ref.py -x -q 2>&1 | tail -3
python bench.py --steps 60 --warmup 3 --no-cpu-baseline > gpurun_out/r2j.json 2> gpurun_out/r2j.err || tail -5 gpurun_out/r2j.err
python -c "
import json;d=json.load(open('gpurun_out/r2j.json'));print(round(d['value']), round(d['e2e']['value']), {k:v['ms'] for k,v in d['stages'].items()})"
bash tools/gpu_ncu_k.sh pyr_ r2j_pyr 8 24
