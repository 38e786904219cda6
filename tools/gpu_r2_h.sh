# round 2, GPU call H: quadtree with 4-barrier passes; bench with the live copy-only ceiling; sim3/mih side bench
python -m pytest tests/test_gpu_parity.py tests/test_gpu_vs_ref.py tests/test_gpu_synthgen.py -x -q 2>&1 | tail -5 > gpurun_out/r2h_pytest.log; cat gpurun_out/r2h_pytest.log
python bench.py --no-cpu-baseline > gpurun_out/r2h_bench.json 2> gpurun_out/r2h_bench.err || tail -5 gpurun_out/r2h_bench.err
python -c "
import json;d=json.load(open('gpurun_out/r2h_bench.json'));print(round(d['value']), round(d['e2e']['value']), d['e2e'].get('copy_only_ceiling'), d['e2e'].get('link_frac'), {k:v['ms'] for k,v in d['stages'].items()})"
python tools/bench_sim3_mih.py > gpurun_out/r2h_sim3_mih.json 2> gpurun_out/r2h_sim3_mih.err || tail -5 gpurun_out/r2h_sim3_mih.err
cat gpurun_out/r2h_sim3_mih.json | cut -c 1-2500
