python bench.py > gpurun_out/r2l_bench.json 2> gpurun_out/r2l_bench.err || tail -5 gpurun_out/r2l_bench.err
python -c "
import json;d=json.load(open('gpurun_out/r2l_bench.json'));print(round(d['value']), round(d['e2e']['value']), d['e2e']['link_frac'], d['roofline']['bound'], d['roofline']['frac'], d['roofline']['int_issue']['frac'], d['roofline']['traffic'], d['clocks'], d['cpu_baseline']['value'], d['cpu_baseline']['kind'])"
python bench.py --config 4 --steps 30 > gpurun_out/r2l_config4.json 2> gpurun_out/r2l_config4.err || tail -3 gpurun_out/r2l_config4.err
python bench.py --config 1 --steps 30 > gpurun_out/r2l_config1.json 2> gpurun_out/r2l_config1.err || tail -3 gpurun_out/r2l_config1.err
