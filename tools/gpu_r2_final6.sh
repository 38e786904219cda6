python bench.py > gpurun_out/r2l_bench.json 2> gpurun_out/r2l_bench.err || tail -5 gpurun_out/r2l_bench.err
python -c "
import json;d=json.load(open('gpurun_out/r2l_bench.json'));print(round(d['value']), round(d['e2e']['value']), d['e2e']['link_frac'], d['roofline']['bound'], d['roofline']['frac'], d['roofline']['int_issue']['frac'], d['roofline']['traffic'], d['clocks'], d['cpu_baseline']['value'])"
for c in 1 1euroc 2tum 2vga 2hd 4; do python bench.py --config $c --steps 30 > gpurun_out/r2l_config$c.json 2> gpurun_out/r2l_config$c.err || tail -3 gpurun_out/r2l_config$c.err; python -c "
import json,sys;d=json.load(open('gpurun_out/r2l_config$c.json'));print('$c', round(d['value']), round(d['e2e']['value']), d['ms_per_step'], d['cpu_baseline']['value'])"; done
python tools/bench_latency.py > gpurun_out/r2l_latency.json 2> gpurun_out/r2l_latency.err || tail -3 gpurun_out/r2l_latency.err; cut -c 1-400 gpurun_out/r2l_latency.json
