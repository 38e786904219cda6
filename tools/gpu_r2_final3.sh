python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python bench.py > gpurun_out/r2l_bench.json 2> gpurun_out/r2l_bench.err || tail -5 gpurun_out/r2l_bench.err
python -c "
import json;d=json.load(open('gpurun_out/r2l_bench.json'));print(round(d['value']), round(d['e2e']['value']), d['e2e']['link_frac'], d['roofline']['bound'], d['roofline']['frac'], d['roofline']['int_issue']['frac'], d['clocks'], d['cpu_baseline']['value'], d['cpu_baseline']['kind'])"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2l_ref.json; cut -c 1-200 gpurun_out/r2l_ref.json
for c in 1euroc 2tum 2vga 2hd 4; do python bench.py --config $c --steps 30 > gpurun_out/r2l_config$c.json 2> gpurun_out/r2l_config$c.err || tail -3 gpurun_out/r2l_config$c.err; python -c "
import json,sys;d=json.load(open('gpurun_out/r2l_config$c.json'));print('$c', round(d['value']), round(d['e2e']['value']), d['ms_per_step'])"; done
