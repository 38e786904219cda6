python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python bench.py > gpurun_out/r2l_bench.json 2> gpurun_out/r2l_bench.err || tail -5 gpurun_out/r2l_bench.err
python -c "
import json;d=json.load(open('gpurun_out/r2l_bench.json'));print(round(d['value']), round(d['e2e']['value']), d['e2e']['link_frac'], d['roofline']['bound'], d['roofline']['frac'], d['roofline']['int_issue'], d['clocks'], d['cpu_baseline']['value'], d['cpu_baseline']['kind'])"
python bench.py --impl reference --steps 3 --warmup 1 | cut -c 1-300
