timeout -s KILL 240 python -m pytest tests/test_gpu_match.py -x -q 2>&1 | tail -2
timeout -s KILL 300 python tools/bench_match.py > gpurun_out/r2tc6_sweep.json 2> gpurun_out/r2tc6_sweep.err || tail -5 gpurun_out/r2tc6_sweep.err
cut -c 1-1600 gpurun_out/r2tc6_sweep.json
ncu --set full --clock-control none --import-source on --kernel-name regex:match_ --launch-skip 3 --launch-count 3 -o gpurun_out/r2tc6_full -f python tools/run_match_once.py 65536 2 > gpurun_out/r2tc6_ncu.log 2>&1; tail -1 gpurun_out/r2tc6_ncu.log
python bench.py --config 3 --steps 20 > gpurun_out/r2_bench_config3.json 2> gpurun_out/r2_bench_config3.err || tail -3 gpurun_out/r2_bench_config3.err
cut -c 1-700 gpurun_out/r2_bench_config3.json
