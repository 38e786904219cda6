timeout -s KILL 300 python tools/bench_match.py > gpurun_out/r2tc6_sweep.json 2> gpurun_out/r2tc6_sweep.err || tail -5 gpurun_out/r2tc6_sweep.err
cut -c 1-300 gpurun_out/r2tc6_sweep.json
python bench.py --config 3 --steps 20 > gpurun_out/r2_bench_config3.json 2> gpurun_out/r2_bench_config3.err || tail -3 gpurun_out/r2_bench_config3.err
cut -c 1-400 gpurun_out/r2_bench_config3.json
ncu --set full --clock-control none --import-source on --kernel-name regex:match_tc --launch-skip 1 --launch-count 1 -o gpurun_out/r2tc9_full -f python tools/run_match_once.py 65536 2 > gpurun_out/r2tc9_ncu.log 2>&1; tail -1 gpurun_out/r2tc9_ncu.log
