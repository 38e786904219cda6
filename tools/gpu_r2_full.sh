# the round-end sequence the driver runs: full GPU suite, smoke(), the default bench line
timeout -s KILL 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
timeout -s KILL 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout -s KILL 600 python bench.py > gpurun_out/r2_bench_default.json 2> gpurun_out/r2_bench_default.err || tail -5 gpurun_out/r2_bench_default.err
cut -c 1-900 gpurun_out/r2_bench_default.json
