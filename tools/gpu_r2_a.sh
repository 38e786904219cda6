# round 2, GPU call A: parity tests, bench line, ncu launch list and one full-set capture of one step's kernels
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/r2a_pytest.log; cat gpurun_out/r2a_pytest.log
python bench.py --steps 20 --warmup 3 > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err || tail -5 gpurun_out/r2a_bench.err
cat gpurun_out/r2a_bench.json | cut -c 1-600
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2a_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2a_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on --launch-skip 42 --launch-count 14 -o gpurun_out/r2a_full -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2a_ncu2.log 2>&1
ls -la gpurun_out/
