# tensor-core Hamming: parity first (under a hard timeout: a wrong barrier would spin), then the sweep
timeout -s KILL 240 python -m pytest tests/test_gpu_match.py -x -q 2>&1 | tail -15 > gpurun_out/r2tc_pytest.log; cat gpurun_out/r2tc_pytest.log
grep -q "failed\|error\|Killed" gpurun_out/r2tc_pytest.log && exit 0
grep -q "passed" gpurun_out/r2tc_pytest.log || exit 0
timeout -s KILL 300 python tools/bench_match.py > gpurun_out/r2tc_sweep.json 2> gpurun_out/r2tc_sweep.err || tail -5 gpurun_out/r2tc_sweep.err
cat gpurun_out/r2tc_sweep.json | cut -c 1-1500
GFORB_NO_TC=1 timeout -s KILL 300 python tools/bench_match.py --max 16384 > gpurun_out/r2tc_sweep_popc.json 2>/dev/null; cut -c 1-600 gpurun_out/r2tc_sweep_popc.json
