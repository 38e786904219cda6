timeout -s KILL 240 python -m pytest tests/test_gpu_match.py -x -q 2>&1 | tail -8 > gpurun_out/r2tc3_pytest.log; cat gpurun_out/r2tc3_pytest.log
grep -q "passed" gpurun_out/r2tc3_pytest.log || exit 0
grep -q "failed" gpurun_out/r2tc3_pytest.log && exit 0
timeout -s KILL 120 python tools/run_match_once.py 65536 4
timeout -s KILL 120 python tools/run_match_once.py 16384 4
