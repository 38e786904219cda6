# round 2, GPU call F: multi-index hashing + the whole GPU suite
python -m pytest tests/test_gpu_mih.py -x -q 2>&1 | tail -25 > gpurun_out/r2f_mih.log; cat gpurun_out/r2f_mih.log
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/r2f_pytest.log; cat gpurun_out/r2f_pytest.log
