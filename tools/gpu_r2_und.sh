timeout -s KILL 600 python -m pytest tests/test_ref_undistort.py tests/test_undistort.py tests/test_gpu_mih.py -x -q -m gpu 2>&1 | tail -4
