timeout -s KILL 900 python -m pytest tests/test_gpu_stress.py -m gpu -x -q 2>&1 | tail -3
