timeout -s KILL 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "second_tier or overflow" 2>&1 | tail -4
