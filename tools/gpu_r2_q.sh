timeout -s KILL 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "roi_and_blur or chunked" 2>&1 | tail -5
