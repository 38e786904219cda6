# memcheck of the extraction + stereo path on small inputs (single-image handles and the batch path)
timeout -s KILL 900 compute-sanitizer --tool memcheck --error-exitcode 9 --print-limit 5 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "not graph and not chunk" 2>&1 | tail -15
echo "rc=$?"
