timeout -s KILL 900 python tools/stress_extract.py 80 31337 2>&1 | tail -2
