python tools/run_match_once.py 65536 3
ncu --set full --clock-control none --import-source on --kernel-name regex:match_tc --launch-skip 1 --launch-count 1 -o gpurun_out/r2tc_full -f python tools/run_match_once.py 65536 2 > gpurun_out/r2tc_ncu.log 2>&1; tail -2 gpurun_out/r2tc_ncu.log
