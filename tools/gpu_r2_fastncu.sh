ncu --set full --clock-control none --import-source on --kernel-name "regex:fast_cells" --launch-skip 3 --launch-count 1 -o gpurun_out/r2_fast -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2_fast.log 2>&1
tail -1 gpurun_out/r2_fast.log | cut -c 1-200
