ncu --set full --clock-control none --import-source on --kernel-name "regex:stereo_kernel|octree_kernel" --launch-skip 6 --launch-count 3 -o gpurun_out/r2_so -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2_so.log 2>&1
tail -1 gpurun_out/r2_so.log | cut -c 1-200
