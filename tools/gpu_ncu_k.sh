# usage: gpu_ncu_k.sh <kernel regex> <tag> [launch-count]: full-set capture of the matching kernels of one bench step
ncu --set full --clock-control none --import-source on --kernel-name "regex:$1" --launch-skip ${4:-24} --launch-count ${3:-8} -o gpurun_out/$2 -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/$2.log 2>&1
tail -2 gpurun_out/$2.log
