# final evidence (after the describe staging and quadtree scheduling changes)
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2f_plain.json 2> gpurun_out/r2f_plain.err || tail -3 gpurun_out/r2f_plain.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2f_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on --launch-skip 42 --launch-count 14 -o gpurun_out/r2_full -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2f_ncu2.log 2>&1
tail -1 gpurun_out/r2f_ncu2.log | cut -c 1-200
