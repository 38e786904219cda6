# ncu evidence of the final round-2 kernels: launch list + full-set capture of one device-resident step (14 launches)
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2f_plain.json 2> gpurun_out/r2f_plain.err || tail -3 gpurun_out/r2f_plain.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2f_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on --launch-skip 42 --launch-count 14 -o gpurun_out/r2_full -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2f_ncu2.log 2>&1
tail -2 gpurun_out/r2f_ncu2.log
