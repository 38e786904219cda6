# round 2, GPU call G: side benches of the widened rows + ncu summaries of the section-8f kernels
python tools/bench_sim3_mih.py > gpurun_out/r2g_sim3_mih.json 2> gpurun_out/r2g_sim3_mih.err || tail -5 gpurun_out/r2g_sim3_mih.err
cat gpurun_out/r2g_sim3_mih.json | cut -c 1-1500
python tools/bench_proj_search.py > gpurun_out/r2g_proj.json 2>gpurun_out/r2g_proj.err || tail -3 gpurun_out/r2g_proj.err
python tools/bench_init_search.py > gpurun_out/r2g_init.json 2>gpurun_out/r2g_init.err || tail -3 gpurun_out/r2g_init.err
python tools/bench_latency.py > gpurun_out/r2g_latency.json 2>gpurun_out/r2g_latency.err || tail -3 gpurun_out/r2g_latency.err
cat gpurun_out/r2g_latency.json | cut -c 1-600
# full-set captures, a handful of launches each (kernel-name filters keep replays short)
ncu --set full --clock-control none --kernel-name "regex:proj_|grid_build" --launch-skip 6 --launch-count 6 -o gpurun_out/r2g_ncu_proj -f python tools/bench_proj_search.py > gpurun_out/r2g_ncu_proj.log 2>&1
ncu --set full --clock-control none --kernel-name "regex:init_|area_query" --launch-skip 4 --launch-count 4 -o gpurun_out/r2g_ncu_init -f python tools/bench_init_search.py > gpurun_out/r2g_ncu_init.log 2>&1
ncu --set full --clock-control none --kernel-name "regex:bow_" --launch-skip 2 --launch-count 4 -o gpurun_out/r2g_ncu_bow -f python tools/bench_bow_search.py > gpurun_out/r2g_ncu_bow.log 2>&1
ncu --set full --clock-control none --kernel-name "regex:match_" --launch-count 6 -o gpurun_out/r2g_ncu_match -f python tools/bench_match.py --max 16384 > gpurun_out/r2g_ncu_match.log 2>&1
ncu --set full --clock-control none --kernel-name "regex:preprocess|remap|undistort" --launch-count 4 -o gpurun_out/r2g_ncu_pre -f python tools/bench_preprocess.py > gpurun_out/r2g_ncu_pre.log 2>&1
ncu --set full --clock-control none --kernel-name "regex:mih_|sim3_" --launch-count 12 -o gpurun_out/r2g_ncu_mih -f python tools/bench_sim3_mih.py > gpurun_out/r2g_ncu_mih.log 2>&1
ls -la gpurun_out/*.ncu-rep
