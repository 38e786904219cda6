# usage: gpu_r2_scale.sh N : config 4 (65,536 distinct frames, one pass) and the default line at N GPUs, torchrun as the driver does
N=$1
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2960$N bench.py --gpus $N --config 4 > gpurun_out/r2_config4_n$N.json 2> gpurun_out/r2_config4_n$N.err || tail -5 gpurun_out/r2_config4_n$N.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2961$N bench.py --gpus $N --steps 100 --warmup 3 > gpurun_out/r2_bench_n$N.json 2> gpurun_out/r2_bench_n$N.err || tail -5 gpurun_out/r2_bench_n$N.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2962$N bench.py --gpus $N --impl reference --steps 3 --warmup 1 > gpurun_out/r2_ref_n$N.json 2> gpurun_out/r2_ref_n$N.err || tail -5 gpurun_out/r2_ref_n$N.err
python - <<PY
import json
for f in ("config4","bench"):
    d=json.load(open(f"gpurun_out/r2_%s_n$N.json"%f)); print(f, d["n_gpus"], round(d["value"]), round(d["e2e"]["value"]), d["e2e"].get("link_frac"), d["steps"], d["scaling"])
d=json.load(open("gpurun_out/r2_ref_n$N.json")); print("ref", round(d["value"],1), d["cpu_baseline"]["cores"], d["cpu_baseline"]["kind"], d["cpu_baseline"]["cpu_model"])
PY
