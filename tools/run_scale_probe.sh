set -x
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
lscpu > gpurun_out/lscpu.txt 2>&1
for n in 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n tools/h2d_ceiling.py --out gpurun_out/h2d_ceiling_multi.json > gpurun_out/h2d_n$n.log 2>&1
  tail -c 600 gpurun_out/h2d_n$n.log
done
for n in 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2952$n bench.py --gpus $n --steps 20 --warmup 3 > gpurun_out/r2_bench0_n$n.json 2> gpurun_out/r2_bench0_n$n.err
  tail -c 300 gpurun_out/r2_bench0_n$n.json
done
