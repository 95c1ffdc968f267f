#!/bin/bash
# tests + the new bench arms.  Outputs -> gpurun_out/
set -u
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
timeout 1200 python -m pytest tests -m gpu -x -q -rA > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --gpus 1 --steps 200 --warmup 20 > gpurun_out/bench_1gpu.json 2> gpurun_out/bench_1gpu.err; echo "bench1 rc=$?"; cut -c1-400 gpurun_out/bench_1gpu.json
timeout 300 python bench.py --impl restated-graph --gpus 1 --steps 100 --warmup 10 > gpurun_out/bench_restated_graph_1gpu.json 2> gpurun_out/bench_restated_graph_1gpu.err; echo "restated-graph rc=$?"; cut -c1-300 gpurun_out/bench_restated_graph_1gpu.json; tail -3 gpurun_out/bench_restated_graph_1gpu.err
timeout 300 python bench.py --impl restated --gpus 1 --steps 50 --warmup 10 > gpurun_out/bench_restated_1gpu.json 2> gpurun_out/bench_restated_1gpu.err; echo "restated rc=$?"; cut -c1-300 gpurun_out/bench_restated_1gpu.json
timeout 600 python bench.py --mode pbt --gpus $N --rounds 4 --steps-per-round 100 > gpurun_out/bench_pbt_cifar_${N}gpu.json 2> gpurun_out/bench_pbt_cifar_${N}gpu.err; echo "pbt cifar rc=$?"; cat gpurun_out/bench_pbt_cifar_${N}gpu.json; tail -5 gpurun_out/bench_pbt_cifar_${N}gpu.err
if [ "$N" -ge 2 ]; then
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_${N}gpu_s20.json 2> gpurun_out/bench_${N}gpu_s20.err; echo "bench$N rc=$?"; cut -c1-300 gpurun_out/bench_${N}gpu_s20.json
  timeout 600 python bench.py --mode pbt --gpus $N --dp-size 2 --rounds 3 --steps-per-round 60 > gpurun_out/bench_pbt_cifar_dp2_${N}gpu.json 2> gpurun_out/bench_pbt_cifar_dp2_${N}gpu.err; echo "pbt dp2 rc=$?"; cat gpurun_out/bench_pbt_cifar_dp2_${N}gpu.json; tail -5 gpurun_out/bench_pbt_cifar_dp2_${N}gpu.err
fi
