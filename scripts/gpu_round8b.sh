#!/bin/bash
# second 8-GPU round: multi-GPU tests, whole PBT runs, library arms with warm NCCL connections
set -u
O=gpurun_out/r8b; mkdir -p $O
N=$(nvidia-smi -L | wc -l); echo "gpus: $N"
timeout 900 python -m pytest tests -m multigpu -q -rA > $O/pytest_multigpu_${N}gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest_multigpu_${N}gpu.log
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --master-port 29533"
timeout 600 python bench.py --mode pbt --model cifar10 --gpus $N --rounds 6 --steps-per-round 100 2> $O/pbt_cifar.err | grep "^{" > $O/pbt_cifar_pop${N}.json; echo "pbt cifar rc=$?"; cut -c1-200 $O/pbt_cifar_pop${N}.json
timeout 600 python bench.py --mode pbt --model mnist --gpus $N --rounds 6 --steps-per-round 100 2> $O/pbt_mnist.err | grep "^{" > $O/pbt_mnist_pop${N}.json; echo "pbt mnist rc=$?"; cut -c1-200 $O/pbt_mnist_pop${N}.json
timeout 600 python bench.py --mode pbt --model cifar10 --gpus $N --dp-size 2 --rounds 6 --steps-per-round 100 2> $O/pbt_dp2.err | grep "^{" > $O/pbt_cifar_pop$((N/2))_dp2.json; echo "pbt dp2 rc=$?"; cut -c1-200 $O/pbt_cifar_pop$((N/2))_dp2.json
timeout 600 $TR --nproc-per-node $N bench.py --impl restated --gpus $N --steps 30 --warmup 5 2> $O/restated.err | grep "^{" > $O/bench_restated_${N}gpu.json; echo "restated rc=$?"; cut -c1-200 $O/bench_restated_${N}gpu.json
timeout 600 $TR --nproc-per-node $N bench.py --impl restated-graph --gpus $N --steps 50 --warmup 5 2> $O/restated_graph.err | grep "^{" > $O/bench_restated_graph_${N}gpu.json; echo "restated-graph rc=$?"; cut -c1-200 $O/bench_restated_graph_${N}gpu.json
timeout 600 $TR --nproc-per-node $N bench.py --gpus $N --dp-size 2 --steps 100 --warmup 10 2> /dev/null | grep "^{" > $O/bench_native_${N}gpu_dp2.json; cut -c1-200 $O/bench_native_${N}gpu_dp2.json
