#!/bin/bash
# The 8-GPU round: scaling at 1/2/4/8, whole PBT runs (BASELINE configs 2, 3, 4), library arms, exploit sweep,
# multi-GPU test log.  Outputs -> gpurun_out/r8/
set -u
O=gpurun_out/r8; mkdir -p $O
N=$(nvidia-smi -L | wc -l); echo "gpus: $N"
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv > $O/clocks_before.csv
timeout 1500 python -m pytest tests -m gpu -q -rA > $O/pytest_gpu_${N}gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest_gpu_${N}gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $O/smoke.log
timeout 300 python bench/peer_copy_bw.py > $O/peer_copy_bw.json 2>&1; cat $O/peer_copy_bw.json
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --master-port 29533"
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 5 > $O/bench_native_1gpu.json 2> $O/bench_native_1gpu.err; cut -c1-200 $O/bench_native_1gpu.json
for G in 2 4 8; do
  [ "$N" -ge "$G" ] || continue
  timeout 600 $TR --nproc-per-node $G bench.py --gpus $G --steps 20 --warmup 5 > $O/bench_native_${G}gpu.json 2> $O/bench_native_${G}gpu.err; echo "native $G rc=$?"; cut -c1-200 $O/bench_native_${G}gpu.json
done
timeout 600 $TR --nproc-per-node $N bench.py --gpus $N --steps 200 --warmup 20 > $O/bench_native_${N}gpu_s200.json 2> /dev/null; cut -c1-200 $O/bench_native_${N}gpu_s200.json
timeout 600 $TR --nproc-per-node $N bench.py --gpus $N --dp-size 2 --steps 100 --warmup 10 > $O/bench_native_${N}gpu_dp2.json 2> $O/bench_native_${N}gpu_dp2.err; echo "dp2 rc=$?"; cut -c1-200 $O/bench_native_${N}gpu_dp2.json
# library arms
timeout 600 $TR --nproc-per-node $N bench.py --impl restated --gpus $N --steps 30 --warmup 5 > $O/bench_restated_${N}gpu.json 2> $O/bench_restated_${N}gpu.err; echo "restated rc=$?"; cut -c1-200 $O/bench_restated_${N}gpu.json
timeout 600 $TR --nproc-per-node $N bench.py --impl restated-graph --gpus $N --steps 50 --warmup 5 > $O/bench_restated_graph_${N}gpu.json 2> $O/bench_restated_graph_${N}gpu.err; echo "restated-graph rc=$?"; cut -c1-200 $O/bench_restated_graph_${N}gpu.json
timeout 300 python bench.py --impl restated-graph --gpus 1 --steps 100 --warmup 10 > $O/bench_restated_graph_1gpu.json 2> /dev/null
timeout 300 python bench.py --impl restated --gpus 1 --steps 50 --warmup 10 > $O/bench_restated_1gpu.json 2> /dev/null
timeout 600 $TR --nproc-per-node 4 bench.py --impl restated --dp-size 2 --gpus 4 --steps 30 --warmup 5 > $O/bench_restated_4gpu_dp2_nccl.json 2> $O/bench_restated_4gpu_dp2_nccl.err; echo "restated dp2 rc=$?"; cut -c1-200 $O/bench_restated_4gpu_dp2_nccl.json
# whole PBT runs through launch.run
timeout 900 python bench.py --mode pbt --model cifar10 --gpus $N --rounds 6 --steps-per-round 100 > $O/pbt_cifar_pop${N}.json 2> $O/pbt_cifar_pop${N}.err; echo "pbt cifar rc=$?"; cut -c1-250 $O/pbt_cifar_pop${N}.json
timeout 900 python bench.py --mode pbt --model mnist --gpus $N --rounds 6 --steps-per-round 100 > $O/pbt_mnist_pop${N}.json 2> $O/pbt_mnist_pop${N}.err; echo "pbt mnist rc=$?"; cut -c1-250 $O/pbt_mnist_pop${N}.json
timeout 900 python bench.py --mode pbt --model cifar10 --gpus $N --dp-size 2 --rounds 6 --steps-per-round 100 > $O/pbt_cifar_pop$((N/2))_dp2.json 2> $O/pbt_cifar_pop$((N/2))_dp2.err; echo "pbt dp2 rc=$?"; cut -c1-250 $O/pbt_cifar_pop$((N/2))_dp2.json
# exploit sweep
timeout 900 $TR --nproc-per-node $N bench/exploit_sweep.py > $O/exploit_sweep_${N}gpu.jsonl 2> $O/exploit_sweep.err; echo "sweep rc=$?"; tail -3 $O/exploit_sweep_${N}gpu.jsonl | cut -c1-300
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv > $O/clocks_after.csv
