#!/bin/bash
# One GPU-box round: tests, smoke, 1-GPU bench and (with >= 2 GPUs) the torchrun benches.  Outputs -> gpurun_out/.
set -u
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
echo "gpus: $N"
timeout 900 python -m pytest tests -m gpu -x -q -rA > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench_1gpu_s20.json 2> gpurun_out/bench_1gpu_s20.err; echo "bench1 rc=$?"; cat gpurun_out/bench_1gpu_s20.json
timeout 300 python bench.py --gpus 1 --steps 200 --warmup 20 > gpurun_out/bench_1gpu.json 2> gpurun_out/bench_1gpu.err; echo "bench1b rc=$?"; cat gpurun_out/bench_1gpu.json
for G in 2 4 8; do
  if [ "$N" -ge "$G" ]; then
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $G --steps 20 --warmup 5 > gpurun_out/bench_${G}gpu_s20.json 2> gpurun_out/bench_${G}gpu_s20.err; echo "bench$G rc=$?"; cat gpurun_out/bench_${G}gpu_s20.json; tail -3 gpurun_out/bench_${G}gpu_s20.err
  fi
done
