#!/bin/bash
# 4 GPUs: whole GPU suite (multi-GPU tests included), data-parallel benches with the in-kernel exchange accounting
# over the device-timed region, single-GPU steps at the per-replica batch for the by-subtraction number.
set -u
O=gpurun_out/r6; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 900 python -m pytest tests -m gpu -q -rA > $O/pytest_gpu_4gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest_gpu_4gpu.log | cut -c1-200; grep -E "^(FAILED|ERROR)" $O/pytest_gpu_4gpu.log | head
timeout 300 $TR --nproc-per-node 4 --master-port 29541 bench.py --gpus 4 --dp-size 2 --steps 200 --warmup 10 --no-e2e 2> $O/dp2.err | grep '^{' > $O/bench_native_4gpu_dp2.json; echo "dp2 rc=$?"; python -c "import json;d=json.load(open('$O/bench_native_4gpu_dp2.json'));print(d['ms_per_step'],d['dp_comm'])"
timeout 300 $TR --nproc-per-node 4 --master-port 29542 bench.py --gpus 4 --dp-size 4 --steps 200 --warmup 10 --no-e2e 2> $O/dp4.err | grep '^{' > $O/bench_native_4gpu_dp4.json; echo "dp4 rc=$?"; python -c "import json;d=json.load(open('$O/bench_native_4gpu_dp4.json'));print(d['ms_per_step'],d['dp_comm'])"
for B in 64 32; do
  timeout 300 python bench.py --gpus 1 --batch $B --steps 200 --warmup 10 --no-e2e 2> $O/b$B.err | grep '^{' > $O/bench_native_1gpu_batch$B.json; python -c "import json;d=json.load(open('$O/bench_native_1gpu_batch$B.json'));print('batch $B', d['ms_per_step'])"
done
tail -3 $O/dp2.err $O/dp4.err | cut -c1-300
