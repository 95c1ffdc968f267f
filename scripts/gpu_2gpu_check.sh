#!/bin/bash
# two GPUs: multi-GPU tests and the 2-GPU headline bench with the own-stream build
O=gpurun_out/check2; mkdir -p $O
timeout 400 python -m pytest tests/test_comm_kernels_gpu.py tests/test_pbt_gpu.py -m gpu -q -x > $O/pytest_2gpu.log 2>&1; echo "2-GPU tests rc=$?"; tail -2 $O/pytest_2gpu.log | cut -c1-200
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus 2 --steps 20 --warmup 5 2> $O/b2.err | grep '^{' > $O/bench_native_2gpu.json; echo "bench2 rc=$?"; cut -c1-220 $O/bench_native_2gpu.json; tail -2 $O/b2.err | cut -c1-200
