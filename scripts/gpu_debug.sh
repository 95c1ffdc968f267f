#!/bin/bash
O=gpurun_out/r7; mkdir -p $O
timeout 600 python -m pytest tests/test_tf32_gpu.py tests/test_kernels_gpu.py -m gpu -q > $O/pytest_tf32.log 2>&1; echo "rc=$?"; tail -3 $O/pytest_tf32.log | cut -c1-300
timeout 300 python bench.py --gpus 1 --steps 50 --warmup 5 --dtype tf32 > $O/bench_native_tf32_1gpu.json 2> $O/tf32.err; echo "tf32 bench rc=$?"; cut -c1-250 $O/bench_native_tf32_1gpu.json; tail -3 $O/tf32.err
