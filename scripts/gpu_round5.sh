#!/bin/bash
# one GPU: tf32 kernels + engine tests first (new code), then the whole GPU suite, then benches.  Outputs -> gpurun_out/r5/
set -u
O=gpurun_out/r5; mkdir -p $O
timeout 600 python -m pytest tests/test_tf32_gpu.py -m gpu -q -rA -x > $O/pytest_tf32.log 2>&1; echo "tf32 rc=$?"; grep -E "passed|failed|Error|error|tf32 resnet" $O/pytest_tf32.log | cut -c1-600 | tail -15
timeout 900 python -m pytest tests -m gpu -q -rA > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest_gpu.log | cut -c1-300; grep -E "^FAILED|^ERROR" $O/pytest_gpu.log | head -20
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 5 > $O/bench_native_1gpu.json 2> $O/bench_native_1gpu.err; echo "bench rc=$?"; cut -c1-250 $O/bench_native_1gpu.json
timeout 300 python bench.py --gpus 1 --steps 30 --warmup 5 --dtype tf32 > $O/bench_native_tf32_1gpu.json 2> $O/bench_native_tf32_1gpu.err; echo "tf32 bench rc=$?"; cut -c1-250 $O/bench_native_tf32_1gpu.json; tail -3 $O/bench_native_tf32_1gpu.err
ls -la $O
