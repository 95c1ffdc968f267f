#!/bin/bash
# development aid: whole GPU suite + headline bench
O=gpurun_out/r7; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q -rA > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest_gpu.log | cut -c1-200
grep -E "^(FAILED|ERROR)" $O/pytest_gpu.log | head -20; grep -E "AssertionError" $O/pytest_gpu.log | cut -c1-400 | head
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 5 > $O/bench_native_1gpu.json 2> $O/bench_native_1gpu.err; echo "bench rc=$?"; cut -c1-250 $O/bench_native_1gpu.json
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 $O/smoke.log
