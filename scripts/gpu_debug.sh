#!/bin/bash
# development aid: the tf32 tests with their failure messages, then the tf32 bench arm
O=gpurun_out/r5; mkdir -p $O
timeout 600 python -m pytest tests/test_tf32_gpu.py -m gpu -q -rA > $O/pytest_tf32.log 2>&1; echo "tf32 rc=$?"
grep -E "^(FAILED|ERROR)|passed|failed" $O/pytest_tf32.log | cut -c1-160 | head -50
grep -E "AssertionError|Error:" $O/pytest_tf32.log | cut -c1-700 | head -40
timeout 300 python bench.py --gpus 1 --steps 50 --warmup 5 --dtype tf32 > $O/bench_native_tf32_1gpu.json 2> $O/bench_native_tf32_1gpu.err; echo "tf32 bench rc=$?"; cut -c1-250 $O/bench_native_tf32_1gpu.json; tail -3 $O/bench_native_tf32_1gpu.err
timeout 300 python bench.py --gpus 1 --steps 50 --warmup 5 --dtype tf32 --no-graphs --no-e2e > /dev/null 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -s 400 -c 120 --csv --log-file $O/launches_tf32.csv python bench.py --steps 2 --warmup 3 --no-graphs --no-e2e --dtype tf32 > $O/ncu_tf32.log 2>&1; echo "ncu rc=$?"
python bench/summarize_launches.py $O/launches_tf32.csv > $O/launches_tf32_summary.md 2>&1; head -20 $O/launches_tf32_summary.md
