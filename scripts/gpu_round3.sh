#!/bin/bash
# one GPU: full GPU test suite (bottleneck engines + wide kernels included), headline bench, bottleneck benches,
# ncu capture of the fused all-reduce + update kernel (world of one).  Outputs -> gpurun_out/r3/
set -u
O=gpurun_out/r3; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q -rA > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 $O/pytest_gpu.log | cut -c1-300
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 5 > $O/bench_native_1gpu.json 2> $O/bench_native_1gpu.err; echo "bench rc=$?"; cut -c1-250 $O/bench_native_1gpu.json
timeout 300 python bench.py --gpus 1 --steps 50 --warmup 5 --bottleneck --resnet-size 29 > $O/bench_bottleneck_v2_29.json 2> $O/bench_bottleneck_v2_29.err; echo "bottleneck rc=$?"; cut -c1-250 $O/bench_bottleneck_v2_29.json; tail -3 $O/bench_bottleneck_v2_29.err
timeout 300 python bench.py --gpus 1 --steps 30 --warmup 5 --bottleneck --resnet-size 29 --impl restated-graph > $O/bench_bottleneck_v2_29_restated_graph.json 2> $O/bench_bottleneck_v2_29_restated_graph.err; echo "bottleneck restated-graph rc=$?"; cut -c1-250 $O/bench_bottleneck_v2_29_restated_graph.json; tail -3 $O/bench_bottleneck_v2_29_restated_graph.err
timeout 300 python bench.py --gpus 1 --steps 50 --warmup 5 --resnet-version 1 > $O/bench_v1_32.json 2> $O/bench_v1_32.err; echo "v1 rc=$?"; cut -c1-250 $O/bench_v1_32.json; tail -3 $O/bench_v1_32.err
COMM="python bench/ncu_comm_targets.py"
timeout 200 $COMM > $O/plain_comm.log 2>&1 && \
timeout 400 ncu --set full --clock-control none --import-source on -k regex:allreduce_update -s 2 -c 2 -o $O/prof_allreduce $COMM > $O/ncu_allreduce.log 2>&1
echo "ncu allreduce rc=$?"; tail -2 $O/plain_comm.log
ncu -i $O/prof_allreduce.ncu-rep --page raw --csv > $O/r2_allreduce_update_raw.csv 2>/dev/null; wc -c $O/r2_allreduce_update_raw.csv
ls -la $O
