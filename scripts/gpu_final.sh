#!/bin/bash
# final check of the tree as committed: GPU suite exit code, smoke, headline bench at 200 steps
O=gpurun_out/final; mkdir -p $O
timeout 300 python -m pytest tests -m gpu -q -x > $O/pytest_gpu.log 2>&1; echo "suite rc=$?"; tail -1 $O/pytest_gpu.log | cut -c1-160
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?"
timeout 120 python bench.py --gpus 1 --steps 200 --warmup 20 > $O/bench_native_1gpu_s200.json 2> $O/bench.err; echo "bench rc=$?"; cut -c1-230 $O/bench_native_1gpu_s200.json
