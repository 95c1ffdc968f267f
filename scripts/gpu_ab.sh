#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_engine_gpu.py -m gpu -x -q > gpurun_out/pytest_kernels.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_kernels.log
TAGS=timing bash scripts/gpu_timeline.sh
{
echo "##### default split search"
timeout 300 python bench/conv_micro.py 2>&1 | head -8
timeout 300 python bench.py --steps 50 --warmup 10 --no-e2e 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('step ms', d['ms_per_step'], 'img/s', d['value'])"
for pct in 42 50 58 66 72 78; do
  echo "##### DTF_BWD_SPLIT=$pct"
  DTF_BWD_SPLIT=$pct timeout 300 python bench/conv_micro.py 2>&1 | grep "bwd fused"
done
echo "##### DTF_WGRAD_BULK=0 DTF_BWD_SPLIT=60"
DTF_WGRAD_BULK=0 DTF_BWD_SPLIT=60 timeout 300 python bench/conv_micro.py 2>&1 | grep "bwd fused"
} 2>&1 | tee gpurun_out/ab.txt
