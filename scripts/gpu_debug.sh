#!/bin/bash
O=gpurun_out/r7; mkdir -p $O
timeout 600 python -m pytest tests/test_engine_gpu.py -m gpu -q -rA -k "shared_relu" > $O/pytest_masks.log 2>&1; echo "rc=$?"
grep -E "shared-mask|passed|failed|Error" $O/pytest_masks.log | cut -c1-3000 | head
