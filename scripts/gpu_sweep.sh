#!/bin/bash
# the reference's own sweep protocol (test_runner.sh) with the round-2 build, one GPU.  Outputs -> gpurun_out/r8/
O=gpurun_out/r8; mkdir -p $O; rm -f $O/pbt_sweep_r2.jsonl $O/sweep.log
timeout 300 python -m pytest tests/test_pbt_gpu.py -m gpu -q -k "polling" 2>&1 | tail -3
for arm_ck in "native 1" "native 0" "native-mf 0"; do
  set -- $arm_ck
  timeout 400 python bench/pbt_sweep.py --arm $1 --n 2 3 --pop 10 30 50 --checkpoint-every $2 2> $O/sweep_$1_$2.err | tee -a $O/sweep.log | grep '^{' >> $O/pbt_sweep_r2.jsonl
  echo "$1 ckpt $2 rc=$?"; tail -2 $O/sweep_$1_$2.err | cut -c1-300
done
grep "^n =" $O/sweep.log
