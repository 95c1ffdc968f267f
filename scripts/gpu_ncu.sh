#!/bin/bash
# ncu evidence (one GPU).  Every command runs once without ncu first; captures land in gpurun_out/ncu/.
set -u
O=gpurun_out/ncu; mkdir -p $O
STEP="python bench.py --steps 2 --warmup 3 --no-graphs --no-e2e"
COMM="python bench/ncu_comm_targets.py"
$STEP > $O/plain_step.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 330 -c 200 --csv --log-file $O/launches_step.csv $STEP > $O/ncu_launches.log 2>&1
echo "launch list rc=$?"
$STEP > $O/plain_step2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:conv3x3_bwd_fused -s 60 -c 3 -o $O/prof_bwd_fused $STEP > $O/ncu_bwd.log 2>&1
echo "bwd fused rc=$?"
$STEP > $O/plain_step3.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:conv3x3_kernel -s 90 -c 3 -o $O/prof_conv3x3 $STEP > $O/ncu_fwd.log 2>&1
echo "conv3x3 rc=$?"
$COMM > $O/plain_comm.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"exploit_copy|allreduce_update" -s 4 -c 4 -o $O/prof_comm $COMM > $O/ncu_comm.log 2>&1
echo "comm rc=$?"
ls -la $O
