#!/bin/bash
# exit codes of the test processes with own streams; fallback configuration if they are not clean
O=gpurun_out/r10; mkdir -p $O
timeout 300 python -X faulthandler -m pytest tests/test_pbt_gpu.py -m gpu -q -x -k "members_in_flight or polling" > $O/own.log 2>&1; echo "own streams subset rc=$?"; tail -25 $O/own.log | cut -c1-160
timeout 600 python -X faulthandler -m pytest tests -m gpu -q -x > $O/pytest_gpu.log 2>&1; echo "own streams suite rc=$?"; tail -2 $O/pytest_gpu.log | cut -c1-200
DTF_POOLED_STREAMS=1 DTF_INFLIGHT_GRAPHS=0 timeout 300 python -X faulthandler -m pytest tests/test_pbt_gpu.py -m gpu -q -x -k "members_in_flight or polling" > $O/pooled.log 2>&1; echo "pooled subset rc=$?"; tail -3 $O/pooled.log | cut -c1-160
