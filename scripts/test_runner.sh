#!/usr/bin/env bash
# The reference's benchmark sweep (test_runner.sh:5-24): PBT wall time for every (n, pop_size) cell, one result
# line "n = N, pop_size = P, time = Ts" per cell appended to test_results.txt.
#   scripts/test_runner.sh [arm] [model]        arm: native | native-mf | restated     model: mnist (sweep protocol)
# Processes instead of mpirun: the in-process fabric for one GPU box, `--transport tcp` for one process per GPU.
set -euo pipefail
ARM="${1:-native}"
cd "$(dirname "$0")/.."
python bench/pbt_sweep.py --arm "$ARM" --n 2 3 4 5 --pop 10 20 30 40 50 | tee -a test_results.txt
