#!/usr/bin/env bash
# Presubmit (reference: official/utils/testing/scripts/presubmit.sh — lint + unit tests under both interpreters).
# Here: byte-compile every Python file, cross-compile every CUDA source for sm_100a, run the CPU test tier; the
# GPU tier runs when a device is visible.
#   scripts/presubmit.sh [lint|build|py_test|gpu_test]...      (default: all that apply)
set -euo pipefail
cd "$(dirname "$0")/.."
steps=("$@"); [ ${#steps[@]} -eq 0 ] && steps=(lint build py_test gpu_test)
for s in "${steps[@]}"; do
  case "$s" in
    lint)     python -m compileall -q distributedtf_b200 tests bench bench.py build_native.py __graft_entry__.py
              if python -c 'import pylint' 2>/dev/null; then python -m pylint --rcfile=scripts/pylint.rcfile distributedtf_b200; fi ;;
    build)    python build_native.py ;;
    py_test)  python -m pytest tests -q -m "not gpu" ;;
    gpu_test) if python -c 'import torch,sys; sys.exit(0 if torch.cuda.is_available() else 1)'; then
                python -m pytest tests -q -m gpu; else echo "no GPU visible: gpu_test skipped"; fi ;;
    *) echo "unknown step $s" >&2; exit 2 ;;
  esac
done
