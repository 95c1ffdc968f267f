#!/bin/bash
# in-kernel timelines of the halo kernels (libraries built with DTF_TIMING=1 DTF_BUILD_TAG=timing*)
mkdir -p gpurun_out
for tag in ${TAGS:-timing}; do
  export DTF_LIB=$PWD/distributedtf_b200/lib/libdtf_b200_$tag.so
  for cfg in "32 16" "16 32" "8 64"; do
    echo "=== $tag fwd $cfg"; python bench/timeline_conv.py $cfg 2>&1 | tail -45
    echo "=== $tag wgrad $cfg"; python bench/timeline_wgrad.py $cfg 2>&1 | tail -40
  done
done > gpurun_out/timelines.txt 2>&1
unset DTF_LIB
