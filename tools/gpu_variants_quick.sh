#!/bin/bash
# per-variant ms/frame of the bench workload, no test suite: tools/gpu_variants_quick.sh TAG variant...
TAG=$1; shift
OUT=gpurun_out/$TAG; mkdir -p $OUT
for v in "$@"; do
  echo "== $v" | tee -a $OUT/variants.log
  if [ "$v" = "base" ]; then timeout 300 python tools/ab_frames.py 7 2>&1 | tee -a $OUT/variants.log
  else LD_LIBRARY_PATH=build/variants/$v timeout 300 python tools/ab_frames.py 7 2>&1 | tee -a $OUT/variants.log; fi
done
