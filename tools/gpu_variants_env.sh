#!/bin/bash
# per-variant per-kernel times + ms/frame of the bench workload: tools/gpu_variants_env.sh TAG variant...
#   variant = base | NAME (build/variants/NAME) | env:VAR=VALUE (in-tree library with that environment)
TAG=$1; shift
OUT=gpurun_out/$TAG; mkdir -p $OUT
for v in "$@"; do
  echo "== $v" | tee -a $OUT/variants.log
  case "$v" in
    base) E="";;
    env:*) E="${v#env:}";;
    *) E="LD_LIBRARY_PATH=build/variants/$v";;
  esac
  env $E timeout 300 python tools/profile_frame.py 8 ${CFG:-c3} 2>&1 | tee -a $OUT/variants.log
  env $E timeout 300 python tools/ab_frames.py 7 2>&1 | tee -a $OUT/variants.log
done
