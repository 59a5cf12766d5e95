#!/bin/bash
# ncu --set full of selected kernels of the bench frame.  usage: gpu_ncu.sh TAG KERNEL_REGEX
TAG=${1:-n}; KRE=${2:-k_geom}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 300 python tools/profile_frame.py 8 2>&1 | tee $OUT/profile_frame.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$KRE" -s 2 -c 2 -o $OUT/full -f \
  python tools/profile_frame.py 4 > $OUT/ncu_full.log 2>&1; echo "ncu full rc=$?"
