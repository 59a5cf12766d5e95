#!/bin/bash
# usage: tools/gpu_variants.sh TAG variant1 variant2 ...   ("base" = the in-tree library): GPU suite once on the base, then
# per variant the bench frame's per-kernel times
TAG=$1; shift
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
tail -3 $OUT/pytest_gpu.log
for v in "$@"; do
  echo "== $v" | tee -a $OUT/variants.log
  if [ "$v" = "base" ]; then timeout 300 python tools/profile_frame.py 8 2>&1 | tee -a $OUT/variants.log; timeout 300 python tools/ab_frames.py 7 2>&1 | tee -a $OUT/variants.log
  else LD_LIBRARY_PATH=build/variants/$v timeout 300 python tools/profile_frame.py 8 2>&1 | tee -a $OUT/variants.log; LD_LIBRARY_PATH=build/variants/$v timeout 300 python tools/ab_frames.py 7 2>&1 | tee -a $OUT/variants.log; fi
done
timeout 200 python tools/fuzz_parity.py ${FUZZ_S:-60} ${FUZZ_SEED:-221} 2>&1 | tail -3 | tee $OUT/fuzz.log
