#!/bin/bash
# A/B of environment knobs on the bench workload: tools/gpu_ab2.sh TAG "ENV1=.." "ENV2=.." ...  ("-" = no knob)
TAG=$1; shift
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
tail -4 $OUT/pytest_gpu.log
for knob in "$@"; do
  echo "== $knob" | tee -a $OUT/ab.log
  if [ "$knob" = "-" ]; then timeout 300 python tools/ab_frames.py 9 2>&1 | tee -a $OUT/ab.log
  else env $knob timeout 300 python tools/ab_frames.py 9 2>&1 | tee -a $OUT/ab.log; fi
done
timeout 300 python tools/profile_frame.py 8 2>&1 | tee $OUT/profile_frame.log
timeout 300 python tools/profile_frame.py 4 c3d 2>&1 | tee $OUT/profile_c3d.log
timeout 600 python tools/profile_frame.py 4 c5 2>&1 | tee $OUT/profile_c5.log
timeout 240 python tools/fuzz_parity.py ${FUZZ_S:-90} ${FUZZ_SEED:-211} 2>&1 | tail -4 | tee $OUT/fuzz.log
