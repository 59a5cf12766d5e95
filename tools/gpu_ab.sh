#!/bin/bash
# Quick A/B iteration: GPU parity suite + ms/frame of the bench workload (c3) + per-kernel split.
# usage: tools/gpu_ab.sh TAG [fuzz-seconds]
TAG=${1:-ab}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
tail -4 $OUT/pytest_gpu.log
timeout 300 python tools/ab_frames.py 9 2>&1 | tee $OUT/ab_frames.log
timeout 300 python tools/profile_frame.py 8 2>&1 | tee $OUT/profile_frame.log
if [ -n "$2" ]; then
  timeout $(( $2 + 120 )) python tools/fuzz_parity.py $2 ${3:-201} 2>&1 | tail -8 | tee $OUT/fuzz.log
fi
if [ -n "$4" ]; then
  timeout $(( $4 + 120 )) python tools/fuzz_parity.py api $4 ${3:-201} 2>&1 | tail -4 | tee $OUT/fuzz_api.log
  VF_TEST_SMALL_BUFFERS=1 timeout $(( $4 + 120 )) python tools/fuzz_parity.py api $4 ${3:-201} 2>&1 | tail -4 | tee -a $OUT/fuzz_api.log
fi
