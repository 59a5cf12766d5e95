#!/bin/bash
# quick GPU iteration: parity tests + per-kernel timing of the bench frame
TAG=${1:-q}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 | tee $OUT/pytest_gpu.log
timeout 300 python tools/profile_frame.py 8 2>&1 | tee $OUT/profile_frame.log
