#!/bin/bash
# One gpurun call: what the driver runs at round end (GPU tests, smoke, both bench arms), no profiler.
# usage: tools/gpu_verify.sh TAG
TAG=${1:-verify}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > $OUT/smi.txt 2>&1
nproc > $OUT/nproc.txt
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
tail -3 $OUT/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/rc.txt
timeout 600 python bench.py --impl reference > $OUT/bench_ref.json 2> $OUT/bench_ref.err; echo "bench ref rc=$?" | tee -a $OUT/rc.txt
timeout 600 python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?" | tee -a $OUT/rc.txt
cat $OUT/bench_ref.json $OUT/bench.json
timeout 300 python tools/profile_frame.py 8 > $OUT/profile_frame.log 2>&1; echo "profile_frame rc=$?" | tee -a $OUT/rc.txt
cat $OUT/profile_frame.log
