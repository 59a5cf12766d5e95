#!/bin/bash
# Round-2 opening call: GPU suite + bench on the round-1 kernels (with the advisor's fixes), then ncu --set full
# of the depth variants (c3 + depth, c5) that round 1 never captured.
TAG=${1:-r02a}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > $OUT/smi.txt 2>&1
nproc > $OUT/nproc.txt
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
tail -3 $OUT/pytest_gpu.log
timeout 600 python bench.py --steps 30 --warmup 3 > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?" | tee -a $OUT/rc.txt
tail -c 1500 $OUT/bench.json
for cfg in c3 c3d c5; do
  timeout 600 python tools/profile_frame.py 4 $cfg > $OUT/profile_$cfg.log 2>&1; echo "profile $cfg rc=$?" | tee -a $OUT/rc.txt
  cat $OUT/profile_$cfg.log
done
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_geom|k_raster' -s 4 -c 2 -o $OUT/full_c3d \
  python tools/profile_frame.py 4 c3d > $OUT/ncu_c3d.log 2>&1; echo "ncu c3d rc=$?" | tee -a $OUT/rc.txt
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'k_geom|k_raster' -s 4 -c 2 -o $OUT/full_c5 \
  python tools/profile_frame.py 4 c5 > $OUT/ncu_c5.log 2>&1; echo "ncu c5 rc=$?" | tee -a $OUT/rc.txt
ls -la $OUT
