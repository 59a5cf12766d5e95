#!/bin/bash
# One gpurun call: GPU tests -> smoke -> both bench arms -> ncu launch list of bench.py -> ncu --set full of the dominant
# kernels at c3, c3 + depth and c5, and k_raster once more without ncu's cache flush.   usage: tools/gpu_round2.sh TAG
TAG=${1:-r02x}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > $OUT/smi.txt 2>&1; nproc > $OUT/nproc.txt
timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
tail -2 $OUT/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/rc.txt
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > $OUT/bench_reference.json 2> $OUT/bench_reference.err; echo "bench ref rc=$?" | tee -a $OUT/rc.txt
timeout 600 python bench.py --steps 30 --warmup 3 > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?" | tee -a $OUT/rc.txt
for cfg in c3 c3d c3s c4 c5; do timeout 600 python tools/profile_frame.py 8 $cfg > $OUT/profile_$cfg.log 2>&1; head -2 $OUT/profile_$cfg.log; done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/launches.csv \
  python bench.py --steps 2 --warmup 1 --frames 8 > $OUT/ncu_launches.log 2>&1; echo "ncu launches rc=$?" | tee -a $OUT/rc.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_geom|k_raster|k_resolve|k_shade' -s 6 -c 6 -o $OUT/full \
  python tools/profile_frame.py 4 > $OUT/ncu_full.log 2>&1; echo "ncu full rc=$?" | tee -a $OUT/rc.txt
timeout 900 ncu --set full --clock-control none --cache-control none --import-source on -k regex:'k_raster|k_resolve|k_shade' -s 4 -c 4 -o $OUT/raster_nocacheflush \
  python tools/profile_frame.py 4 > $OUT/ncu_raster.log 2>&1; echo "ncu raster (no cache flush) rc=$?" | tee -a $OUT/rc.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_geom|k_raster|k_resolve|k_shade' -s 6 -c 3 -o $OUT/full_c3d \
  python tools/profile_frame.py 4 c3d > $OUT/ncu_c3d.log 2>&1; echo "ncu c3d rc=$?" | tee -a $OUT/rc.txt
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'k_geom|k_raster|k_resolve|k_shade' -s 6 -c 3 -o $OUT/full_c5 \
  python tools/profile_frame.py 4 c5 > $OUT/ncu_c5.log 2>&1; echo "ncu c5 rc=$?" | tee -a $OUT/rc.txt
ls -la $OUT | head -40
