#!/bin/bash
# One gpurun call: GPU tests -> smoke -> bench -> ncu launch list -> ncu --set full of the top kernels.
# usage: tools/gpu_round.sh TAG [skip-tests]
TAG=${1:-r01x}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > $OUT/smi.txt 2>&1
if [ "$2" != "skip-tests" ]; then
  timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
  timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/rc.txt
fi
timeout 600 python bench.py --steps 30 --warmup 3 > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?" | tee -a $OUT/rc.txt
tail -c 3000 $OUT/bench.json
timeout 300 python tools/profile_frame.py 8 > $OUT/profile_frame.log 2>&1; echo "profile_frame rc=$?" | tee -a $OUT/rc.txt
cat $OUT/profile_frame.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/launches.csv \
  python bench.py --steps 2 --warmup 1 --frames 8 > $OUT/ncu_launches.log 2>&1; echo "ncu launches rc=$?" | tee -a $OUT/rc.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_geom|k_raster' -s 4 -c 4 -o $OUT/full \
  python tools/profile_frame.py 4 > $OUT/ncu_full.log 2>&1; echo "ncu full rc=$?" | tee -a $OUT/rc.txt
ls -la $OUT
