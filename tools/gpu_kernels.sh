#!/bin/bash
# per-kernel time / instruction / issue metrics of the bench frame (ncu, light metric set)
TAG=${1:-k}
mkdir -p gpurun_out/$TAG
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active --clock-control none -c 60 --csv --log-file gpurun_out/$TAG/l.csv python tools/profile_frame.py 4 > /dev/null 2>&1; echo "ncu rc=$?"
