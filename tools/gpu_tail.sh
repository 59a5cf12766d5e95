#!/bin/bash
# A/B of the task-size schedule of k_geom_march (taper factor x shortest task), one process each
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for cfg in "0 4" "1 4" "1 2" "1 6" "1 8" "2 8"; do
  set -- $cfg
  echo -n "taper=$1 min_rows=$2  "; VF_GM_TAPER=$1 VF_GM_MIN_ROWS=$2 python tools/ab_frames.py 7 2>&1 | tail -1
done
