#!/bin/bash
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
for cfg in "1 4" "0 4" "1 4"; do
  set -- $cfg
  echo -n "taper=$1 min_rows=$2  "; VF_GM_TAPER=$1 VF_GM_MIN_ROWS=$2 python tools/ab_frames.py 7 2>&1 | tail -1
done
python tools/ab_c5.py 16384
