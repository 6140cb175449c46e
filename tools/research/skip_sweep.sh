#!/bin/bash
# sweep of the skip parameters (research)
for cfg in "1 3" "2 2" "2 1" "3 1" "3 2" "4 1" "5 1"; do
  set -- $cfg
  echo "== min_rad $1 max_iter $2"
  ASSCPP_SKIP_MIN_RAD=$1 ASSCPP_SKIP_MAX_ITER=$2 timeout 300 python tools/skip_check.py 10000000 10000 2>&1 | head -3
done
