#!/bin/bash
# one compute-sanitizer tool per gpurun call (B200_PROFILING.md): $1 = memcheck | racecheck | synccheck
set -u
TOOL=$1
O=gpurun_out/r02_san
mkdir -p $O
python tools/sanitize_small.py > $O/plain_$TOOL.log 2>&1 || { echo "plain run failed"; tail -5 $O/plain_$TOOL.log; exit 1; }
timeout 1500 compute-sanitizer --tool $TOOL --print-limit 20 python tools/sanitize_small.py > $O/$TOOL.log 2>&1
echo "rc=$?" >> $O/$TOOL.log
tail -15 $O/$TOOL.log
