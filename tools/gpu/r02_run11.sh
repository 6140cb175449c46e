#!/bin/bash
# Round 2, GPU call: what stalls a SMALL batch?  ncu --set full of one eval_views_small_kernel launch (40 views)
set -u
O=gpurun_out/${RUN_TAG:-r02l}
mkdir -p $O
timeout 300 python tools/research/latency.py > $O/latency_plain.txt 2>&1 || exit 1
timeout 900 ncu --set full --import-source on --clock-control none -k regex:eval_views_small -s 800 -c 1 -o $O/small40 python tools/research/latency.py > $O/ncu_small.log 2>&1
ls -la $O
