#!/bin/bash
# Round 2, GPU call: ncu launch list and --set full of a 100 000-view launch of the FINAL library (refreshes profiles/traffic.json)
set -u
O=gpurun_out/${RUN_TAG:-r02ap}
mkdir -p $O
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > $O/plain.log 2>&1 || exit 1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv $CMD > $O/ncu_launches.log 2>&1
timeout 1500 ncu --set full --import-source on --clock-control none -k regex:eval_views_kernel -s 6 -c 1 -o $O/eval100k $CMD > $O/ncu_full.log 2>&1
ls -la $O
