#!/bin/bash
# Round 2, GPU call: ncu --set full with source correlation of one 10 000-view launch of the current kernel; walk-threshold sweep
set -u
O=gpurun_out/${RUN_TAG:-r02r}
mkdir -p $O
for t in 8 12 16 20 24; do
  ASSCPP_WALK_THRESH=$t timeout 600 python bench.py --steps 6 --warmup 3 --no-cpu-baseline > $O/bench_t$t.json 2> $O/bench_t$t.err
  python -c "import json;d=json.loads(open('$O/bench_t$t.json').read().strip().splitlines()[-1]);print('walk_thresh $t',round(d['value']))"
done | tee $O/thresh_sweep.txt
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --views-per-step 10000"
$CMD > $O/plain.log 2>&1 && timeout 1200 ncu --set full --import-source on --clock-control none -k regex:eval_views_kernel -s 6 -c 1 -o $O/eval10k $CMD > $O/ncu.log 2>&1
ls -la $O
