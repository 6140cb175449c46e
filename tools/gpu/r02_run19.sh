#!/bin/bash
# Round 2, GPU call: final kernel -- full GPU suite, default bench, ncu launch list and --set full of a 100 000-view launch (traffic.json)
set -u
O=gpurun_out/${RUN_TAG:-r02v}
mkdir -p $O
( time timeout 1500 python -m pytest tests -m gpu -q -x ) > $O/pytest_gpu_all.txt 2>&1
tail -6 $O/pytest_gpu_all.txt
timeout 900 python bench.py > $O/bench_c3_default.json 2> $O/bench_c3_default.err
python -c "import json;d=json.loads(open('$O/bench_c3_default.json').read().strip().splitlines()[-1]);print('c3',round(d['value']),round(d['e2e']['value']),d['gpu_launches'],d['clocks'])"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err
tail -c 600 $O/bench_reference.json
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > $O/plain.log 2>&1 || exit 1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv $CMD > $O/ncu_launches.log 2>&1
timeout 1500 ncu --set full --import-source on --clock-control none -k regex:eval_views_kernel -s 6 -c 1 -o $O/eval100k $CMD > $O/ncu_full.log 2>&1
ls -la $O
