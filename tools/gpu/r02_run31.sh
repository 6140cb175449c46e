#!/bin/bash
set -u
O=gpurun_out/${RUN_TAG:-r02ah}
mkdir -p $O
timeout 1500 python bench.py --config 5 > $O/bench_c5_n1.json 2> $O/bench_c5_n1.err
python -c "import json;d=json.loads(open('$O/bench_c5_n1.json').read().strip().splitlines()[-1]);print('c5 N=1',round(d['value']),round(d['e2e']['value']),d['cpu_baseline'])"
