#!/bin/bash
# Round 2, GPU call: full GPU suite with the small-batch path and lane groups; latency; bench lines of configs 3, 2 and 5
set -u
O=gpurun_out/${RUN_TAG:-r02n}
mkdir -p $O
( time timeout 1500 python -m pytest tests -m gpu -q -x ) > $O/pytest_gpu_all.txt 2>&1
tail -6 $O/pytest_gpu_all.txt
timeout 300 python tools/research/latency.py > $O/latency.txt 2>&1; cat $O/latency.txt
timeout 900 python bench.py > $O/bench_c3_default.json 2> $O/bench_c3_default.err
python -c "import json;d=json.loads(open('$O/bench_c3_default.json').read().strip().splitlines()[-1]);print('c3',round(d['value']),round(d['e2e']['value']),d['roofline'],d['cpu_baseline'])"
timeout 600 python bench.py --config 2 --no-cpu-baseline > $O/bench_c2.json 2> $O/bench_c2.err
python -c "import json;d=json.loads(open('$O/bench_c2.json').read().strip().splitlines()[-1]);print('c2',round(d['value']),round(d['e2e']['value']))"
timeout 1200 python bench.py --config 5 --no-cpu-baseline > $O/bench_c5.json 2> $O/bench_c5.err
python -c "import json;d=json.loads(open('$O/bench_c5.json').read().strip().splitlines()[-1]);print('c5',round(d['value']),round(d['e2e']['value']))"
