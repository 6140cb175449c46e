#!/bin/bash
set -u
O=gpurun_out/${RUN_TAG:-r02ab}
mkdir -p $O
( time timeout 1200 python -m pytest tests/test_boundary.py -m gpu -q -x ) > $O/pytest_boundary.txt 2>&1
tail -6 $O/pytest_boundary.txt
timeout 600 python bench.py --steps 8 --warmup 3 --no-cpu-baseline > $O/bench_c3.json 2> $O/bench_c3.err
python -c "import json;d=json.loads(open('$O/bench_c3.json').read().strip().splitlines()[-1]);print('c3',round(d['value']),round(d['e2e']['value']))"
