#!/bin/bash
set -u
O=gpurun_out/${RUN_TAG:-r02af}
mkdir -p $O
( timeout 900 python -m pytest tests/test_search.py tests/test_gpu_parity.py -m gpu -q -x ) > $O/pytest_parity.txt 2>&1
tail -3 $O/pytest_parity.txt
timeout 300 python tools/research/latency.py > $O/latency.txt 2>&1; cat $O/latency.txt
timeout 300 python tools/research/latency_phases.py > $O/latency_phases.txt 2>&1; cat $O/latency_phases.txt
timeout 600 python bench.py --config 2 --no-cpu-baseline > $O/bench_c2.json 2> $O/bench_c2.err
python -c "import json;d=json.loads(open('$O/bench_c2.json').read().strip().splitlines()[-1]);print('c2',round(d['value']),round(d['e2e']['value']))"
timeout 600 python bench.py --steps 8 --warmup 3 --no-cpu-baseline > $O/bench_c3.json 2> $O/bench_c3.err
python -c "import json;d=json.loads(open('$O/bench_c3.json').read().strip().splitlines()[-1]);print('c3',round(d['value']),round(d['e2e']['value']))"
