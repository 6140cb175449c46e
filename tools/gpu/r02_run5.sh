#!/bin/bash
# Round 2, GPU call: split-group mode (parity + latency), full GPU suite
set -u
O=gpurun_out/r02e
mkdir -p $O
( python -m pytest tests/test_search.py tests/test_boundary.py -m gpu -q ) > $O/pytest_search.txt 2>&1
tail -4 $O/pytest_search.txt
python tools/research/latency.py > $O/latency_split.txt 2>&1; cat $O/latency_split.txt
ASSCPP_SPLIT_GROUP_MAX=0 python tools/research/latency.py > $O/latency_nosplit.txt 2>&1; cat $O/latency_nosplit.txt
( time python -m pytest tests -m gpu -q -x ) > $O/pytest_gpu_all.txt 2>&1
tail -5 $O/pytest_gpu_all.txt
python bench.py --steps 8 --warmup 3 --no-cpu-baseline > $O/bench_c3.json 2> $O/bench_c3.err
python -c "import json;d=json.loads(open('$O/bench_c3.json').read().strip().splitlines()[-1]);print(round(d['value']),round(d['e2e']['value']))"
