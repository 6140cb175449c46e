#!/bin/bash
# Round 2, GPU call: the round-end sequence as the driver runs it -- full GPU suite (-x), smoke(), default bench, reference arm
set -u
O=gpurun_out/${RUN_TAG:-r02ae}
mkdir -p $O
( time timeout 1800 python -m pytest tests -x -q -m gpu ) > $O/pytest_gpu_all.txt 2>&1
tail -6 $O/pytest_gpu_all.txt
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.txt 2>&1; echo "rc=$?" >> $O/smoke.txt; tail -2 $O/smoke.txt
timeout 900 python bench.py > $O/bench_default.json 2> $O/bench_default.err
python -c "import json;d=json.loads(open('$O/bench_default.json').read().strip().splitlines()[-1]);print('c3',round(d['value']),round(d['e2e']['value']),d['gpu_launches'],d['clocks'],d['roofline']['issue']['frac'],d['roofline']['dram']['frac'])"
