#!/bin/bash
# Round 2, GPU call: shared-memory request sized by the model's largest possible grid -- parity, latency, config 2 and 3
set -u
O=gpurun_out/${RUN_TAG:-r02ad}
mkdir -p $O
( timeout 900 python -m pytest tests/test_search.py tests/test_gpu_parity.py -m gpu -q -x ) > $O/pytest_parity.txt 2>&1
tail -3 $O/pytest_parity.txt
for v in sized full; do
  if [ $v = full ]; then export ASSCPP_FULL_SMEM=1; else unset ASSCPP_FULL_SMEM; fi
  timeout 300 python tools/research/latency.py > $O/latency_$v.txt 2>&1; cat $O/latency_$v.txt
  timeout 600 python bench.py --config 2 --no-cpu-baseline > $O/bench_c2_$v.json 2> $O/bench_c2_$v.err
  python -c "import json;d=json.loads(open('$O/bench_c2_$v.json').read().strip().splitlines()[-1]);print('$v c2',round(d['value']),round(d['e2e']['value']))"
done
unset ASSCPP_FULL_SMEM
timeout 600 python bench.py --steps 8 --warmup 3 --no-cpu-baseline > $O/bench_c3.json 2> $O/bench_c3.err
python -c "import json;d=json.loads(open('$O/bench_c3.json').read().strip().splitlines()[-1]);print('c3',round(d['value']),round(d['e2e']['value']))"
