#!/bin/bash
# Round 2, GPU call: by-chunk cull for coarse lattices -- parity (automatic and forced on every model), latency, config 2
set -u
O=gpurun_out/${RUN_TAG:-r02ai}
mkdir -p $O
( timeout 900 python -m pytest tests/test_search.py tests/test_gpu_parity.py tests/test_fuzz.py -m gpu -q -x ) > $O/pytest_parity.txt 2>&1
tail -3 $O/pytest_parity.txt
( ASSCPP_CHUNK_CULL=1 timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_fuzz.py -m gpu -q -x -k "not 10m and not 50m and not config2_full" ) > $O/pytest_forced.txt 2>&1
tail -3 $O/pytest_forced.txt
for v in 1 0; do
  export ASSCPP_CHUNK_CULL=$v
  timeout 300 python tools/research/latency.py > $O/latency_chunk$v.txt 2>&1; cat $O/latency_chunk$v.txt
  timeout 600 python bench.py --config 2 --no-cpu-baseline > $O/bench_c2_chunk$v.json 2> $O/bench_c2_chunk$v.err
  python -c "import json;d=json.loads(open('$O/bench_c2_chunk$v.json').read().strip().splitlines()[-1]);print('chunk_cull $v c2',round(d['value']),round(d['e2e']['value']))"
done
unset ASSCPP_CHUNK_CULL
timeout 300 python tools/research/latency_phases.py > $O/latency_phases.txt 2>&1; cat $O/latency_phases.txt
