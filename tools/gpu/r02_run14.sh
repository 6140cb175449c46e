#!/bin/bash
# Round 2, GPU call: neighbourhood walk inside the service pass -- parity, throughput, phase split; 2 CTAs per SM A/B
set -u
O=gpurun_out/${RUN_TAG:-r02p}
mkdir -p $O
( timeout 900 python -m pytest tests/test_search.py tests/test_gpu_parity.py -m gpu -q -x ) > $O/pytest_parity.txt 2>&1
tail -4 $O/pytest_parity.txt
timeout 600 python bench.py --steps 8 --warmup 3 --no-cpu-baseline > $O/bench_c3.json 2> $O/bench_c3.err
python -c "import json;d=json.loads(open('$O/bench_c3.json').read().strip().splitlines()[-1]);print('c3',round(d['value']),round(d['e2e']['value']))"
timeout 600 python tools/phase_profile.py > $O/phase.txt 2>&1; tail -6 $O/phase.txt
ASSCPP_CTA_THREADS=512 timeout 600 python bench.py --steps 8 --warmup 3 --no-cpu-baseline > $O/bench_c3_512.json 2> $O/bench_c3_512.err
python -c "import json;d=json.loads(open('$O/bench_c3_512.json').read().strip().splitlines()[-1]);print('c3 2 CTAs/SM',round(d['value']),round(d['e2e']['value']))"
timeout 300 python tools/research/latency.py > $O/latency.txt 2>&1; cat $O/latency.txt
