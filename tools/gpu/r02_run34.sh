#!/bin/bash
# Round 2, GPU call: templated cull (coarse form out of line) -- same-box A/B on config 3, parity, coarse-lattice numbers
set -u
O=gpurun_out/${RUN_TAG:-r02al}
mkdir -p $O
for rep in 1 2; do for v in base prechunk; do
  if [ $v = base ]; then export ASSCPP_LIB=$PWD/asscpp_b200/lib/libasscpp_b200.so; else export ASSCPP_LIB=$PWD/asscpp_b200/lib/variants/lib_$v.so; fi
  timeout 600 python bench.py --steps 8 --warmup 3 --no-cpu-baseline > $O/bench_${v}_$rep.json 2> $O/bench_${v}_$rep.err
  python -c "import json;d=json.loads(open('$O/bench_${v}_$rep.json').read().strip().splitlines()[-1]);print('$v rep $rep',round(d['value']),round(d['e2e']['value']))"
done; done | tee $O/ab.txt
unset ASSCPP_LIB
( timeout 900 python -m pytest tests/test_search.py tests/test_gpu_parity.py tests/test_fuzz.py -m gpu -q -x -k "not 10m and not 50m" ) > $O/pytest_parity.txt 2>&1
tail -3 $O/pytest_parity.txt
( ASSCPP_CHUNK_CULL=1 timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_fuzz.py -m gpu -q -x -k "not 10m and not 50m and not config2_full" ) > $O/pytest_forced.txt 2>&1
tail -3 $O/pytest_forced.txt
timeout 300 python tools/research/latency.py > $O/latency.txt 2>&1; cat $O/latency.txt
timeout 600 python bench.py --config 2 --no-cpu-baseline > $O/bench_c2.json 2> $O/bench_c2.err
python -c "import json;d=json.loads(open('$O/bench_c2.json').read().strip().splitlines()[-1]);print('c2',round(d['value']),round(d['e2e']['value']))"
