#!/bin/bash
# Round 2, GPU call: guided claim sizes in the cull's P2 -- same-box A/B on config 3
set -u
O=gpurun_out/${RUN_TAG:-r02ao}
mkdir -p $O
for rep in 1 2; do for v in base guided; do
  if [ $v = base ]; then export ASSCPP_LIB=$PWD/asscpp_b200/lib/libasscpp_b200.so; else export ASSCPP_LIB=$PWD/asscpp_b200/lib/variants/lib_$v.so; fi
  timeout 600 python bench.py --steps 8 --warmup 3 --no-cpu-baseline > $O/bench_${v}_$rep.json 2> $O/bench_${v}_$rep.err
  python -c "import json;d=json.loads(open('$O/bench_${v}_$rep.json').read().strip().splitlines()[-1]);print('$v rep $rep',round(d['value']),round(d['e2e']['value']))"
done; done | tee $O/ab.txt
export ASSCPP_LIB=$PWD/asscpp_b200/lib/variants/lib_guided.so
timeout 300 python tools/phase_profile.py 2>&1 | head -4 | tee $O/phase_guided.txt
