#!/bin/bash
# Round 2, GPU call: A/B of cull claim sizes and chunk sizes (same call, same box), device-timed config 3 + phase split
set -u
O=gpurun_out/${RUN_TAG:-r02q}
mkdir -p $O
for v in base claim128 claim64 chunk64 chunk64c128 chunk256; do
  if [ $v = base ]; then export ASSCPP_LIB=$PWD/asscpp_b200/lib/libasscpp_b200.so; else export ASSCPP_LIB=$PWD/asscpp_b200/lib/variants/lib_$v.so; fi
  timeout 600 python bench.py --steps 6 --warmup 3 --no-cpu-baseline > $O/bench_$v.json 2> $O/bench_$v.err
  python -c "import json;d=json.loads(open('$O/bench_$v.json').read().strip().splitlines()[-1]);print('$v',round(d['value']),round(d['e2e']['value']))"
  timeout 300 python tools/phase_profile.py 2>&1 | head -3 | tee $O/phase_$v.txt
done
