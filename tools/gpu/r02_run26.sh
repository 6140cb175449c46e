#!/bin/bash
# Round 2, GPU call: sweep of the ray-setup iteration knobs with the final kernel (config 3, device-timed)
set -u
O=gpurun_out/${RUN_TAG:-r02ac}
mkdir -p $O
for s in 1 2 3; do for h in 0 3 6; do
  ASSCPP_SETUP_ITERS=$s ASSCPP_HALO_ITERS=$h timeout 600 python bench.py --steps 6 --warmup 3 --no-cpu-baseline > $O/b.json 2> $O/b.err
  python -c "import json;d=json.loads(open('$O/b.json').read().strip().splitlines()[-1]);print('setup_iters $s halo_iters $h',round(d['value']))"
done; done | tee $O/setup_sweep.txt
