#!/bin/bash
# Round 2, GPU call: dynamic warp-per-voxel threshold; A/B of the persisting-L2 window over the ray queues (speed + DRAM bytes)
set -u
O=gpurun_out/${RUN_TAG:-r02k}
mkdir -p $O
( timeout 900 python -m pytest tests/test_search.py tests/test_gpu_parity.py -m gpu -q -x ) > $O/pytest_parity.txt 2>&1
tail -4 $O/pytest_parity.txt
timeout 300 python tools/research/latency.py > $O/latency.txt 2>&1; cat $O/latency.txt
timeout 300 python tools/research/latency_phases.py > $O/latency_phases.txt 2>&1; cat $O/latency_phases.txt
for v in 1 0; do
  export ASSCPP_L2_PERSIST=$v
  timeout 600 python bench.py --steps 8 --warmup 3 --no-cpu-baseline > $O/bench_c3_l2persist$v.json 2> $O/bench_c3_l2persist$v.err
  python -c "import json;d=json.loads(open('$O/bench_c3_l2persist$v.json').read().strip().splitlines()[-1]);print('l2persist $v',round(d['value']),round(d['e2e']['value']))"
done
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --views-per-step 10000"
for v in 1 0; do
  export ASSCPP_L2_PERSIST=$v
  $CMD > $O/ncu_plain_$v.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,smsp__inst_executed.sum --clock-control none -k regex:eval_views_kernel -s 6 -c 1 --csv --log-file $O/ncu_l2persist$v.csv $CMD > $O/ncu_$v.log 2>&1
  grep -v "^==" $O/ncu_l2persist$v.csv | cut -d, -f13-15 | tail -6
done
