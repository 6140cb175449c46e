#!/bin/bash
# Round 2, GPU call: new host-loop test; A/B of the streaming-point-loads variant (throughput + DRAM bytes by ncu)
set -u
O=gpurun_out/r02d
mkdir -p $O
( python -m pytest tests/test_gpu_parity.py -m gpu -q -k "host_loop or aliasing or multi_gpu" ) > $O/pytest_hostloop.txt 2>&1
tail -3 $O/pytest_hostloop.txt
for v in base ldcs; do
  if [ $v = base ]; then export ASSCPP_LIB=$PWD/asscpp_b200/lib/libasscpp_b200.so; else export ASSCPP_LIB=$PWD/asscpp_b200/lib/variants/lib_$v.so; fi
  python bench.py --steps 8 --warmup 3 --no-cpu-baseline > $O/bench_$v.json 2> $O/bench_$v.err
  python -c "import json;d=json.loads(open('$O/bench_$v.json').read().strip().splitlines()[-1]);print('$v',round(d['value']),round(d['e2e']['value']),d['ms_per_step'])"
done
python tools/research/latency.py > $O/latency.txt 2>&1; cat $O/latency.txt
python tools/phase_profile.py > $O/phase.txt 2>&1; tail -5 $O/phase.txt
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --views-per-step 10000"
for v in base ldcs; do
  if [ $v = base ]; then export ASSCPP_LIB=$PWD/asscpp_b200/lib/libasscpp_b200.so; else export ASSCPP_LIB=$PWD/asscpp_b200/lib/variants/lib_$v.so; fi
  $CMD > $O/ncu_plain_$v.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,smsp__inst_executed.sum --clock-control none -k regex:eval_views_kernel -s 6 -c 1 --csv --log-file $O/ncu_$v.csv $CMD > $O/ncu_$v.log 2>&1
  grep -v "^==" $O/ncu_$v.csv | cut -d, -f13-15 | tail -6
done
