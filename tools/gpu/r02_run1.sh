#!/bin/bash
# Round 2, GPU call 1: GPU test suite, the three bench configs, the reference arm, ncu launch list + one full capture.
set -u
O=gpurun_out/r02a
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.txt 2>&1
( time python -m pytest tests -m gpu -x -q ) > $O/pytest_gpu.txt 2>&1
echo "pytest rc=$?" >> $O/pytest_gpu.txt
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.txt 2>&1
python bench.py --steps 20 --warmup 5 > $O/bench_c3_n1.json 2> $O/bench_c3_n1.err
python bench.py --impl reference --steps 20 --warmup 5 > $O/bench_c3_reference.json 2> $O/bench_c3_reference.err
python bench.py --config 2 --warmup 5 > $O/bench_c2_n1.json 2> $O/bench_c2_n1.err
python bench.py --config 3 --brute --steps 3 --warmup 3 --views-per-step 10000 --no-cpu-baseline > $O/bench_c3_brute.json 2> $O/bench_c3_brute.err
python bench.py --config 5 --steps 2 --warmup 3 --cpu-seconds 8 > $O/bench_c5_n1.json 2> $O/bench_c5_n1.err
python tools/research/latency.py > $O/latency.txt 2>&1
# ncu: launch list, then one full capture of the batch launch (same command, exited 0 just before)
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > $O/ncu_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/ncu_launches.csv $CMD > $O/ncu_launches.log 2>&1
$CMD > $O/ncu_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:eval_views_kernel -s 6 -c 1 -o $O/prof_eval_views $CMD > $O/ncu_full.log 2>&1
ls -la $O
