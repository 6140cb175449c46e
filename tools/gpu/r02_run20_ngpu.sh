#!/bin/bash
# Round 2, N-GPU call (N = $1): NCCL legs, single-process multi-GPU tests, bench of config 3 and config 5 at N
set -u
N=${1:-2}
O=gpurun_out/${RUN_TAG:-r02w}_n$N
mkdir -p $O
nvidia-smi -L > $O/smi.txt
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29611 tools/nccl_check.py > $O/nccl_check.txt 2>&1; echo "rc=$?" >> $O/nccl_check.txt
( timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "multi_gpu or nccl or host_loop" ) > $O/pytest_multi.txt 2>&1; echo "rc=$?" >> $O/pytest_multi.txt
timeout 900 $TR --master-port 29612 bench.py --gpus $N --steps 10 --warmup 3 > $O/bench_c3.json 2> $O/bench_c3.err
timeout 1500 $TR --master-port 29613 bench.py --gpus $N --config 5 --no-cpu-baseline > $O/bench_c5.json 2> $O/bench_c5.err
tail -n 3 $O/nccl_check.txt; tail -n 3 $O/pytest_multi.txt
for c in c3 c5; do python -c "import json;d=json.loads(open('$O/bench_$c.json').read().strip().splitlines()[-1]);print('$c N=$N',round(d['value']),round(d['e2e']['value']),d.get('collective_ms'))"; done
