#!/bin/bash
# Round 2, 2-GPU call: the final library under torchrun (NCCL legs + bench of config 3)
set -u
O=gpurun_out/${RUN_TAG:-r02aq}
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29611 tools/nccl_check.py > $O/nccl_check.txt 2>&1; echo "rc=$?" >> $O/nccl_check.txt; tail -n 2 $O/nccl_check.txt
timeout 900 $TR --master-port 29612 bench.py --gpus 2 --steps 10 --warmup 3 > $O/bench_c3_n2.json 2> $O/bench_c3_n2.err
python -c "import json;d=json.loads(open('$O/bench_c3_n2.json').read().strip().splitlines()[-1]);print('c3 N=2',round(d['value']),round(d['e2e']['value']),d.get('collective_ms'))"
