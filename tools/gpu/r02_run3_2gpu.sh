#!/bin/bash
# Round 2, 2-GPU call: NCCL legs (gain allgather, count check, OR-allreduce, sharded adaptive loop), the
# single-process multi-GPU evaluator, and the bench at N=2 with both collectives on the timed path.
set -u
O=gpurun_out/r02c
mkdir -p $O
nvidia-smi -L > $O/smi.txt
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 tools/nccl_check.py > $O/nccl_check_n2.txt 2>&1
echo "rc=$?" >> $O/nccl_check_n2.txt
( python -m pytest tests/test_gpu_parity.py -m gpu -q -k "multi_gpu or nccl" ) > $O/pytest_multi.txt 2>&1
echo "rc=$?" >> $O/pytest_multi.txt
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus 2 --steps 10 --warmup 3 > $O/bench_c3_n2.json 2> $O/bench_c3_n2.err
python bench.py --gpus 1 --steps 10 --warmup 3 --no-cpu-baseline > $O/bench_c3_n1.json 2> $O/bench_c3_n1.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29613 bench.py --impl reference --gpus 2 --steps 3 --warmup 3 --cpu-seconds 4 > $O/bench_ref_n2.json 2> $O/bench_ref_n2.err
tail -3 $O/nccl_check_n2.txt $O/pytest_multi.txt; tail -c 600 $O/bench_c3_n2.json; tail -c 300 $O/bench_c3_n2.err
