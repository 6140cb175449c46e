#!/bin/bash
# Round 2, GPU call: the allocation-free entry point -- alone in a fresh process, then after every other GPU test
set -u
O=gpurun_out/r02f
mkdir -p $O
( python -m pytest tests/test_zz_no_alloc.py -m gpu -q ) > $O/pytest_noalloc_alone.txt 2>&1
tail -4 $O/pytest_noalloc_alone.txt
( time python -m pytest tests -m gpu -q -x ) > $O/pytest_gpu_all.txt 2>&1
tail -5 $O/pytest_gpu_all.txt
