#!/bin/bash
# Round 2, GPU call: allocation-free entry point alone and after the others (only the last two files, the full suite ran in r02f);
# phase profile of small batches
set -u
O=gpurun_out/r02g
mkdir -p $O
( python -m pytest tests/test_zz_no_alloc.py -m gpu -q ) > $O/pytest_noalloc_alone.txt 2>&1
tail -4 $O/pytest_noalloc_alone.txt
( python -m pytest tests/test_search.py tests/test_zz_no_alloc.py -m gpu -q -x ) > $O/pytest_noalloc_after_others.txt 2>&1
tail -4 $O/pytest_noalloc_after_others.txt
python tools/research/latency_phases.py > $O/latency_phases.txt 2>&1; cat $O/latency_phases.txt
