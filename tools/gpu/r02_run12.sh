#!/bin/bash
# Round 2, GPU call: tiny-batch host planes + completion-flag wait; latency with and without the spin
set -u
O=gpurun_out/${RUN_TAG:-r02m}
mkdir -p $O
( timeout 900 python -m pytest tests/test_search.py tests/test_gpu_parity.py -m gpu -q -x ) > $O/pytest_parity.txt 2>&1
tail -4 $O/pytest_parity.txt
timeout 300 python tools/research/latency.py > $O/latency.txt 2>&1; cat $O/latency.txt
ASSCPP_SMALL_SPIN_US=0 timeout 300 python tools/research/latency.py > $O/latency_nospin.txt 2>&1; cat $O/latency_nospin.txt
timeout 300 python tools/research/latency_phases.py > $O/latency_phases.txt 2>&1; cat $O/latency_phases.txt
