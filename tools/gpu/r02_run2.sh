#!/bin/bash
# Round 2, GPU call: the boundary tests (reference planners over the plugin), the shim, the new ABI entry points
set -u
O=gpurun_out/r02b
mkdir -p $O
( time python -m pytest tests/test_boundary.py tests/test_cpp_shim.py -m gpu -q ) > $O/pytest_boundary_all.txt 2>&1
( time python -m pytest tests/test_gpu_parity.py -m gpu -q -k "aliasing or multi_gpu or eval_ex or uncovered" ) > $O/pytest_new.txt 2>&1
tail -5 $O/pytest_boundary_all.txt $O/pytest_new.txt
