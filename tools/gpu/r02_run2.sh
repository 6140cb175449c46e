#!/bin/bash
# Round 2, GPU call: the boundary tests (reference planners over the plugin), the shim, the new ABI entry points
set -u
O=gpurun_out/r02b
mkdir -p $O
( time python -m pytest tests/test_boundary.py tests/test_cpp_shim.py tests/test_ref_pin.py -m gpu -q -x ) > $O/pytest_boundary.txt 2>&1
echo "rc=$?" >> $O/pytest_boundary.txt
( time python -m pytest tests/test_boundary.py tests/test_cpp_shim.py -m gpu -q ) > $O/pytest_boundary_all.txt 2>&1
tail -30 $O/pytest_boundary_all.txt
