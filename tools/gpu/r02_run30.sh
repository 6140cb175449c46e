#!/bin/bash
set -u
O=gpurun_out/${RUN_TAG:-r02ag}
mkdir -p $O
( timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "view_cells or carry" ) > $O/pytest_cells.txt 2>&1
tail -12 $O/pytest_cells.txt
