#!/bin/bash
set -u
O=gpurun_out/${RUN_TAG:-r02z}
mkdir -p $O
( time timeout 1200 python -m pytest tests/test_boundary.py -m gpu -q -x -k "info_gain" ) > $O/pytest_infogain_astar.txt 2>&1
tail -15 $O/pytest_infogain_astar.txt
