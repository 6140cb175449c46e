#!/bin/bash
# Round 2, GPU call: per-view coverage cells (ASSCPP_EVAL_VIEW_CELLS) and InfoGainVolumetricH in its intended form
set -u
O=gpurun_out/${RUN_TAG:-r02y}
mkdir -p $O
( timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_boundary.py tests/test_search.py -m gpu -q -x -k "view_cells or info_gain or eval_groups or small_batch" ) > $O/pytest_infogain.txt 2>&1
tail -15 $O/pytest_infogain.txt
