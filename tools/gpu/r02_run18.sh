#!/bin/bash
set -u
O=gpurun_out/${RUN_TAG:-r02t}
mkdir -p $O
timeout 600 python tools/phase_profile.py > $O/phase.txt 2>&1; cat $O/phase.txt
