#!/bin/bash
set -u
O=gpurun_out/${RUN_TAG:-r02x}
mkdir -p $O
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.txt 2>&1; echo "rc=$?" >> $O/smoke.txt; cat $O/smoke.txt
