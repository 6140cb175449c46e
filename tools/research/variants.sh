#!/bin/bash
# Developer aid: phase split + throughput of kernel variants built into asscpp_b200/lib/variants/ (not product).
for f in asscpp_b200/lib/variants/lib_*.so; do
  echo "== $f"
  ASSCPP_LIB=$PWD/$f python tools/phase_profile.py 2>&1 | head -3
done
