#!/bin/bash
# Developer aid: throughput of the benchmark workload against the walk threshold / setup iterations (not product).
for t in 8 12 16 20 24; do
  echo "== ASSCPP_WALK_THRESH=$t"; ASSCPP_WALK_THRESH=$t python tools/phase_profile.py 2>&1 | sed -n 3p | cut -c1-120
done
for h in 0 1 2 3 5; do
  echo "== ASSCPP_HALO_ITERS=$h"; ASSCPP_HALO_ITERS=$h python tools/phase_profile.py 2>&1 | sed -n 3p | cut -c1-120
done
