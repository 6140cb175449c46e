"""Write tests/golden/etihad_oracle_golden.npz: the oracle's outputs on the bundled model for a fixed
subset of the fixture viewpoints.  The reference publishes no outputs for this path (PARITY UNPINNED),
so these vectors pin the ORACLE against regressions and give the GPU tests a travelling fixture.

  python tools/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402

G = os.path.join(ROOT, "tests", "golden")
xyz = np.load(os.path.join(G, "etihad_model.npz"))["xyz"]
ss = np.load(os.path.join(G, "etihad_searchspace_2p5m.npz"))
poses = np.concatenate([ss["cam0"], ss["cam1"]])
o = O.Oracle(xyz)
all_st = o.eval_batch(poses, memo=False, nthreads=os.cpu_count())
nz = np.nonzero(all_st[:, O.ST_SURVIVORS] > 0)[0]
pick = np.sort(np.concatenate([nz[:: max(1, len(nz) // 96)], np.arange(0, len(poses), 245)]))
pick = np.unique(pick)
st = o.eval_batch(poses[pick], memo=True, nthreads=1)
list_view = int(np.argmax(st[:, O.ST_VISIBLE]))
r = o.eval_view(poses[pick[list_view]])
np.savez_compressed(os.path.join(G, "etihad_oracle_golden.npz"), view_index=pick, stats=st, list_view=list_view,
                    visible_list=r["visible"], survivor_list=r["survivors"])
print("golden:", len(pick), "views; list view", pick[list_view], "visible", len(r["visible"]))
