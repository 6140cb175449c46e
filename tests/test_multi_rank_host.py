"""CPU test of the N>1 path (world_size 2, gloo): sharded evaluation + OR-allreduce + score allgather give
exactly the single-process result."""
import os
import subprocess
import sys

import numpy as np

from conftest import ROOT


def test_shard_range_partitions():
    from asscpp_b200.adaptive import shard_range
    for n in (0, 1, 7, 100, 1225):
        for world in (1, 2, 3, 8):
            parts = [shard_range(n, world, r) for r in range(world)]
            assert parts[0][0] == 0 and parts[-1][1] == n
            assert all(parts[i][1] == parts[i + 1][0] for i in range(world - 1))
            assert max(h - l for l, h in parts) - min(h - l for l, h in parts) <= 1


def test_world2_gloo_matches_single_process(tmp_path, oracle_mod, etihad, space_2p5):
    out = str(tmp_path / "w2.npz")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    subprocess.check_call([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                           "--master-addr", "127.0.0.1", "--master-port", "29531",
                           os.path.join(ROOT, "tests", "gloo_worker.py"), out], env=env, timeout=600)
    w2 = np.load(out)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from asscpp_b200 import adaptive as AD
    from backends import OracleBackend
    xyz, _ = etihad
    orc = oracle_mod.Oracle(xyz)
    trace = AD.adaptive_loop(OracleBackend(orc, oracle_mod, nthreads=4), AD.VoxelModel(xyz, 0.5), (0, 0, 4.5),
                             (36, 50, 9), max_iters=2)
    assert np.array_equal(w2["bits"], np.stack([t["bits"] for t in trace]))
    assert list(w2["accepted"]) == [t["accepted"] for t in trace]
    assert list(w2["candidates"]) == [t["candidates"] for t in trace]
    st = orc.eval_batch(space_2p5["cam0"][600:641], memo=True)
    assert np.array_equal(w2["gathered"], st)
