"""Rank of the world_size-2 gloo test: the N>1 host logic (viewpoint sharding, covered-bitset
OR-allreduce as allgather + OR, acceptance-count sum) with the CPU oracle as the evaluator."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from asscpp_b200 import adaptive as AD  # noqa: E402
from backends import OracleBackend  # noqa: E402
from oracle import oracle as O  # noqa: E402


def main(out_path):
    dist.init_process_group("gloo")
    world, rank = dist.get_world_size(), dist.get_rank()
    xyz = np.load(os.path.join(ROOT, "tests", "golden", "etihad_model.npz"))["xyz"]
    orc = O.Oracle(xyz)
    model = AD.VoxelModel(xyz, 0.5)
    be = OracleBackend(orc, O, nthreads=2)

    def allreduce_or(bits):
        # same shape as the NCCL path: NCCL has no bitwise OR, so gather every rank's bitset and OR locally
        t = torch.from_numpy(bits.view(np.int64).copy())
        parts = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(parts, t)
        acc = parts[0].numpy().view(np.uint64).copy()
        for p in parts[1:]:
            acc |= p.numpy().view(np.uint64)
        return acc

    def gather_counts(n):
        t = torch.tensor([n], dtype=torch.int64)
        dist.all_reduce(t)
        return int(t.item())

    be.allreduce = allreduce_or
    trace = AD.adaptive_loop(be, model, (0, 0, 4.5), (36, 50, 9), world=world, rank=rank, gather_counts=gather_counts,
                             max_iters=2)
    # sharding of a plain batch + allgather of per-view scores in rank order
    poses = np.load(os.path.join(ROOT, "tests", "golden", "etihad_searchspace_2p5m.npz"))["cam0"][600:641]
    lo, hi = AD.shard_range(len(poses), world, rank)
    st = orc.eval_batch(poses[lo:hi], memo=True)
    sizes = [AD.shard_range(len(poses), world, r) for r in range(world)]
    pad = max(h - l for l, h in sizes)
    mine = np.zeros((pad, st.shape[1]), np.int64)
    mine[: hi - lo] = st
    parts = [torch.empty(pad, st.shape[1], dtype=torch.int64) for _ in range(world)]
    dist.all_gather(parts, torch.from_numpy(mine))
    gathered = np.concatenate([p.numpy()[: h - l] for p, (l, h) in zip(parts, sizes)])
    if rank == 0:
        np.savez(out_path, bits=np.stack([t["bits"] for t in trace]), accepted=[t["accepted"] for t in trace],
                 candidates=[t["candidates"] for t in trace], gathered=gathered)
    dist.destroy_process_group()


if __name__ == "__main__":
    main(sys.argv[1])
