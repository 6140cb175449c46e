"""Multi-GPU check (run under torchrun, one rank per GPU): the NCCL legs of the C ABI --
asscpp_allgather_results and asscpp_allreduce_or -- against the same work done on one GPU."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import asscpp_b200 as A  # noqa: E402
from asscpp_b200 import adaptive as AD  # noqa: E402


def main():
    world, rank, local = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    g = os.path.join(ROOT, "tests", "golden")
    xyz = np.load(os.path.join(g, "etihad_model.npz"))["xyz"]
    ss = np.load(os.path.join(g, "etihad_searchspace_2p5m.npz"))
    poses = np.ascontiguousarray(np.concatenate([ss["cam0"][500:756], ss["cam1"][500:756]]))
    ctx = A.CoverageContext(xyz, A.default_params(), device=local)
    uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
    if rank == 0:
        uid.copy_(torch.from_numpy(A.CoverageContext.comm_unique_id()))
    dist.broadcast(uid, 0)
    ctx.comm_init(world, rank, uid.cpu().numpy())
    # single-GPU truth on this rank's own device
    full = ctx.eval_batch(poses, flags=A.EVAL_HASH)
    s_full = ctx.new_set()
    s_full.or_views(poses, 5)
    truth_pts, truth_vox = s_full.export()
    # sharded: equal blocks (allgather needs equal counts per rank)
    per = len(poses) // world
    lo, hi = rank * per, (rank + 1) * per
    ctx.eval_async(poses[lo:hi], None, A.EVAL_HASH)
    gathered = ctx.allgather_results(per)
    ok = True
    for f in ("survivors", "visible", "gain_pts", "gain_vox", "vis_hash", "steps"):
        if not np.array_equal(gathered[f], full[f][: per * world]):
            ok = False
            print("rank %d: allgather mismatch in %s" % (rank, f))
    ctx.timer_begin()
    ctx.eval_async(poses[lo:hi], None, 0)
    gp, gv = ctx.allgather_gain(per)
    ctx.timer_end()
    (ag_ms, ag_n), _ = ctx.collective_ms()
    if not (np.array_equal(gp, full["gain_pts"][: per * world]) and np.array_equal(gv, full["gain_vox"][: per * world])):
        ok = False
        print("rank %d: gain allgather mismatch" % rank)
    if ag_n != 1 or not ag_ms > 0:
        ok = False
        print("rank %d: collective timing missing (%r, %r)" % (rank, ag_ms, ag_n))
    # unequal counts are refused, not hung on
    ctx.eval_async(poses[lo:hi - rank], None, 0)
    try:
        ctx.allgather_gain(per - rank)
        if world > 1:
            ok = False
            print("rank %d: unequal allgather counts were accepted" % rank)
    except A.AsscppError as e:
        if e.code != -1:
            ok = False
            print("rank %d: unexpected error %r" % (rank, e))
    s = ctx.new_set()
    s.or_views(poses[lo:hi], 5)
    if rank == world - 1 and per * world < len(poses):
        s.or_views(poses[per * world:], 5)
    ctx.allreduce_or(s)
    pts, vox = s.export()
    if not (np.array_equal(pts, truth_pts) and np.array_equal(vox, truth_vox)):
        ok = False
        print("rank %d: OR-allreduce mismatch" % rank)
    npts, nvox = s.popcount()
    # config 4: the adaptive loop with the candidates sharded over the ranks and one covered-bitset OR-allreduce
    # per iteration, against the same loop run on this rank's GPU alone
    mg = np.load(os.path.join(g, "etihad_mesh.npz"))
    mesh = A.Mesh(mg["vertices"], mg["triangles"], device=local)
    model = AD.VoxelModel(xyz, 0.5)

    def gather_counts(n):
        t = torch.tensor([n], device="cuda", dtype=torch.int64)
        dist.all_reduce(t)
        return int(t.item())

    solo = AD.adaptive_loop(AD.GpuBackend(ctx, comm=False, mesh=mesh), model, (0, 0, 4.5), (36, 50, 9))
    shard = AD.adaptive_loop(AD.GpuBackend(ctx, comm=True, mesh=mesh), model, (0, 0, 4.5), (36, 50, 9), world=world,
                             rank=rank, gather_counts=gather_counts)
    if len(solo) != len(shard):
        ok = False
        print("rank %d: adaptive loop length %d vs %d" % (rank, len(shard), len(solo)))
    for a, b in zip(shard, solo):
        if not (np.array_equal(a["bits"], b["bits"]) and a["accepted"] == b["accepted"]
                and a["candidates"] == b["candidates"] and a["clusters"] == b["clusters"]):
            ok = False
            print("rank %d: adaptive loop iteration %d differs" % (rank, a["iter"]))
    if rank == 0:
        print("adaptive loop world=%d: %d iterations, uncovered %s" % (
            world, len(shard), " ".join("%.3f%%" % t["uncovered_pct"] for t in shard)))
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("nccl_check world=%d: %s (covered %d points, %d voxels, %d views gathered)"
              % (world, "OK" if int(flag.item()) else "FAILED", npts, nvox, per * world))
    ctx.close()
    dist.destroy_process_group()
    return 0 if int(flag.item()) else 1


if __name__ == "__main__":
    sys.exit(main())
