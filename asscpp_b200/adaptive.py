"""Host side of the ASSCPP adaptive sampling loop (BASELINE.json config 4) and the viewpoint sharding
used by every multi-GPU entry point.

Restates the skeleton of PathPlanner::dynamicNodesGenerationAndConnection (sspp/src/pathplanner.cpp:
729-850) around the hot path:
  * generate_grid()      = generateRegularGrid's z, x, y, yaw loops (pathplanner.cpp:122-147) and the two
                           sensor poses per robot pose (sensors.cpp:47-90, mounts coverage_heuristic_test.cpp:94-95)
  * evaluate + accept    = isFilteringConditionSatisfied's coverage part (heuristic.cpp:339-357), batched
  * uncovered percentage = pointCloudDiff (heuristic.cpp:88-145) on the covered voxel bitset
  * clusters + boxes     = clusteringPointCloud / findClusterBB (heuristic.cpp:147-246)
  * sample filters      = isFilteringConditionSatisfied's geometric part (heuristic.cpp:317-335) through the
                           backend's triangle mesh (asscpp_mesh_filter); a backend built without a mesh falls
                           back to a stand-off test against the voxel lattice (synthetic models have no mesh)
Viewpoint evaluation, the covered-set union, the uncovered
voxel extraction, the Euclidean clustering (tolerance 0.8 m, sizes 5..50000) and the cluster boxes run on the
GPUs (asscpp_uncovered_voxels / asscpp_uncovered_clusters), viewpoints sharded over ranks with ONE
covered-bitset OR-allreduce per iteration.

The `backend` is duck-typed (GpuBackend here, OracleBackend in tests/): the same loop drives the CUDA
path and the CPU oracle, which is how the per-iteration parity of config 4 is asserted.
"""
import numpy as np

MOUNTS = (((0.0, 0.022, 0.065), (0.0, -0.349, 0.0)), ((0.0, 0.022, -0.065), (0.0, 0.349, 0.0)))


def shard_range(n, world, rank):
    """Contiguous block of [0, n) owned by `rank`."""
    base, extra = divmod(n, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def generate_grid(center, size, res, robot2sensor, yaws=8, mounts=MOUNTS):
    """pathplanner.cpp:122-147: z, x, y loops with `<=` bounds and += res, 8 yaws of 45 degrees.
    Returns robot poses [n, 7] and sensor poses [n, len(mounts), 7]."""
    def axis(c, s):
        out, v = [], c - s / 2.0
        while v <= c + s / 2.0:
            out.append(v)
            v += res
        return out
    robots, sensors = [], []
    for z in axis(center[2], size[2]):
        for x in axis(center[0], size[0]):
            for y in axis(center[1], size[1]):
                yaw = 0.0
                for _ in range(yaws):
                    rp = np.array([x, y, z, 0.0, 0.0, np.sin(yaw * 0.5), np.cos(yaw * 0.5)])  # tf createQuaternionFromYaw
                    yaw += 45.0 * np.pi / 180.0
                    robots.append(rp)
                    sensors.append([robot2sensor(rp, off, rpy) for off, rpy in mounts])
    if not robots:
        return np.zeros((0, 7)), np.zeros((0, len(mounts), 7))
    return np.array(robots), np.array(sensors)


class VoxelModel:
    """Host view of the model's coverage lattice: occupied voxel coordinates in compact-id order."""

    def __init__(self, xyz, cov_leaf):
        inv = np.float32(1.0) / np.float32(cov_leaf)
        ijk = np.floor(xyz.astype(np.float32) * inv).astype(np.int64)
        self.lo = ijk.min(0)
        self.div = ijk.max(0) - self.lo + 1
        rel = ijk - self.lo
        key = rel[:, 0] + self.div[0] * (rel[:, 1] + self.div[1] * rel[:, 2])
        self.keys = np.unique(key)  # ascending (k, j, i): the compact ids of the C ABI and the oracle
        k = self.keys
        self.ijk = np.stack([k % self.div[0], (k // self.div[0]) % self.div[1], k // (self.div[0] * self.div[1])], 1) + self.lo
        self.leaf = float(cov_leaf)
        self.centres = (self.ijk + 0.5) * self.leaf
        self.n = len(k)

    def standoff_ok(self, positions, dmin=1.0, dmax=4.0):
        """Stand-in for the CGAL distance filter (heuristic.cpp:328-331): distance from the waypoint to the
        nearest occupied voxel centre within [dmin, dmax]."""
        ok = np.zeros(len(positions), bool)
        for s in range(0, len(positions), 512):
            d = np.sqrt(((positions[s:s + 512, None, :] - self.centres[None, :, :]) ** 2).sum(-1)).min(1)
            ok[s:s + 512] = (d >= dmin) & (d <= dmax)
        return ok


def bits_to_ids(bits, n):
    return np.nonzero(np.unpackbits(bits.view(np.uint8), bitorder="little")[:n])[0]


class GpuBackend:
    """CUDA path: evaluation, union and reduction all through the C ABI."""

    def __init__(self, ctx, comm=False, mesh=None):
        self.ctx, self.comm, self.mesh = ctx, comm, mesh
        self.set = ctx.new_set()
        self.n_cov = ctx.n_cov_voxels

    def robot2sensor(self, rp, off, rpy):
        from . import robot2sensor
        return robot2sensor(rp, off, rpy)

    def filter_samples(self, positions):
        """heuristic.cpp:317-335: outside the mesh and 1..4 m from it; None when there is no mesh."""
        return None if self.mesh is None else self.mesh.filter(positions, (100.0, 10.0, 56.0), 1.0, 4.0)[2]

    def evaluate_and_accumulate(self, poses, min_visible=5):
        """visible count per pose; every view with more than min_visible points joins the covered set."""
        if len(poses) == 0:
            return np.zeros(0, np.int64)
        self.set.or_views(poses, min_visible)          # heuristic.cpp:344-346
        return self.ctx.results(len(poses))["visible"].astype(np.int64)

    def reduce(self):
        if self.comm:
            self.ctx.allreduce_or(self.set)            # covered-bitset OR-allreduce between iterations

    def covered_voxels(self):
        return self.set.export()[1]

    def uncovered_regions(self):
        """(number of uncovered coverage voxels, [(grid_start, grid_size)] per cluster, largest first)."""
        ids, _ = self.set.uncovered()                       # pointCloudDiff, heuristic.cpp:88-145
        _, cl = self.set.uncovered_clusters(0.8, 5, 50000)  # clusteringPointCloud :147-190 + findClusterBB :218-246
        return len(ids), [(c["grid_start"].copy(), c["grid_size"].copy()) for c in cl]


def adaptive_loop(backend, model, grid_center, grid_size, start_res=4.5, res_dec=1.5, world=1, rank=0,
                  gather_counts=None, max_iters=8, log=None):
    """dynamicNodesGenerationAndConnection (pathplanner.cpp:729-850) around the hot path.

    Every rank generates the same candidate list, evaluates its contiguous shard, and the covered bitsets
    are OR-reduced once per iteration; acceptance counts are summed through `gather_counts` (identity for
    one rank).  Returns the per-iteration trace used by the parity tests.
    """
    trace = []
    res = start_res
    boxes = [(np.asarray(grid_center, float), np.asarray(grid_size, float))]
    for it in range(max_iters):
        accepted_total, cand_total = 0, 0
        for center, size in boxes:
            robots, sensors = generate_grid(center, size, res, backend.robot2sensor)
            if len(robots) == 0:
                continue
            keep = backend.filter_samples(robots[:, :3])  # heuristic.cpp:320-335
            if keep is None:
                keep = model.standoff_ok(robots[:, :3])
            robots, sensors = robots[keep], sensors[keep]
            lo, hi = shard_range(len(robots), world, rank)
            flat = sensors[lo:hi].reshape(-1, 7)
            vis = backend.evaluate_and_accumulate(flat, 5).reshape(-1, sensors.shape[1]) if hi > lo else np.zeros((0, 2), np.int64)
            accepted_total += int((vis.sum(1) > 5).sum())  # heuristic.cpp:357
            cand_total += len(robots)
        backend.reduce()
        if gather_counts is not None:
            accepted_total = gather_counts(accepted_total)
        bits = backend.covered_voxels()
        covered = bits_to_ids(bits, model.n)
        n_uncovered, regions = backend.uncovered_regions()
        pct = (n_uncovered / float(model.n)) * 100.0  # heuristic.cpp:143
        trace.append({"iter": it, "res": res, "candidates": cand_total, "accepted": accepted_total,
                      "covered_voxels": int(len(covered)), "uncovered_pct": pct, "bits": bits.copy(),
                      "clusters": len(regions)})
        if log:
            log("iter %d res %.2f candidates %d accepted %d covered %d uncovered %.3f%% clusters %d"
                % (it, res, cand_total, accepted_total, len(covered), pct, len(regions)))
        if pct <= 0.01:          # pathplanner.cpp:776
            break
        res -= res_dec           # pathplanner.cpp:792
        if res <= 0 or not regions:
            break
        # findClusterBB's gridStart is handed to generateRegularGrid, which treats it as the grid CENTRE
        # (pathplanner.cpp:122-126, 802-806): kept as the reference does it
        boxes = regions
    return trace
