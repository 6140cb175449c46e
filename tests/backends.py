"""Oracle-side backend of asscpp_b200.adaptive.adaptive_loop (test infrastructure)."""
import numpy as np


class OracleBackend:
    def __init__(self, orc, O, nthreads=1, mesh=None):
        self.o, self.O, self.nthreads, self.mesh = orc, O, nthreads, mesh
        self.cp = np.zeros(orc.pts_words(), np.uint64)
        self.cv = np.zeros(orc.vox_words(), np.uint64)
        self.n_cov = orc.n_cov
        self.allreduce = None  # set by the gloo test

    def robot2sensor(self, rp, off, rpy):
        return self.O.robot2sensor(rp, off, rpy)

    def filter_samples(self, positions):
        if self.mesh is None:
            return None
        return self.mesh.filter(positions, (100.0, 10.0, 56.0), 1.0, 4.0, nthreads=self.nthreads)[2]

    def evaluate_and_accumulate(self, poses, min_visible=5):
        if len(poses) == 0:
            return np.zeros(0, np.int64)
        st = self.o.eval_batch(poses, memo=True, nthreads=self.nthreads)
        self.o.accumulate(poses, self.cp, self.cv, min_visible=min_visible, nthreads=self.nthreads)
        return st[:, self.O.ST_VISIBLE].astype(np.int64)

    def reduce(self):
        if self.allreduce is not None:
            self.cv = self.allreduce(self.cv)
            self.cp = self.allreduce(self.cp)

    def covered_voxels(self):
        return self.cv

    def uncovered_regions(self):
        ids, ctr = self.o.uncovered(self.cv)
        labels, k, _ = self.O.euclidean_clusters(ctr, 0.8, 5, 50000)
        return len(ids), [self.O.cluster_bb(ctr[labels == c], 0.5) for c in range(k)]


class OracleScorer:
    """Oracle-side scorer of asscpp_b200.search.astar_search: calculateHeuristic's SurfaceCoverageH blocks
    (heuristic.cpp:439-445, 470-488) literally -- visible cloud of every sensor, union with the parent's covered
    voxels, coverage percent -- one child at a time."""

    def __init__(self, orc, O, accumulate=False):
        self.o, self.O, self.accumulate = orc, O, accumulate
        self.ids = orc.cov_ids()
        self.cache = {}

    def _voxels(self, pose):
        key = pose.tobytes()
        if key not in self.cache:
            self.cache[key] = np.unique(self.ids[self.o.eval_view(pose)["visible"]])
        return self.cache[key]

    def score(self, parent_state, sensors):
        from asscpp_b200.search import coverage_percent
        cov, states = [], []
        for sen in np.asarray(sensors, np.float64):
            own = [self._voxels(p) for p in sen]
            base = parent_state if (self.accumulate and parent_state is not None) else np.zeros(0, np.int64)
            u = np.unique(np.concatenate([base] + own))
            cov.append(coverage_percent(len(u), self.o.n_cov))
            states.append(u)
        return cov, states
