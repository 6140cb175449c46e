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
