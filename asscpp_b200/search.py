"""Host side of the SSPP search around the hot path (BASELINE.json config 1): the search-space
wire formats, node connection, and the A* loop with ONE batched scoring call per expansion.

Restates, with the reference's list semantics kept literally:
  * SearchSpace.insert        = SearchSpace::insertNode (sspp/src/searchspace.cpp:64-105): duplicates dropped, HEAD
                                insertion -- list order is the reverse of insertion order
  * load_regular_grid()       = PathPlanner::loadRegularGrid (sspp/src/pathplanner.cpp:204-267): three text files,
                                `x y z qx qy qz qw` per line (robot, sensor 0, sensor 1)
  * save() / load()           = saveSearchSpace / loadSearchSpace (:269-361): one line per connection, 14 numbers,
                                fixed notation with 3 decimals; sensors recomputed with robot2sensor on load
  * connect()                 = connectNodes (:534-568): all pairs within the connection radius, collision test on
                                the segment between different positions (isConnectionConditionSatisfied,
                                cscpp/src/coverage_path_planning_heuristic.cpp:300-315 -> asscpp_mesh_segments_free)
  * astar_search()            = Astar::astarSearch / findRoot / makeChildrenNodes (sspp/src/astar.cpp:74-133,
                                220-373, 375-458) with LList::add / find / remove (sspp/src/llist.cpp:41-146)
  * SurfaceCoverageH scoring  = calculateHeuristic (heuristic.cpp:425-488, 763-781, root 1227-1240),
                                terminateConditionReached (:253-283)

Quirk Q10 (kept by default, `accumulate=False`): the node popped from the open list is COPIED with
`new Node(openList->getHead())` (astar.cpp:262) and Node::Node(Node*) gives the copy a fresh, empty cloud_filtered
(sspp/src/node.cpp:43-68); children hang off the copy (astar.cpp:440), so `node->parent->cloud_filtered`
(heuristic.cpp:473) is empty for every child: a node's coverage is the coverage of its OWN sensors' visible cloud,
while `c = node->coverage - node->parent->coverage` still subtracts the parent's number.  `accumulate=True` gives
the evidently intended behaviour (the covered set grows along the path) through device-resident covered sets.

The `scorer` is duck-typed (GpuScorer here, OracleScorer in tests/backends.py): the same loop drives the CUDA path
and the CPU oracle, which is how the parity of config 1 is asserted.
"""
import numpy as np


def _pose_equal(a, b):
    """isPositionEqual && isOrientationEqual (sspp/include/sspp/utils.h:241-259)."""
    return bool((np.abs(a - b) < 1e-12).all())


def dist(a, b):
    """Dist(Pose, Pose), utils.h:356-359."""
    return float(np.sqrt((a[0] - b[0]) * (a[0] - b[0]) + (a[1] - b[1]) * (a[1] - b[1]) + (a[2] - b[2]) * (a[2] - b[2])))


class SearchSpace:
    """Waypoints in LIST order (head = last inserted): robots [n, 7], sensors [n, S, 7], ids [n], children[i] =
    indices of the nodes reachable from node i, in the order connectNodes pushed them."""

    def __init__(self):
        self._robots, self._sensors, self._ids = [], [], []
        self._seen = {}
        self.children = None
        self.n_connections = 0

    # -- construction -------------------------------------------------------------------------------
    def insert(self, robot, sensors):
        """insertNode: returns False when a node with this pose exists (nodeExists, searchspace.cpp:72)."""
        robot = np.asarray(robot, np.float64)
        key = robot.tobytes()  # poses closer than 1e-12 are, for files and generated grids, identical doubles
        if key in self._seen:
            return False
        self._seen[key] = len(self._robots)
        self._robots.append(robot)
        self._sensors.append(np.asarray(sensors, np.float64).reshape(-1, 7))
        self._ids.append(len(self._ids))  # idCount++ (searchspace.cpp:66)
        self.children = None
        return True

    @classmethod
    def from_poses(cls, robots, sensors):
        s = cls()
        for r, c in zip(np.asarray(robots, np.float64), np.asarray(sensors, np.float64)):
            s.insert(r, c)
        return s

    @property
    def n(self):
        return len(self._robots)

    @property
    def robots(self):  # list order: head first
        return np.array(self._robots[::-1]).reshape(-1, 7)

    @property
    def sensors(self):
        return np.array(self._sensors[::-1])

    @property
    def ids(self):
        return np.array(self._ids[::-1], np.int64)

    # -- connectNodes (pathplanner.cpp:534-568) -----------------------------------------------------
    def connect(self, radius, segments_free=None, block=256):
        """children[i] for every node; `segments_free(a[m, 3], b[m, 3]) -> bool[m]` is the collision test between
        different positions (None: no mesh, every segment is free).  One batched call per block of nodes."""
        R = self.robots
        pos = np.ascontiguousarray(R[:, :3])
        n = len(R)
        self.children = [None] * n
        total = 0
        # candidate pairs from a k-d tree (a superset: slightly larger radius), then the reference's own test
        # Dist(S, temp) <= radius in double on every candidate
        try:
            from scipy.spatial import cKDTree
            tree = cKDTree(pos)
        except ImportError:  # all pairs of the block instead
            tree = None
        for lo in range(0, n, block):
            hi = min(lo + block, n)
            if tree is not None:
                cand = tree.query_ball_point(pos[lo:hi], radius * (1.0 + 1e-9) + 1e-9)
                ti = np.repeat(np.arange(hi - lo), [len(c) for c in cand])
                si = np.fromiter((j for c in cand for j in c), np.int64, count=len(ti))
            else:
                ti, si = (a.ravel() for a in np.meshgrid(np.arange(hi - lo), np.arange(n), indexing="ij"))
            d = pos[lo + ti] - pos[si]
            dd = np.sqrt(d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1] + d[:, 2] * d[:, 2])  # Dist()
            sel = (dd <= radius) & (si != lo + ti)  # S != temp
            ti, si, d = ti[sel], si[sel], d[sel]
            same = (d == 0).all(-1)
            ok = same.copy()  # same position, other orientation: always connected (:553-561)
            if segments_free is not None and (~same).any():
                ok[~same] = np.asarray(segments_free(pos[lo + ti[~same]], pos[si[~same]]), bool)
            else:
                ok[:] = True
            ti, si = ti[ok], si[ok]
            order = np.lexsort((si, ti))  # children of a node in ascending S = list order
            ti, si = ti[order], si[order]
            cuts = np.searchsorted(ti, np.arange(hi - lo + 1))
            for t in range(hi - lo):
                self.children[lo + t] = si[cuts[t]:cuts[t + 1]]
            total += len(si)
        self.n_connections = total  # MAXNODES (:568)
        return total

    # -- wire formats -------------------------------------------------------------------------------
    def save(self, path):
        """saveSearchSpace (:269-299): `parent pose, child pose` per connection, fixed, 3 decimals."""
        R = self.robots
        with open(path, "w") as f:
            for i in range(self.n):
                for c in self.children[i]:
                    f.write(" ".join("%.3f" % v for v in np.concatenate([R[i], R[c]])) + "\n")

    @classmethod
    def load(cls, path, robot2sensor, mounts):
        """loadSearchSpace (:301-361): both ends inserted (sensors from robot2sensor), child pushed per line."""
        s = cls()
        edges = []
        with open(path) as f:
            for line in f:
                if line.startswith("#"):
                    continue
                v = np.array(line.split(), np.float64)
                if len(v) < 14:
                    continue
                ends = []
                for p in (v[:7], v[7:14]):
                    s.insert(p, [robot2sensor(p, off, rpy) for off, rpy in mounts])
                    ends.append(s._seen[p.tobytes()])
                edges.append(ends)
        n = s.n
        kids = [[] for _ in range(n)]
        for a, b in edges:  # insertion indices -> list indices
            kids[n - 1 - a].append(n - 1 - b)
        s.children = [np.array(k, np.int64) for k in kids]
        s.n_connections = len(edges)
        return s


def load_regular_grid(robot_file, sensor0_file, sensor1_file):
    """loadRegularGrid (pathplanner.cpp:204-267)."""
    cols = [np.loadtxt(p, dtype=np.float64, ndmin=2)[:, :7] for p in (robot_file, sensor0_file, sensor1_file)]
    n = min(len(c) for c in cols)
    return SearchSpace.from_poses(cols[0][:n], np.stack([cols[1][:n], cols[2][:n]], 1))


def save_regular_grid(space, robot_file, sensor0_file, sensor1_file):
    """The `SearchSpaceUAV_* / SearchSpaceCam_*_{0,1}` triplet (cscpp/txt/, written by the reference's
    coverage_heuristic_test.cpp with operator<< on doubles), in insertion order."""
    R, S = space.robots[::-1], space.sensors[::-1]
    for path, arr in ((robot_file, R), (sensor0_file, S[:, 0]), (sensor1_file, S[:, 1])):
        with open(path, "w") as f:
            for p in arr:
                f.write(" ".join("%.6g" % v for v in p) + "\n")


# ---------------------------------------------------------------------------------------------------
# scoring
# ---------------------------------------------------------------------------------------------------
def coverage_percent(n_voxels, n_cov):
    """calcCoveragePercent's float32 arithmetic (occlusion_culling.cpp:401, 431-432)."""
    return float(np.float32(np.float32(np.float32(n_voxels) / np.float32(n_cov)) * np.float32(100)))


class GpuScorer:
    """SurfaceCoverageH child scoring through asscpp_eval_groups: all children of an expansion, all their sensors,
    one launch.  A node's `state` is its covered set (accumulate=True) or None (quirk Q10)."""

    def __init__(self, ctx, accumulate=False):
        self.ctx, self.accumulate = ctx, accumulate
        self.n_cov = ctx.n_cov_voxels
        self.launches = 0
        self.slow_groups = 0

    def _exact(self, parent_set, sensors):
        """clone + OR (the per-child path of the C++ shim): exact also when a view took the Q5 replay."""
        s = self.ctx.new_set() if parent_set is None else parent_set.clone()
        s.or_views(sensors, -1)
        return s

    def _materialise(self, state):
        """covered set of a node whose scoring only produced a count: parent's set OR its own views"""
        if state is None or state.get("set") is not None:
            return state
        parent = self._materialise(state["parent"])
        state["set"] = self._exact(None if parent is None else parent["set"], state["sensors"])
        return state

    def score(self, parent_state, sensors):
        """sensors [n, S, 7] -> (coverage % per candidate as the reference's float, states)."""
        sensors = np.asarray(sensors, np.float64)
        n, S = sensors.shape[0], sensors.shape[1]
        from . import VS_ORDERED
        parent_set, base = None, 0
        if self.accumulate and parent_state is not None:
            parent_set = self._materialise(parent_state)["set"]
            base = parent_set.popcount()[1]
        res = self.ctx.eval_groups(sensors.reshape(-1, 7), S, covered=parent_set)
        self.launches += 1
        gain = res["gain_vox"].astype(np.int64).reshape(n, S).sum(1)
        replay = ((res["status"] & VS_ORDERED) != 0).reshape(n, S).any(1)
        cov, states = [], []
        for i in range(n):
            st = {"parent": parent_state if self.accumulate else None, "sensors": sensors[i], "set": None}
            nv = base + int(gain[i])
            if replay[i]:
                self.slow_groups += 1
                st["set"] = self._exact(parent_set, sensors[i])
                nv = st["set"].popcount()[1]
            cov.append(coverage_percent(nv, self.n_cov))
            states.append(st)
        return cov, states


# ---------------------------------------------------------------------------------------------------
# A*
# ---------------------------------------------------------------------------------------------------
class _Node:
    __slots__ = ("id", "ss", "pose", "parent", "depth", "f", "distance", "coverage", "state")


def _llist_add(lst, node):
    """LList::add for a reward (isCost() false): before the first element with f <= the new f (llist.cpp:83-124)."""
    for k, p in enumerate(lst):
        if p.f <= node.f:
            lst.insert(k, node)
            return
    lst.append(node)


def _llist_find(lst, node):
    """LList::find: first element with an equal pose (Node::operator==, node.cpp:75-78).  Every node carries the
    pose of a search-space waypoint and those are unique (insertNode drops duplicates), so equal pose <=> same
    waypoint index."""
    for k, p in enumerate(lst):
        if p.ss == node.ss:
            return k
    return -1


def astar_search(space, start, scorer, coverage_target, coverage_tolerance, max_expansions=None, log=None):
    """Astar::astarSearch with SurfaceCoverageH.  Returns {"path": [search-space index], "coverage": [...],
    "f": [...], "distance": total, "expanded": n, "created": n} or None (no path / MAXNODES, astar.cpp:358-372)."""
    R, SEN = space.robots, space.sensors
    start = np.asarray(start, np.float64)
    # findRoot (:74-133): nearest node, first one wins ties
    if space.n == 0:
        return None
    dx, dy, dz = R[:, 0] - start[0], R[:, 1] - start[1], R[:, 2] - start[2]
    best = int(np.argmin(np.sqrt(dx * dx + dy * dy + dz * dz)))  # first minimum = strict `<` in list order
    root = _Node()
    root.id, root.ss, root.pose, root.parent, root.depth, root.distance = 0, best, R[best], None, 0, 0.0
    cov, st = scorer.score(None, SEN[best][None])
    root.coverage, root.state = cov[0], st[0]
    root.f = 0.0  # root: coverage from its own views (heuristic.cpp:1227-1240), f zeroed (:1243-1247)
    open_list, closed = [root], []
    global_id, expanded = 0, 0
    while open_list:
        current = open_list.pop(0)          # :262-263 head = highest reward
        _llist_add(closed, current)         # :266
        expanded += 1
        # terminateConditionReached (:253-283); `current != root` compares a fresh copy's address: always true
        if coverage_target - current.coverage <= coverage_tolerance:
            path, p = [], current
            while p is not None:
                path.append(p)
                p = p.parent
            path.reverse()
            return {"path": [p.ss for p in path], "coverage": [p.coverage for p in path], "f": [p.f for p in path],
                    "distance": current.distance, "expanded": expanded, "created": global_id}
        # makeChildrenNodes (:375-458): locate the node by pose, children linked in reverse
        # (the reference scans the list for the first node with this pose: poses are unique, so it is the node itself)
        kids = list(space.children[current.ss])
        nodes = []
        for c in kids:
            nd = _Node()
            nd.ss, nd.pose, nd.parent, nd.depth, nd.id = int(c), R[c], current, current.depth + 1, global_id
            global_id += 1
            nodes.append(nd)
        if nodes:
            covs, states = scorer.score(current.state, SEN[[nd.ss for nd in nodes]])  # ONE launch per expansion
            for nd, cv, st in zip(nodes, covs, states):
                d = dist(nd.pose, current.pose)
                nd.distance = current.distance + d                       # heuristic.cpp:452
                nd.coverage, nd.state = cv, st
                c = nd.coverage - current.coverage                        # :483
                nd.f = current.f + ((1 / d) * c) if d != 0.0 else current.f + c  # :487 / :780
        for nd in reversed(nodes):          # the child list is walked from its head = last child made
            cond = current.f >= nd.f        # :339
            if _llist_find(open_list, nd) >= 0 and cond:
                continue
            if _llist_find(closed, nd) >= 0 and cond:
                continue
            k = _llist_find(open_list, nd)  # :356 remove any similar node from the open list
            if k >= 0:
                open_list.pop(k)
            _llist_add(open_list, nd)
        if log:
            log("expanded %d id %d ss %d coverage %.4f f %.6f open %d" % (expanded, current.id, current.ss,
                                                                        current.coverage, current.f, len(open_list)))
        if current.id > space.n_connections:  # :361 MAXNODES = numConnections
            return None
        if max_expansions is not None and expanded >= max_expansions:
            return None
    return None
