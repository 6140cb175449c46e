// hostloop.cpp -- the C++ host side of the ASSCPP adaptive sampling loop, batched:
//   PathPlanner::generateRegularGrid               sspp/src/pathplanner.cpp:106-201  (candidate lattice, loop order z x y yaw)
//   PathPlanner::dynamicNodesGenerationAndConnection  :729-850                         (stages, clusters, cluster boxes)
//   SearchSpace::insertNode / nodeExists           sspp/src/searchspace.cpp:62-175   (prepend, duplicates dropped)
// The reference evaluates one candidate per isFilteringConditionSatisfied call (:160).  Whether a candidate is accepted
// does not depend on the candidates before it (the mesh filters and "its sensors see more than 5 points",
// cscpp/src/coverage_path_planning_heuristic.cpp:317-364); only the covered set accumulates, and it is read once per
// stage (:771).  So a whole stage -- every candidate of every cluster grid -- is ONE mesh launch and ONE evaluation
// launch sharded over all GPUs (asscpp_multi_eval, accumulating into the covered set), followed by the uncovered-region
// extraction on the device.  The waypoints come out in the search space's order (newest first), equal row for row to
// what the reference's own loop produces (tests/test_gpu_parity.py::test_adaptive_sampling_host_loop).
// Node connection (connectClustersInternalNodes, connectToNN) is graph building outside the coverage path; see
// asscpp_mesh_segments_free for its geometric predicate.
#include <cmath>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/asscpp_gpu.h"

namespace {

struct Candidate { double pose[7]; };

// pathplanner.cpp:122-147: for z, for x, for y (inclusive upper bounds, += gridResolution held as float), 360/orientationRes
// yaws starting at 0; tf::createQuaternionFromYaw(yaw) = (0, 0, sin(yaw/2), cos(yaw/2)) in <tf setRPY> arithmetic
void regular_grid(const double start[3], const double size[3], float res, float orientation_res, std::vector<Candidate>& out) {
  const int n_yaw = (int)(360.0f / orientation_res);
  for (double z = (start[2] - size[2] / 2.0); z <= (start[2] + size[2] / 2.0); z += res)
    for (double x = (start[0] - size[0] / 2.0); x <= (start[0] + size[0] / 2.0); x += res)
      for (double y = (start[1] - size[1] / 2.0); y <= (start[1] + size[1] / 2.0); y += res) {
        double yaw = 0.0;
        for (int i = 0; i < n_yaw; ++i) {
          const double hy = yaw * 0.5, cy = std::cos(hy), sy = std::sin(hy);
          const double cp = std::cos(0.0), sp = std::sin(0.0), cr = std::cos(0.0), sr = std::sin(0.0);
          Candidate c;
          c.pose[0] = x; c.pose[1] = y; c.pose[2] = z;
          c.pose[3] = sr * cp * cy - cr * sp * sy;
          c.pose[4] = cr * sp * cy + sr * cp * sy;
          c.pose[5] = cr * cp * sy - sr * sp * cy;
          c.pose[6] = cr * cp * cy + sr * sp * sy;
          out.push_back(c);
          yaw += (orientation_res * M_PI / 180.0f);
        }
      }
}

// sspp/include/sspp/utils.h:386-390 samePosition: every component within 1e-12
bool same_pose(const double* a, const double* b) {
  for (int k = 0; k < 7; ++k)
    if (!(std::fabs(a[k] - b[k]) < 1E-12)) return false;
  return true;
}

}  // namespace

extern "C" {

int64_t asscpp_grid_candidates(const double grid_start[3], const double grid_size[3], double res,
                               double orientation_res_deg, double* robot_poses, int64_t cap) {
  if (!grid_start || !grid_size || !(res > 0) || !(orientation_res_deg > 0)) return ASSCPP_EINVAL;
  std::vector<Candidate> c;
  regular_grid(grid_start, grid_size, (float)res, (float)orientation_res_deg, c);
  for (size_t i = 0; i < c.size() && (int64_t)i < cap && robot_poses; ++i) std::memcpy(robot_poses + 7 * i, c[i].pose, 56);
  return (int64_t)c.size();
}

int asscpp_adaptive_sampling(AsscppMulti* m, AsscppMesh* mesh, const AsscppSamplingOpts* o, int global_set,
                             double* robot_poses, double* sensor_poses, int64_t cap, int64_t* n_out,
                             AsscppSamplingStage* stages, int stage_cap, int* n_stages) {
  if (!m || !mesh || !o || !n_out || cap < 0 || o->n_sensors < 1 || o->n_sensors > 4) return ASSCPP_EINVAL;
  AsscppCtx* c0 = asscpp_multi_context(m, 0);
  AsscppModelInfo info;
  asscpp_model_info(c0, &info);
  const int ns = o->n_sensors;
  std::vector<Candidate> space;                 // insertion order (the search space is its reverse)
  std::vector<std::vector<double> > space_sensors;
  int scratch = -1, rc = ASSCPP_OK, nst = 0;
  if ((rc = asscpp_multi_covered_new(m, &scratch))) return rc;  // mirrored like every set of an AsscppMulti
  double res = o->start_res;
  std::vector<std::pair<std::vector<double>, std::vector<double> > > grids;  // (start, size) of the stage's grids
  grids.push_back(std::make_pair(std::vector<double>(o->grid_start, o->grid_start + 3),
                                 std::vector<double>(o->grid_size, o->grid_size + 3)));
  double uncovered_pct = 100.0;
  for (;;) {
    // ---- one stage: all candidates of all grids ------------------------------------------------------------
    std::vector<Candidate> cand;
    for (size_t g = 0; g < grids.size(); ++g)
      regular_grid(grids[g].first.data(), grids[g].second.data(), (float)res, (float)o->orientation_res_deg, cand);
    AsscppSamplingStage st;
    std::memset(&st, 0, sizeof(st));
    st.res = res; st.candidates = (int64_t)cand.size(); st.grids = (int64_t)grids.size(); st.uncovered_pct_before = uncovered_pct;
    // mesh filters (heuristic.cpp:319-335), one launch for the stage
    std::vector<double> pos(3 * cand.size());
    for (size_t i = 0; i < cand.size(); ++i) std::memcpy(&pos[3 * i], cand[i].pose, 24);
    std::vector<uint8_t> pass(cand.size());
    const double through[3] = {100.0, 10.0, 56.0};
    if (!cand.empty() && (rc = asscpp_mesh_filter(mesh, pos.data(), (int64_t)cand.size(), through, o->min_dist, o->max_dist,
                                                  nullptr, nullptr, pass.data()))) break;
    std::vector<size_t> keep;
    for (size_t i = 0; i < cand.size(); ++i)
      if (pass[i]) keep.push_back(i);
    st.passed_mesh = (int64_t)keep.size();
    // sensor poses (pathplanner.cpp:142-146), then ONE evaluation over all GPUs: visible counts + union of the clouds
    // of more than 5 points into the covered set (heuristic.cpp:339-354)
    std::vector<double> rp(7 * keep.size()), sp(7 * keep.size() * ns);
    for (size_t k = 0; k < keep.size(); ++k) std::memcpy(&rp[7 * k], cand[keep[k]].pose, 56);
    std::vector<double> one(7 * keep.size());
    for (int s = 0; s < ns && !keep.empty(); ++s) {
      if ((rc = asscpp_robot2sensor_batch(rp.data(), (int64_t)keep.size(), o->sensor_offset[s], o->sensor_rpy[s], one.data()))) break;
      for (size_t k = 0; k < keep.size(); ++k) std::memcpy(&sp[7 * (k * ns + s)], &one[7 * k], 56);
    }
    if (rc) break;
    std::vector<AsscppViewResult> res_v(keep.size() * ns);
    if (!keep.empty()) {
      AsscppEvalOpts eo;
      std::memset(&eo, 0, sizeof(eo));
      eo.covered_set = -1; eo.accum_set = global_set; eo.accum_min_visible = 5; eo.group_size = 1;
      if ((rc = asscpp_multi_eval(m, sp.data(), (int64_t)(keep.size() * ns), &eo, res_v.data(), nullptr, nullptr))) break;
    }
    for (size_t k = 0; k < keep.size(); ++k) {
      unsigned long long total = 0;
      for (int s = 0; s < ns; ++s) total += res_v[k * ns + s].visible;
      if (total <= 5) continue;                                              // heuristic.cpp:357
      ++st.accepted;
      bool dup = false;                                                      // searchspace.cpp:62-104 via nodeExists
      for (size_t j = 0; j < space.size() && !dup; ++j) dup = same_pose(space[j].pose, &rp[7 * k]);
      if (dup) continue;
      Candidate c;
      std::memcpy(c.pose, &rp[7 * k], 56);
      space.push_back(c);
      space_sensors.push_back(std::vector<double>(sp.begin() + 7 * k * ns, sp.begin() + 7 * (k + 1) * ns));
    }
    // ---- uncovered part (pathplanner.cpp:769-786): pointCloudDiff with its index quirk, clusters, cluster boxes ----
    const int64_t nc = info.n_cov_voxels;
    std::vector<float> ctr(3 * (size_t)(nc > 0 ? nc : 1));
    int64_t nun = 0;
    if ((rc = asscpp_uncovered_voxels_ex(c0, global_set, ASSCPP_DIFF_PCL_ALIASING, nullptr, ctr.data(), nc, &nun))) break;
    uncovered_pct = nc > 0 ? ((double)nun / (double)nc) * 100 : 0.0;
    st.uncovered_pct_after = uncovered_pct;
    std::vector<AsscppCluster> cl((size_t)(nc > 0 ? nc : 1));
    int64_t ncl = 0;
    if (uncovered_pct > o->stop_uncovered_pct && nun > 0) {
      if ((rc = asscpp_covered_clear(c0, scratch))) break;
      if ((rc = asscpp_covered_or_cloud(c0, scratch, ctr.data(), nun))) break;
      if ((rc = asscpp_covered_invert_voxels(c0, scratch))) break;
      if ((rc = asscpp_uncovered_clusters(c0, scratch, 0.8, 5, 50000, nullptr, cl.data(), nc, &ncl))) break;
    }
    st.clusters = ncl;
    if (stages && nst < stage_cap) stages[nst] = st;
    ++nst;
    if (uncovered_pct <= o->stop_uncovered_pct) break;                       // :776
    res -= o->res_decrement;                                                 // :792
    if (!(res > 0)) break;                                                   // :795, :843
    grids.clear();
    for (int64_t k = 0; k < ncl; ++k)                                        // :799-804 findClusterBB per cluster
      grids.push_back(std::make_pair(std::vector<double>(cl[(size_t)k].grid_start, cl[(size_t)k].grid_start + 3),
                                     std::vector<double>(cl[(size_t)k].grid_size, cl[(size_t)k].grid_size + 3)));
    // findClusterBB leaves gridStart.position.z at the cluster's lowest cell centre (:244) and generateRegularGrid
    // centres its z range on it (:122): keep exactly that
  }
  asscpp_multi_covered_free(m, scratch);
  if (rc) return rc;
  if (n_stages) *n_stages = nst;
  *n_out = (int64_t)space.size();
  for (size_t i = 0; i < space.size() && (int64_t)i < cap; ++i) {
    const size_t src = space.size() - 1 - i;   // search-space order: newest first
    if (robot_poses) std::memcpy(robot_poses + 7 * i, space[src].pose, 56);
    if (sensor_poses) std::memcpy(sensor_poses + 7 * ns * i, space_sensors[src].data(), 56 * (size_t)ns);
  }
  return ASSCPP_OK;
}

}  // extern "C"
