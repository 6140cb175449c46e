// coverage_path_planning_heuristic_b200_impl.h -- member definitions of SSPP::CoveragePathPlanningHeuristic
// (included by coverage_path_planning_heuristic_b200.h; citations are into cscpp/src/coverage_path_planning_heuristic.cpp
// unless another file is named).
#pragma once

namespace SSPP {

inline void CoveragePathPlanningHeuristic::init_defaults() {
  occlussionCulling = NULL;
  mesh_ = NULL;
  coverageTarget = 0; coverageTolerance = 0;
  accuracySum = 0.0; extraCovSum = 0.0;                 // :46-47
  volumetricVoxelRes = 0.5;                             // :49
  accW = 0.8; distW = 0.5; covW = 1; angleW = 0.2;      // :50-53
  selectedPointsNum = 0;                                // :54
  voxelResForConn = 0.5;                                // :55
  maxConnRadius = std::sqrt((4.5 * 4.5) + (4.5 * 4.5)) + 0.01;  // :56
  maxDepth = 0;
  sibling_source_ = NULL;
  cache_parent_ = NULL;
  parent_set_valid_for_ = NULL;
  parent_vox_cached_ = 0;
  parent_set_ = global_set_ = scratch_set_ = -1;
  launches_scored_ = nodes_scored_ = 0;
  fillCloudFiltered = false;
  appendGlobalCloud = true;
  info_cells_ = 0; modelVolume = 0; modelTotalEntroby = 0;
  intended_info_gain_ = false;
  info_cache_parent_ = NULL;
}

inline void CoveragePathPlanningHeuristic::finish_init() {
  if (heuristicType == SurfaceAreaCoverageH)
    throw std::runtime_error("CoveragePathPlanningHeuristic: SurfaceAreaCoverageH is commented out in the reference");
  check(asscpp_covered_new(ctx(), &parent_set_));
  check(asscpp_covered_new(ctx(), &global_set_));
  check(asscpp_covered_new(ctx(), &scratch_set_));
}

inline void CoveragePathPlanningHeuristic::set_mesh(const std::vector<float>& vertices, const std::vector<int32_t>& triangles) {
  // cgalTree = new Tree1(triangles.begin(), triangles.end())  (:31)
  if (asscpp_mesh_create(OcclusionCulling::device_from_env(), vertices.data(), (int64_t)vertices.size() / 3, triangles.data(),
                         (int64_t)triangles.size() / 3, &mesh_) != ASSCPP_OK)
    throw std::runtime_error(std::string("asscpp_mesh_create: ") + asscpp_mesh_last_error(NULL));
}

inline void CoveragePathPlanningHeuristic::loadOBJFile(const char* filename, std::vector<float>& vertices,
                                                       std::vector<int32_t>& triangles) {
  FILE* file = std::fopen(filename, "rb");
  if (!file) { std::fprintf(stderr, "file not exist\n"); return; }  // :1358-1362
  bool has_normal = false, has_texture = false;
  char line_buffer[2000];
  while (std::fgets(line_buffer, 2000, file)) {
    char* first_token = std::strtok(line_buffer, "\r\n\t ");
    if (!first_token || first_token[0] == '#' || first_token[0] == 0) continue;
    if (first_token[0] == 'v') {
      if (first_token[1] == 'n') { std::strtok(NULL, "\t "); std::strtok(NULL, "\t "); std::strtok(NULL, "\t "); has_normal = true; }
      else if (first_token[1] == 't') { std::strtok(NULL, "\t "); std::strtok(NULL, "\t "); has_texture = true; }
      else {
        float x = (float)std::atof(std::strtok(NULL, "\t "));
        float y = (float)std::atof(std::strtok(NULL, "\t "));
        float z = (float)std::atof(std::strtok(NULL, "\t "));
        vertices.push_back(x); vertices.push_back(y); vertices.push_back(z);
      }
    } else if (first_token[0] == 'f') {
      char* data[30];
      int n = 0;
      while (n < 29 && (data[n] = std::strtok(NULL, "\t \r\n")) != NULL)
        if (std::strlen(data[n])) n++;
      for (int t = 0; t < (n - 2); ++t) {
        int32_t idx[3];
        if (!has_texture && !has_normal) {
          idx[0] = std::atoi(data[0]) - 1; idx[1] = std::atoi(data[1]) - 1; idx[2] = std::atoi(data[2]) - 1;  // :1404-1406
        } else {
          for (int i = 0; i < 3; ++i) idx[i] = std::atoi(i == 0 ? data[0] : data[t + i]) - 1;              // :1416-1427
        }
        // collinear triangles are dropped by asscpp_mesh_create, as CGAL::collinear does here (:1409, :1430)
        triangles.push_back(idx[0]); triangles.push_back(idx[1]); triangles.push_back(idx[2]);
      }
    }
  }
  std::fclose(file);
}

// <tf Quaternion::angleShortestPath>
inline double CoveragePathPlanningHeuristic::shortest_angle(const geometry_msgs::Pose& pp, const geometry_msgs::Pose& pn) {
  const double l2p = pp.orientation.x * pp.orientation.x + pp.orientation.y * pp.orientation.y +
                     pp.orientation.z * pp.orientation.z + pp.orientation.w * pp.orientation.w;
  const double l2n = pn.orientation.x * pn.orientation.x + pn.orientation.y * pn.orientation.y +
                     pn.orientation.z * pn.orientation.z + pn.orientation.w * pn.orientation.w;
  const double sq = std::sqrt(l2p * l2n);
  const double dt = pp.orientation.x * pn.orientation.x + pp.orientation.y * pn.orientation.y +
                    pp.orientation.z * pn.orientation.z + pp.orientation.w * pn.orientation.w;
  return dt < 0 ? std::acos(-dt / sq) * 2.0 : std::acos(dt / sq) * 2.0;
}

// The set a child is scored against = the coverage voxels of node->parent->cloud_filtered (:473, :499, :766, ...).
// Empty cloud (what the reference's A* always hands over, quirk Q10) -> -1, the empty set.
inline int CoveragePathPlanningHeuristic::parent_covered_set(const Node* parent, int64_t* parent_vox) {
  *parent_vox = 0;
  if (!parent || !parent->cloud_filtered || parent->cloud_filtered->points.empty()) return -1;
  const std::vector<pcl::PointXYZ>& pts = parent->cloud_filtered->points;
  std::vector<float> xyz(3 * pts.size());
  for (size_t i = 0; i < pts.size(); ++i) { xyz[3 * i] = pts[i].x; xyz[3 * i + 1] = pts[i].y; xyz[3 * i + 2] = pts[i].z; }
  check(asscpp_covered_clear(ctx(), parent_set_));
  check(asscpp_covered_or_cloud(ctx(), parent_set_, xyz.data(), (int64_t)pts.size()));
  int64_t np = 0;
  check(asscpp_covered_popcount(ctx(), parent_set_, &np, parent_vox));
  return parent_set_;
}

// one launch for the sensors of all the given nodes (they share the parent, hence the covered set)
inline void CoveragePathPlanningHeuristic::score_nodes(const std::vector<Node*>& nodes, int covered_set,
                                                       std::vector<Score>& out) {
  out.assign(nodes.size(), Score());
  if (nodes.empty()) return;
  const size_t per = nodes[0]->senPoses.size();
  bool uniform = per > 0;
  for (size_t i = 0; i < nodes.size(); ++i) uniform = uniform && nodes[i]->senPoses.size() == per;
  if (!uniform) {  // nodes with differing sensor counts (never produced by the planners): one launch each
    for (size_t i = 0; i < nodes.size(); ++i) {
      std::vector<Score> one;
      if (nodes[i]->senPoses.empty()) { Score z; z.gain_vox = 0; z.visible = 0; z.err_sum_z = 0; z.exact = true; out[i] = z; continue; }
      score_nodes(std::vector<Node*>(1, nodes[i]), covered_set, one);
      out[i] = one[0];
    }
    return;
  }
  std::vector<double> p(7 * per * nodes.size());
  for (size_t i = 0; i < nodes.size(); ++i)
    for (size_t s = 0; s < per; ++s) OcclusionCulling::pack(nodes[i]->senPoses[s].p, &p[7 * (i * per + s)]);
  AsscppEvalOpts o;
  std::memset(&o, 0, sizeof(o));
  o.covered_set = covered_set; o.accum_set = -1; o.group_size = (int32_t)per;
  o.flags = needs_accuracy() ? (uint32_t)ASSCPP_EVAL_ACCURACY : 0u;
  const int64_t nviews = (int64_t)(per * nodes.size());
  check(asscpp_eval_ex(ctx(), p.data(), nviews, &o));
  std::vector<AsscppViewResult> res((size_t)nviews);
  check(asscpp_get_results(ctx(), res.data(), nviews));
  std::vector<AsscppViewAccuracy> acc;
  if (needs_accuracy()) {
    acc.resize((size_t)nviews);
    check(asscpp_get_accuracy(ctx(), acc.data(), nviews));
  }
  launches_scored_ += 1;
  for (size_t i = 0; i < nodes.size(); ++i) {
    Score sc;
    sc.gain_vox = 0; sc.visible = 0; sc.err_sum_z = 0; sc.exact = true;
    for (size_t s = 0; s < per; ++s) {
      const AsscppViewResult& r = res[i * per + s];
      sc.gain_vox += r.gain_vox;
      sc.visible += r.visible;
      if (r.status & ASSCPP_VS_ORDERED) sc.exact = false;
      if (needs_accuracy()) sc.err_sum_z += acc[i * per + s].err_sum_z;
    }
    out[i] = sc;
  }
}

// exact union for a node whose group saw a Q5 replay: clone + OR, popcount difference
inline CoveragePathPlanningHeuristic::Score CoveragePathPlanningHeuristic::score_exact(Node* node, int covered_set,
                                                                                       int64_t parent_vox) {
  std::vector<Score> one;
  score_nodes(std::vector<Node*>(1, node), covered_set, one);
  Score sc = one[0];
  check(asscpp_covered_clear(ctx(), scratch_set_));
  if (covered_set >= 0) check(asscpp_covered_or_set(ctx(), scratch_set_, covered_set));
  std::vector<double> p(7 * node->senPoses.size());
  for (size_t s = 0; s < node->senPoses.size(); ++s) OcclusionCulling::pack(node->senPoses[s].p, &p[7 * s]);
  check(asscpp_covered_or_views(ctx(), scratch_set_, p.data(), (int64_t)node->senPoses.size(), -1));
  int64_t np = 0, nv = 0;
  check(asscpp_covered_popcount(ctx(), scratch_set_, &np, &nv));
  sc.gain_vox = (unsigned long long)(nv - parent_vox);
  sc.exact = true;
  return sc;
}

// Node::cloud_filtered = VoxelGrid(0.5) of parent's cloud + visibleCloud (:473-480): here one point per covered voxel
inline void CoveragePathPlanningHeuristic::fill_cloud_filtered(Node* node, int covered_set) {
  check(asscpp_covered_clear(ctx(), scratch_set_));
  if (covered_set >= 0) check(asscpp_covered_or_set(ctx(), scratch_set_, covered_set));
  std::vector<double> p(7 * node->senPoses.size());
  for (size_t s = 0; s < node->senPoses.size(); ++s) OcclusionCulling::pack(node->senPoses[s].p, &p[7 * s]);
  if (!node->senPoses.empty())
    check(asscpp_covered_or_views(ctx(), scratch_set_, p.data(), (int64_t)node->senPoses.size(), -1));
  check(asscpp_covered_invert_voxels(ctx(), scratch_set_));  // the voxels the set HAS become the ones it "lacks"
  AsscppModelInfo info;
  asscpp_model_info(ctx(), &info);
  std::vector<float> ctr(3 * (size_t)std::max<int64_t>(info.n_cov_voxels, 1));
  int64_t n = 0;
  check(asscpp_uncovered_voxels(ctx(), scratch_set_, NULL, ctr.data(), info.n_cov_voxels, &n));
  node->cloud_filtered->points.resize((size_t)n);
  for (int64_t i = 0; i < n; ++i) node->cloud_filtered->points[(size_t)i] = pcl::PointXYZ(ctr[3 * i], ctr[3 * i + 1], ctr[3 * i + 2]);
  node->cloud_filtered->width = (uint32_t)n;
}

}  // namespace SSPP

#include "coverage_path_planning_heuristic_b200_impl2.h"
