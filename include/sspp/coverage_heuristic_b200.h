// coverage_heuristic_b200.h -- the node-scoring side of the boundary: SurfaceCoverageH of
// SSPP::CoveragePathPlanningHeuristic (cscpp/src/coverage_path_planning_heuristic.cpp:425-488, root
// :1227-1240), the acceptance rule of isFilteringConditionSatisfied (:339-357) and pointCloudDiff's
// uncovered percentage (:88-145), all in covered-bitset form on top of the C ABI.
//
// What changes against the reference: Node::cloud_filtered (a re-voxelised point cloud per node) becomes
// a device-resident covered set handle; the per-child call pattern of Astar (sspp/src/astar.cpp:311-357)
// is fed by ViewpointBatcher::prefetch(), which pushes the sensor poses of ALL children of an expansion
// through one asscpp_eval_batch launch and caches the per-view results by node id.
#pragma once
#include <algorithm>
#include <climits>
#include <cmath>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

#include "../cscpp/occlusion_culling_b200.h"

namespace SSPP {

// Minimal stand-in of SSPP::Node (sspp/include/sspp/node.h:48-71) with the fields the coverage
// heuristic reads and writes.  cloud_filtered is replaced by covered_set.
struct NodeB200 {
  int id = 0;
  NodeB200* parent = nullptr;
  geometry_msgs::Pose pose;                   // robot pose (Node::pose.p)
  std::vector<geometry_msgs::Pose> senPoses;  // one per sensor (Node::senPoses[i].p)
  double f_value = 0, g_value = 0, h_value = 0, distance = 0, coverage = 0;
  int covered_set = -1;                       // replaces Node::cloud_filtered
};

class ViewpointBatcher {
 public:
  explicit ViewpointBatcher(OcclusionCulling* oc) : oc_(oc) {}
  // One launch for every sensor pose of every child, scored against the parent's covered set
  // (asscpp_eval_groups: a child's gain_vox sum is the gain of the union of its sensors' visible clouds).
  // Children must carry the same number of sensors (they do: Robot::sensors, sspp/src/astar.cpp:420-433).
  void prefetch(NodeB200* const* children, int n, int parent_set) {
    std::vector<double> p;
    std::vector<std::pair<int, int> > owner;  // (node id, sensor)
    const size_t per = n > 0 ? children[0]->senPoses.size() : 0;
    for (int i = 0; i < n; ++i)
      for (size_t s = 0; s < children[i]->senPoses.size(); ++s) {
        p.resize(p.size() + 7);
        OcclusionCulling::pack(children[i]->senPoses[s], &p[p.size() - 7]);
        owner.push_back(std::make_pair(children[i]->id, (int)s));
      }
    cache_.clear();  // results of an earlier expansion were scored against another set (and handle ids are recycled)
    parent_set_ = -2;
    if (owner.empty() || per == 0) return;
    oc_->check(asscpp_eval_groups(oc_->context(), p.data(), (int64_t)owner.size(), (int)per, parent_set, 0));
    std::vector<AsscppViewResult> res(owner.size());
    oc_->check(asscpp_get_results(oc_->context(), res.data(), (int64_t)res.size()));
    for (size_t k = 0; k < res.size(); ++k) cache_[owner[k]] = res[k];
    parent_set_ = parent_set;
    launches_++;
  }
  bool lookup(int node_id, int sensor, AsscppViewResult* out) const {
    std::map<std::pair<int, int>, AsscppViewResult>::const_iterator it = cache_.find(std::make_pair(node_id, sensor));
    if (it == cache_.end()) return false;
    *out = it->second;
    return true;
  }
  // voxels the node adds to the set it was prefetched against; false when not cached or when one of its views
  // took the ordered replay (quirk Q5), whose group sum is not union-exact
  bool groupGain(const NodeB200* node, int parent_set, unsigned long long* gain) const {
    if (parent_set != parent_set_ || node->senPoses.empty()) return false;
    unsigned long long g = 0;
    for (size_t s = 0; s < node->senPoses.size(); ++s) {
      AsscppViewResult r;
      if (!lookup(node->id, (int)s, &r) || (r.status & ASSCPP_VS_ORDERED)) return false;
      g += r.gain_vox;
    }
    *gain = g;
    return true;
  }
  void clear() { cache_.clear(); parent_set_ = -2; }
  int launches() const { return launches_; }

 private:
  OcclusionCulling* oc_;
  std::map<std::pair<int, int>, AsscppViewResult> cache_;
  int parent_set_ = -2;
  int launches_ = 0;
};

// heuristic types of cscpp/include/cscpp/coverage_path_planning_heuristic.h:72 that live on the hot path
enum { SurfaceCoverageH = 0, SurfaceCoveragewithOrientationH = 1, SurfaceCoveragewithAccuracyH = 2, VolumetricCoverageH = 4 };

class CoverageHeuristicB200 {
 public:
  // accumulate = false (default) is what the reference's A* computes: Astar copies the node it pops (astar.cpp:262) and
  // Node::Node(Node*) gives the copy an empty cloud_filtered (node.cpp:43-68), so every child is scored against the
  // EMPTY set and node->coverage is that of its own views, while c still subtracts the parent's number (quirk Q10).
  // accumulate = true is the evidently intended behaviour: the parent's covered set is carried along the path.
  // The verbatim drop-in (class SSPP::CoveragePathPlanningHeuristic, include/cscpp/coverage_path_planning_heuristic_b200.h)
  // reads node->parent->cloud_filtered like the reference and so behaves either way depending on the caller.
  CoverageHeuristicB200(OcclusionCulling* oc, double coverageTarget, double coverageTolerance,
                        int heuristicType = SurfaceCoverageH, bool accumulate = false)
      : oc_(oc), batcher_(oc), target_(coverageTarget), tol_(coverageTolerance), type_(heuristicType),
        accumulate_(accumulate) {}
  bool accumulate() const { return accumulate_; }
  // the set the children of `parent` are scored against (pass it to batcher().prefetch)
  int scoringSet(const NodeB200* parent) const { return accumulate_ && parent ? parent->covered_set : -1; }
  // free the device set of a node that is discarded (closed-list clean-up)
  void release(NodeB200* node) {
    if (node && node->covered_set >= 0) { asscpp_covered_free(oc_->context(), node->covered_set); node->covered_set = -1; }
  }

  // setMaxMinSensorAccuracy (coverage_path_planning_heuristic.cpp:399-403): needed before SurfaceCoveragewithAccuracyH
  void setMaxMinSensorAccuracy(std::vector<geometry_msgs::PoseArray> sensorsViewsSS) {
    oc_->SSMaxMinAccuracy(sensorsViewsSS);
    maxDepth = std::sqrt(oc_->maxAccuracyError / 0.0000285);
  }
  double maxDepth = 0;
  double accuracySum = 0;  // :497, :837

  ViewpointBatcher& batcher() { return batcher_; }

  // coverage_path_planning_heuristic.cpp:425-488 (SurfaceCoverageH) / :1227-1240 (root)
  void calculateHeuristic(NodeB200* node) {
    if (!node) return;
    AsscppCtx* c = oc_->context();
    // children of an expansion that went through batcher().prefetch() are scored from the cached group gains:
    // no launch here, and the child's covered set is only made if the node is expanded later (ensureSet)
    unsigned long long cachedGain = 0;
    const bool cached = node->parent && type_ != SurfaceCoveragewithAccuracyH && type_ != VolumetricCoverageH &&
                        node->covered_set < 0 && (!accumulate_ || node->parent->covered_set >= 0) &&
                        batcher_.groupGain(node, scoringSet(node->parent), &cachedGain);
    if (!cached) ensureSet(node);
    double a = 1.0;
    if ((type_ == SurfaceCoveragewithAccuracyH || type_ == VolumetricCoverageH) && node->parent) {
      // a = (maxAccuracyError - calcAvgAccuracy(visibleCloud)) / maxAccuracyError (:493-494, :836-838); the
      // one-argument overload averages 0.0000285 z^2 over the concatenated sensor clouds
      std::vector<double> p(7 * node->senPoses.size());
      for (size_t s = 0; s < node->senPoses.size(); ++s) OcclusionCulling::pack(node->senPoses[s], &p[7 * s]);
      oc_->check(asscpp_eval_batch(c, p.data(), (int64_t)node->senPoses.size(), -1, ASSCPP_EVAL_ACCURACY));
      std::vector<AsscppViewAccuracy> acc(node->senPoses.size());
      oc_->check(asscpp_get_accuracy(c, acc.data(), (int64_t)acc.size()));
      double sum = 0, cnt = 0;
      for (size_t s = 0; s < acc.size(); ++s) { sum += acc[s].err_sum_z; cnt += acc[s].visible; }
      const double avgAcc = sum / cnt;  // 0/0 = NaN for an empty cloud, as in the reference
      accuracySum += avgAcc;
      a = (oc_->maxAccuracyError - avgAcc) / oc_->maxAccuracyError;
    }
    float pct = 0;
    if (cached) {
      int64_t npts = 0, nvox = 0;
      if (accumulate_) oc_->check(asscpp_covered_popcount(c, node->parent->covered_set, &npts, &nvox));
      AsscppModelInfo info;
      asscpp_model_info(c, &info);
      pct = (float)(nvox + (int64_t)cachedGain) / (float)info.n_cov_voxels * 100.0f;  // occlusion_culling.cpp:431-432
    } else {
      oc_->check(asscpp_coverage_percent(c, node->covered_set, &pct));  // :482
    }
    node->coverage = pct;
    if (type_ == VolumetricCoverageH) {
      // :581-641 (moved), :911-971 (same position), root :1111-1126.  node->voxels = VoxelGrid(volumetricVoxelRes) of
      // the parent's voxels + the visible cloud: in covered-set form the voxel bits of the set, on the context's
      // coverage lattice (cov_leaf = volumetricVoxelRes = 0.5).  Counts and percentages are doubles here.
      AsscppModelInfo info;
      asscpp_model_info(c, &info);
      int64_t npts = 0, nvox = 0, pvox = 0;
      oc_->check(asscpp_covered_popcount(c, node->covered_set, &npts, &nvox));
      node->coverage = ((double)nvox / (double)info.n_cov_voxels) * 100;  // :629
      if (node->parent) {
        if (accumulate_) oc_->check(asscpp_covered_popcount(c, node->parent->covered_set, &npts, &pvox));
        const double dx = node->pose.position.x - node->parent->pose.position.x,
                     dy = node->pose.position.y - node->parent->pose.position.y,
                     dz = node->pose.position.z - node->parent->pose.position.z;
        const double d = std::sqrt(dx * dx + dy * dy + dz * dz);
        node->distance = node->parent->distance + d;
        node->h_value = 0; node->g_value = 0;
        const double extraVoxelsNum = (double)nvox - (double)pvox;             // :615
        const double normAngle = 1 - shortestAngle(node->parent->pose, node->pose) / (2 * M_PI);  // :584-588
        const double normDist = (2.5 - d) / 2.5;                               // :631, 2.5 = max connection radius
        node->f_value = d != 0.0 ? node->parent->f_value + ((accW * a) + (distW * normDist) + (angleW * normAngle)) * extraVoxelsNum  // :640
                                 : node->parent->f_value + ((accW * a) + (angleW * normAngle)) * extraVoxelsNum;                       // :970
      } else {
        node->distance = 0; node->h_value = 0; node->g_value = 0; node->f_value = 0;  // :1243-1247
      }
      return;
    }
    if (node->parent) {
      double dx = node->pose.position.x - node->parent->pose.position.x,
             dy = node->pose.position.y - node->parent->pose.position.y,
             dz = node->pose.position.z - node->parent->pose.position.z;
      double d = std::sqrt(dx * dx + dy * dy + dz * dz);  // Dist(), sspp/include/sspp/utils.h
      node->distance = node->parent->distance + d;         // :452
      double cgain = node->coverage - node->parent->coverage;  // :483
      node->h_value = 0; node->g_value = 0;
      double normAngle = 1.0;
      if (d == 0.0 && type_ != SurfaceCoverageH) {
        // turning angle between parent and node orientation (:785-790, :813-817)
        normAngle = 1 - shortestAngle(node->parent->pose, node->pose) / (2 * M_PI);
      }
      if (type_ == SurfaceCoveragewithAccuracyH)
        node->f_value = d != 0.0 ? node->parent->f_value + ((1 / d) * cgain * a)   // :512
                                 : node->parent->f_value + a * cgain * normAngle;  // :840
      else if (type_ == SurfaceCoveragewithOrientationH)
        node->f_value = d != 0.0 ? node->parent->f_value + (1 / d) * cgain         // :487 (shared with SurfaceCoverageH)
                                 : node->parent->f_value + normAngle * cgain;      // :807
      else
        node->f_value = d != 0.0 ? node->parent->f_value + (1 / d) * cgain  // :487
                                 : node->parent->f_value + cgain;            // :780
    } else {
      node->distance = 0; node->h_value = 0; node->g_value = 0;
      node->f_value = 0;  // root: coverage from its own views (:1227-1240), f and g zeroed (:1243-1247)
    }
  }

  // <tf Quaternion::angleShortestPath>
  static double shortestAngle(const geometry_msgs::Pose& pp, const geometry_msgs::Pose& pn) {
    const double l2p = pp.orientation.x * pp.orientation.x + pp.orientation.y * pp.orientation.y +
                       pp.orientation.z * pp.orientation.z + pp.orientation.w * pp.orientation.w;
    const double l2n = pn.orientation.x * pn.orientation.x + pn.orientation.y * pn.orientation.y +
                       pn.orientation.z * pn.orientation.z + pn.orientation.w * pn.orientation.w;
    const double sq = std::sqrt(l2n * l2p);
    const double dt = pp.orientation.x * pn.orientation.x + pp.orientation.y * pn.orientation.y +
                      pp.orientation.z * pn.orientation.z + pp.orientation.w * pn.orientation.w;
    return dt < 0 ? std::acos(-dt / sq) * 2.0 : std::acos(dt / sq) * 2.0;
  }
  double accW = 0.8, distW = 0.5, angleW = 0.2;  // heuristic.cpp:50-53

  // Node::cloud_filtered in covered-set form: the parent's set OR the node's own sensor views (:439-445, :473-480).
  // calculateHeuristic calls it for nodes that were not prefetched; a prefetched node gets its set here when it is
  // taken from the open list for expansion (sspp/src/astar.cpp:262), one launch per EXPANDED node instead of one per
  // child.  The parent chain must still be alive (it is: closed-list nodes are kept until the search ends).
  void ensureSet(NodeB200* node) {
    if (node->covered_set >= 0) return;
    AsscppCtx* c = oc_->context();
    if (node->parent && accumulate_) {
      ensureSet(node->parent);
      oc_->check(asscpp_covered_clone(c, node->parent->covered_set, &node->covered_set));  // :473 parent's cloud
    } else {
      oc_->check(asscpp_covered_new(c, &node->covered_set));
    }
    std::vector<double> p(7 * node->senPoses.size());
    for (size_t s = 0; s < node->senPoses.size(); ++s) OcclusionCulling::pack(node->senPoses[s], &p[7 * s]);
    // visibleCloud += extractVisibleSurface(senPoses[i]) for every sensor (:439-445); union into the set (:474-480)
    oc_->check(asscpp_covered_or_views(c, node->covered_set, p.data(), (int64_t)node->senPoses.size(), -1));
  }

  // coverage_path_planning_heuristic.cpp:253-283
  bool terminateConditionReached(const NodeB200* node) const {
    double delta = target_ - node->coverage;
    return delta <= tol_;
  }

  // acceptance rule of isFilteringConditionSatisfied (:339-357) for candidates whose geometric filters
  // (CGAL inside test / stand-off distance, :320-335) passed on the host: accept iff the sensors see more
  // than 5 points in total; each sensor cloud with more than 5 points joins the global covered set.
  // One batched launch for all candidates; accepted[i] is written per candidate.
  void filterCandidates(const std::vector<std::vector<geometry_msgs::Pose> >& sensorPoses, int global_set,
                        std::vector<char>& accepted) {
    std::vector<geometry_msgs::Pose> flat;
    for (size_t i = 0; i < sensorPoses.size(); ++i) flat.insert(flat.end(), sensorPoses[i].begin(), sensorPoses[i].end());
    std::vector<AsscppViewResult> res;
    std::vector<AsscppViewAccuracy> acc;
    oc_->evalBatch(flat, res, -1, 0, &acc);
    // :347-352: sensor clouds of more than 5 points whose average accuracy error reaches the threshold become
    // extra clusters of the next resampling round (their poses are kept; clusterBoxOfView gives the box)
    for (size_t k = 0; k < flat.size(); ++k)
      if (res[k].visible > 5 && acc[k].err_sum_depth / (double)acc[k].visible >= accuracyThreshhold)
        accuracyClusterViews.push_back(flat[k]);
    std::vector<double> p(7 * flat.size());
    for (size_t k = 0; k < flat.size(); ++k) OcclusionCulling::pack(flat[k], &p[7 * k]);
    oc_->check(asscpp_covered_or_views(oc_->context(), global_set, p.data(), (int64_t)flat.size(), 5));  // :344-346
    accepted.assign(sensorPoses.size(), 0);
    size_t k = 0;
    for (size_t i = 0; i < sensorPoses.size(); ++i) {
      unsigned long long total = 0;
      for (size_t s = 0; s < sensorPoses[i].size(); ++s) total += res[k++].visible;
      accepted[i] = total > 5;  // :357
    }
  }

  double accuracyThreshhold = 1e300;                      // heuristic.cpp accuracyThreshhold; off unless set
  std::vector<geometry_msgs::Pose> accuracyClusterViews;  // :350 accuracyClusters, as the views that produced them

  // clusteringPointCloud (:147-190) + findClusterBB (:218-246) of the uncovered part of the model, on the device
  void clusterUncovered(int global_set, std::vector<AsscppCluster>& clusters, double tolerance = 0.8, int minSize = 5,
                        int maxSize = 50000) {
    AsscppModelInfo info;
    asscpp_model_info(oc_->context(), &info);
    clusters.resize((size_t)std::max<int64_t>(info.n_cov_voxels, 1));
    int64_t n = 0;
    oc_->check(asscpp_uncovered_clusters(oc_->context(), global_set, tolerance, minSize, maxSize, nullptr,
                                         clusters.data(), (int64_t)clusters.size(), &n));
    clusters.resize((size_t)n);
  }
  // findClusterBB (:218-246) of one view's visible cloud (an accuracy cluster, pathplanner.cpp:783-786)
  void clusterBoxOfView(const geometry_msgs::Pose& view, double gridStart[3], double gridSize[3], float leaf = 0.5f) {
    std::vector<int32_t> idx;
    oc_->visibleIndices(view, idx);
    const float inv = 1.0f / leaf;
    int mn[3] = {INT_MAX, INT_MAX, INT_MAX}, mx[3] = {INT_MIN, INT_MIN, INT_MIN};
    for (size_t k = 0; k < idx.size(); ++k) {
      const pcl::PointXYZ& p = oc_->cloudCopy->points[(size_t)idx[k]];
      const float v[3] = {p.x, p.y, p.z};
      for (int a = 0; a < 3; ++a) {
        int c = (int)std::floor(v[a] * inv);
        mn[a] = std::min(mn[a], c); mx[a] = std::max(mx[a], c);
      }
    }
    float cmin[3], cmax[3];
    for (int a = 0; a < 3; ++a) {
      const float b_min = (float)mn[a] * leaf;
      cmin[a] = (b_min + leaf * 0.5f) + (float)0 * leaf;
      cmax[a] = (b_min + leaf * 0.5f) + (float)(mx[a] - mn[a]) * leaf;
    }
    gridSize[0] = std::abs(cmax[0] - cmin[0]) + 5.0; gridSize[1] = std::abs(cmax[1] - cmin[1]) + 5.0;
    gridSize[2] = std::abs(cmax[2] - cmin[2]) + 3.0;
    gridStart[0] = cmin[0] - 2.5; gridStart[1] = cmin[1] - 2.5; gridStart[2] = cmin[2];
  }

  // loadOBJFile (:1379-1437) hands vertices and faces here instead of building the CGAL AABB tree (:35-37)
  void setMesh(const std::vector<float>& vertices_xyz, const std::vector<int32_t>& triangles, int device = 0) {
    if (mesh_) asscpp_mesh_destroy(mesh_);
    mesh_ = nullptr;
    oc_->check(asscpp_mesh_create(device, vertices_xyz.data(), (int64_t)vertices_xyz.size() / 3, triangles.data(),
                                  (int64_t)triangles.size() / 3, &mesh_));
  }
  ~CoverageHeuristicB200() { if (mesh_) asscpp_mesh_destroy(mesh_); }

  // geometric part of isFilteringConditionSatisfied (:319-335) for a batch of robot positions: the sample is
  // outside the mesh (ray towards (100, 10, 56) crosses it an even number of times) and minDist..maxDist from it
  void filterPositions(const std::vector<geometry_msgs::Pose>& robotPoses, double minDist, double maxDist,
                       std::vector<char>& pass) {
    std::vector<double> p(3 * robotPoses.size());
    for (size_t i = 0; i < robotPoses.size(); ++i) {
      p[3 * i] = robotPoses[i].position.x; p[3 * i + 1] = robotPoses[i].position.y; p[3 * i + 2] = robotPoses[i].position.z;
    }
    const double through[3] = {100.0, 10.0, 56.0};  // :321
    std::vector<uint8_t> ok(robotPoses.size());
    if (asscpp_mesh_filter(mesh_, p.data(), (int64_t)robotPoses.size(), through, minDist, maxDist, nullptr, nullptr,
                           ok.data()) != ASSCPP_OK)
      throw std::runtime_error(std::string("asscpp_mesh_filter: ") + asscpp_mesh_last_error(mesh_));
    pass.assign(ok.begin(), ok.end());
  }

  // isConnectionConditionSatisfied (:300-315) for a batch of candidate edges a[i] -> b[i]
  void connectionsFree(const std::vector<geometry_msgs::Pose>& a, const std::vector<geometry_msgs::Pose>& b,
                       std::vector<char>& free_out) {
    std::vector<double> pa(3 * a.size()), pb(3 * a.size());
    for (size_t i = 0; i < a.size(); ++i) {
      pa[3 * i] = a[i].position.x; pa[3 * i + 1] = a[i].position.y; pa[3 * i + 2] = a[i].position.z;
      pb[3 * i] = b[i].position.x; pb[3 * i + 1] = b[i].position.y; pb[3 * i + 2] = b[i].position.z;
    }
    std::vector<uint8_t> ok(a.size());
    if (asscpp_mesh_segments_free(mesh_, pa.data(), pb.data(), (int64_t)a.size(), ok.data()) != ASSCPP_OK)
      throw std::runtime_error(std::string("asscpp_mesh_segments_free: ") + asscpp_mesh_last_error(mesh_));
    free_out.assign(ok.begin(), ok.end());
  }
  bool isConnectionConditionSatisfied(const geometry_msgs::Pose& a, const geometry_msgs::Pose& b) {
    std::vector<char> f;
    connectionsFree(std::vector<geometry_msgs::Pose>(1, a), std::vector<geometry_msgs::Pose>(1, b), f);
    return f[0] != 0;
  }

  // pointCloudDiff (:88-145): uncovered model voxels / model voxels * 100 (double arithmetic, :143)
  double uncoveredPercent(int global_set) {
    int64_t npts = 0, nvox = 0;
    oc_->check(asscpp_covered_popcount(oc_->context(), global_set, &npts, &nvox));
    AsscppModelInfo info;
    asscpp_model_info(oc_->context(), &info);
    return ((double)(info.n_cov_voxels - nvox) / (double)info.n_cov_voxels) * 100;
  }

 private:
  OcclusionCulling* oc_;
  ViewpointBatcher batcher_;
  double target_, tol_;
  int type_;
  bool accumulate_;
  AsscppMesh* mesh_ = nullptr;
};

}  // namespace SSPP
