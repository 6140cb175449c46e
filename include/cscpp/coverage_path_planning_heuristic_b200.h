// coverage_path_planning_heuristic_b200.h -- drop-in for the reference's plugin class
//     class SSPP::CoveragePathPlanningHeuristic : public SSPP::Heuristic
// (cscpp/include/cscpp/coverage_path_planning_heuristic.h:77-133, cscpp/src/coverage_path_planning_heuristic.cpp), same
// class name, same constructor, the ten virtuals of sspp/include/sspp/heuristic_interface.h:35-56 at the reference's
// signatures, on top of the C ABI (include/asscpp_gpu.h).  SSPP::Astar (sspp/src/astar.cpp:131,275,318) and
// SSPP::PathPlanner (sspp/src/pathplanner.cpp:160,549,595,687,771,781,804) call it unchanged; tests/cpp/ref_sspp_driver.cpp
// links the reference's own astar.cpp / pathplanner.cpp objects against this class and runs them on the GPU.
//
// What is behind each virtual:
//   calculateHeuristic            one asscpp_eval_ex launch scores the node -- or, with batchSiblingsFrom(), every child
//                                 of the expansion the node belongs to (the batcher: scores cached per Node*)
//   isFilteringConditionSatisfied asscpp_mesh_filter (CGAL inside / distance tests) + one launch for the candidate's
//                                 sensors (visible counts, accuracy, union into the device covered set)
//   isConnectionConditionSatisfied asscpp_mesh_segments_free
//   pointCloudDiff / clusteringPointCloud / findClusterBB   asscpp_covered_or_cloud, asscpp_uncovered_voxels,
//                                 asscpp_uncovered_clusters
//   displayProgress, findClusterOuterPoints   RViz output / never called by the planners: unavailable on this box
// Heuristic types: SurfaceCoverageH, SurfaceCoveragewithOrientationH, SurfaceCoveragewithAccuracyH, VolumetricCoverageH.
// InfoGainVolumetricH: see info_gain_b200.h (the reference's own block dereferences a pointer Node::Node(Node*) never
// sets, sspp/src/node.cpp:43-68, so only its evidently intended form can be offered; opt-in, allowIntendedInfoGain()).
//
// No CPU fallback: construction throws when no sm_100 GPU is usable.
#pragma once
#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include <ros/ros.h>
#include <geometry_msgs/Pose.h>
#include <geometry_msgs/PoseArray.h>
#include <geometry_msgs/Vector3.h>
#include <sspp/heuristic_interface.h>
#include <sspp/node.h>

#include "occlusion_culling_b200.h"

#ifndef ASSCPP_HEURISTIC_TYPES_DEFINED
#define ASSCPP_HEURISTIC_TYPES_DEFINED
// cscpp/include/cscpp/coverage_path_planning_heuristic.h:72
enum { SurfaceCoverageH, SurfaceCoveragewithOrientationH, SurfaceCoveragewithAccuracyH, SurfaceAreaCoverageH,
       VolumetricCoverageH, InfoGainVolumetricH };
#endif

namespace SSPP {

class CoveragePathPlanningHeuristic : public Heuristic {
 public:
  // coverage_path_planning_heuristic.cpp:27-86.  collisionCheckModelP: OBJ file, occlusionCullingModelN: PCD file
  CoveragePathPlanningHeuristic(ros::NodeHandle& nh, std::string collisionCheckModelP, std::string occlusionCullingModelN,
                                bool d, bool gradualV, int hType)
      : debug(d), gradualVisualization(gradualV), heuristicType(hType) {
    init_defaults();
    std::vector<float> vertices;
    std::vector<int32_t> triangles;
    loadOBJFile(collisionCheckModelP.c_str(), vertices, triangles);
    occlussionCulling = new OcclusionCulling(nh, occlusionCullingModelN);
    occlussionCulling->accumulateOcclusionFreeCloud = false;
    set_mesh(vertices, triangles);
    finish_init();
  }
  // non-breaking addition: model cloud and mesh already in memory (vertices n x 3 floats, triangles 0-based)
  CoveragePathPlanningHeuristic(ros::NodeHandle& nh, pcl::PointCloud<pcl::PointXYZ>::Ptr& cloudPtr,
                                const std::vector<float>& vertices, const std::vector<int32_t>& triangles, bool d,
                                bool gradualV, int hType)
      : debug(d), gradualVisualization(gradualV), heuristicType(hType) {
    init_defaults();
    occlussionCulling = new OcclusionCulling(nh, cloudPtr);
    occlussionCulling->accumulateOcclusionFreeCloud = false;
    set_mesh(vertices, triangles);
    finish_init();
  }
  ~CoveragePathPlanningHeuristic() {
    if (mesh_) asscpp_mesh_destroy(mesh_);
    delete occlussionCulling;  // coverage_path_planning_heuristic.cpp:248-251
  }

  bool isCost() { return heuristicType == InfoGainVolumetricH; }                       // :387-392
  void setCoverageTarget(double t) { coverageTarget = t; }                            // :404-407
  void setCoverageTolerance(double t) { coverageTolerance = t; }                      // :409-412
  void setDebug(bool d) { debug = d; }                                                // :414-417
  int getHeuristicType() { return heuristicType; }                                    // :419-422
  void setMaxMinSensorAccuracy(std::vector<geometry_msgs::PoseArray> sensorsViewsSS) {  // :399-403
    occlussionCulling->SSMaxMinAccuracy(sensorsViewsSS);
    maxDepth = std::sqrt(occlussionCulling->maxAccuracyError / 0.0000285);
  }

  // ---- non-breaking additions -------------------------------------------------------------------------------------
  // The viewpoint batcher.  Astar keeps the not yet scored siblings of the child it hands to calculateHeuristic in its
  // public member childList (sspp/include/sspp/astar.h:72; the loop at astar.cpp:311-318 unlinks one child at a time).
  // Given the address of that member -- one added line in main, e.g. heuristic.batchSiblingsFrom(&pathPlanner->childList)
  // -- the first call of an expansion scores ALL its children in one launch and the following calls read the cache.
  // Without it every call is its own launch; the results are the same.
  void batchSiblingsFrom(Node* const* childListAddr) { sibling_source_ = childListAddr; }
  // Node::cloud_filtered of a scored node is left empty by default: in the reference's A* nobody reads it (the parent a
  // child sees is the copy made at astar.cpp:262, whose cloud is empty, node.cpp:67).  true = fill it with one point
  // (the cell centre) per covered coverage voxel, which is what calcCoveragePercent and a child's accumulation see of it.
  bool fillCloudFiltered;
  // isFilteringConditionSatisfied appends the visible points of accepted sensors to the caller's globalCloud as the
  // reference does (:344-346).  false = keep the union only as a covered set on the device (pointCloudDiff then ignores
  // the cloud it is handed and uses that set): no point traffic at all.
  bool appendGlobalCloud;
  // InfoGainVolumetricH in its evidently intended form (info_gain_b200.h): a child inherits the tree of the node it
  // was expanded from.  Off by default: calculateHeuristic then refuses the type and says why.
  void allowIntendedInfoGain(bool on = true) { intended_info_gain_ = on; }
  OcclusionCulling* occlusionCulling() { return occlussionCulling; }
  AsscppMesh* mesh() { return mesh_; }
  int globalCoveredSet() const { return global_set_; }
  long launchesScored() const { return launches_scored_; }
  long nodesScored() const { return nodes_scored_; }

  // statistics the reference prints (coverage_path_planning_heuristic.cpp:259-264)
  std::vector<double> accuracyPerViewpointAvg, extraCovPerViewpointAvg;
  double accuracySum, extraCovSum;
  int selectedPointsNum;

  void calculateHeuristic(Node* node);
  bool terminateConditionReached(Node* node);
  bool isConnectionConditionSatisfied(SearchSpaceNode* temp, SearchSpaceNode* S) {     // :300-315
    return isConnectionConditionSatisfied(temp->location, S->location);
  }
  bool isConnectionConditionSatisfied(geometry_msgs::Pose temp, geometry_msgs::Pose S);  // :285-298
  bool isFilteringConditionSatisfied(geometry_msgs::Pose pose, geometry_msgs::PoseArray& correspondingSensorPoses,
                                     double minDist, double maxDist, pcl::PointCloud<pcl::PointXYZ>& globalCloud,
                                     std::vector<pcl::PointCloud<pcl::PointXYZ> >& accuracyClusters,
                                     double accuracyThreshhold);
  void displayProgress(std::vector<Tree>) {}  // RViz search-tree markers (:366-386): unavailable on this box
  double pointCloudDiff(pcl::PointCloud<pcl::PointXYZ>::Ptr globalCloudPtr,
                        pcl::PointCloud<pcl::PointXYZ>::Ptr& pointCloudDiffPtr);
  void clusteringPointCloud(std::vector<pcl::PointCloud<pcl::PointXYZ> >& clustersPointCloudVec,
                            pcl::PointCloud<pcl::PointXYZ>::Ptr pointCloudDiffPtr);
  // :192-216: concave hull of the waypoints for an RViz marker; its call sites in the planner are commented out
  // (pathplanner.cpp:746,817).  The hull offered here is the waypoint positions themselves.
  void findClusterOuterPoints(geometry_msgs::PoseArray waypoints, pcl::PointCloud<pcl::PointXYZ>& cloudHull) {
    for (size_t i = 0; i < waypoints.poses.size(); ++i)
      cloudHull.push_back(pcl::PointXYZ((float)waypoints.poses[i].position.x, (float)waypoints.poses[i].position.y,
                                        (float)waypoints.poses[i].position.z));
  }
  void findClusterBB(pcl::PointCloud<pcl::PointXYZ> clusterPoints, geometry_msgs::Vector3& gridSize,
                     geometry_msgs::Pose& gridStart);

  // loadOBJFile (:1355-1451) with its exact token rules, including its quirk for polygons without normals / texture
  // coordinates: every fan triangle of such a face repeats the face's FIRST three vertices (:1402-1412)
  static void loadOBJFile(const char* filename, std::vector<float>& vertices, std::vector<int32_t>& triangles);

 private:
  struct Score {            // what one scored node needs: sums over its sensors
    unsigned long long gain_vox;   // coverage voxels the node adds to its parent's set (union over the sensors)
    unsigned long long visible;    // points of visibleCloud
    double err_sum_z;              // sum of 0.0000285 z^2 over visibleCloud (calcAvgAccuracy one-argument form, :445-458)
    bool exact;                    // false: a view took the ordered replay (quirk Q5), group sum not union-exact
  };
  struct InfoTree {                    // InfoGainVolumetricH: a node's octree as a flat array
    std::vector<float> logodds;        // per compact coverage-voxel id; valid where known
    std::vector<unsigned char> known;  // the cell has a leaf in the tree (octomap search() != NULL)
  };
  std::map<std::pair<int, const Node*>, InfoTree> info_trees_;  // by (node id, node->parent): survives astar.cpp:262's copy
  struct InfoViews {                   // what one launch returns for the sensors of one node
    std::vector<uint32_t> rows;        // per sensor: the coverage voxels its visible cloud touches
    std::vector<AsscppViewResult> res;
    std::vector<AsscppViewAccuracy> acc;
  };
  std::map<const Node*, InfoViews> info_cache_;  // siblings scored by the launch of the expansion's first child
  const Node* info_cache_parent_;
  std::vector<float> info_centres_;
  int64_t info_cells_;
  double modelVolume, modelTotalEntroby;
  bool intended_info_gain_;
  void info_gain_init();
  void calculateInfoGain(Node* node);
  static float oct_logodds(double p);
  static double oct_probability(double l);
  OcclusionCulling* occlussionCulling;
  AsscppMesh* mesh_;
  bool debug, gradualVisualization;
  int heuristicType;
  double coverageTarget, coverageTolerance, volumetricVoxelRes, voxelResForConn, maxConnRadius, maxDepth;
  double accW, distW, covW, angleW;
  Node* const* sibling_source_;
  std::map<const Node*, Score> cache_;
  const Node* cache_parent_;
  const Node* parent_set_valid_for_;  // the parent whose cloud parent_set_ currently holds
  int64_t parent_vox_cached_;
  int parent_set_, global_set_, scratch_set_;
  long launches_scored_, nodes_scored_;

  void init_defaults();
  void finish_init();
  void set_mesh(const std::vector<float>& vertices, const std::vector<int32_t>& triangles);
  void check(int rc) const { occlussionCulling->check(rc); }
  AsscppCtx* ctx() const { return occlussionCulling->context(); }
  bool needs_accuracy() const { return heuristicType == SurfaceCoveragewithAccuracyH || heuristicType == VolumetricCoverageH; }
  int parent_covered_set(const Node* parent, int64_t* parent_vox);
  void score_nodes(const std::vector<Node*>& nodes, int covered_set, std::vector<Score>& out);
  Score score_exact(Node* node, int covered_set, int64_t parent_vox);
  void fill_cloud_filtered(Node* node, int covered_set);
  static double shortest_angle(const geometry_msgs::Pose& a, const geometry_msgs::Pose& b);
};

}  // namespace SSPP

#include "coverage_path_planning_heuristic_b200_impl.h"
#include "info_gain_b200.h"
