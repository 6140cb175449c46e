// coverage_path_planning_heuristic_b200_impl2.h -- the virtuals (see coverage_path_planning_heuristic_b200.h).
#pragma once

namespace SSPP {

// coverage_path_planning_heuristic.cpp:425-1300
inline void CoveragePathPlanningHeuristic::calculateHeuristic(Node* node) {
  if (node == NULL) return;
  if (heuristicType == InfoGainVolumetricH) {
    if (!intended_info_gain_)
      throw std::runtime_error("CoveragePathPlanningHeuristic: the reference's InfoGainVolumetricH block copies "
                               "*node->parent->octree (coverage_path_planning_heuristic.cpp:649), a pointer "
                               "Node::Node(Node*) never sets (sspp/src/node.cpp:43-68): undefined behaviour, no result to "
                               "reproduce.  allowIntendedInfoGain() enables the evidently intended form (info_gain_b200.h)");
    calculateInfoGain(node);
    return;
  }
  const bool volumetric = heuristicType == VolumetricCoverageH;
  // the parent's accumulated cloud: cloud_filtered (:473), for VolumetricCoverageH its twin Node::voxels (:549)
  Node* parent = node->parent;
  int64_t parent_vox = 0;
  int covered = -1;
  Score sc;
  std::map<const Node*, Score>::iterator hit = cache_.find(node);
  if (hit != cache_.end() && cache_parent_ == parent) {
    sc = hit->second;
    cache_.erase(hit);
    // the set the batch was scored against is still in parent_set_ when the parent handed over a cloud
    if (parent && parent_set_valid_for_ == parent) { covered = parent_set_; parent_vox = parent_vox_cached_; }
  } else {
    if (parent) {
      if (volumetric) {
        // same accumulation through Node::voxels: borrow the cloud_filtered slot of a stand-in parent view
        Node tmp;
        tmp.cloud_filtered->points = parent->voxels.points;
        covered = parent_covered_set(&tmp, &parent_vox);
      } else {
        covered = parent_covered_set(parent, &parent_vox);
      }
    }
    parent_set_valid_for_ = covered >= 0 ? parent : NULL;
    parent_vox_cached_ = parent_vox;
    // the batcher: this child + its not yet scored siblings (Astar::childList), one launch
    std::vector<Node*> group(1, node);
    if (sibling_source_ && parent)
      for (Node* s = *sibling_source_; s != NULL; s = s->next)
        if (s != node && s->parent == parent) group.push_back(s);
    std::vector<Score> scores;
    score_nodes(group, covered, scores);
    cache_.clear();
    cache_parent_ = parent;
    for (size_t i = 1; i < group.size(); ++i) cache_[group[i]] = scores[i];
    sc = scores[0];
  }
  if (!sc.exact) sc = score_exact(node, covered, parent_vox);
  nodes_scored_ += 1;

  AsscppModelInfo info;
  asscpp_model_info(ctx(), &info);
  const int64_t node_vox = parent_vox + (int64_t)sc.gain_vox;
  // calcCoveragePercent: float counters, float ratio, * 100 (occlusion_culling.cpp:401,431-432)
  const float pct = ((float)node_vox / occlussionCulling->OriginalVoxelsSize) * 100;
  // a = (maxAccuracyError - calcAvgAccuracy(visibleCloud)) / maxAccuracyError (:493-494); 0/0 = NaN for an empty cloud
  double a = 1.0, avgAcc = 0.0;
  if (needs_accuracy() && parent) {
    avgAcc = sc.err_sum_z / (double)sc.visible;
    a = (occlussionCulling->maxAccuracyError - avgAcc) / occlussionCulling->maxAccuracyError;
    accuracyPerViewpointAvg.push_back(a);
    accuracySum += avgAcc;
  }
  if (fillCloudFiltered) {
    fill_cloud_filtered(node, covered);
    if (volumetric) node->voxels.points = node->cloud_filtered->points;
  }

  if (parent) {
    node->h_value = 0;
    node->g_value = 0;
    double f = 0, c = 0;
    const double d = Dist(node->pose.p, parent->pose.p);                      // :448
    node->distance = parent->distance + d;                                    // :450
    const double normAngle = 1 - shortest_angle(parent->pose.p, node->pose.p) / (2 * M_PI);
    if (volumetric) {
      // :529-580 (moved), :833-877 (same position): voxel counts and percentages in double
      const double extraVoxelsNum = (double)node_vox - (double)parent_vox;
      const double extraVoxelPercent = (extraVoxelsNum / (double)info.n_cov_voxels) * 100;
      extraCovPerViewpointAvg.push_back(extraVoxelPercent);
      extraCovSum += extraVoxelPercent;
      node->coverage = ((double)node_vox / (double)info.n_cov_voxels) * 100;
      const double normDist = (2.5 - d) / 2.5;
      f = d != 0.0 ? parent->f_value + ((accW * a) + (distW * normDist) + (angleW * normAngle)) * extraVoxelsNum
                   : parent->f_value + ((accW * a) + (angleW * normAngle)) * extraVoxelsNum;
    } else {
      node->coverage = pct;                                                   // :482
      c = node->coverage - parent->coverage;                                  // :483
      extraCovPerViewpointAvg.push_back(c);
      extraCovSum += c;
      if (d != 0.0) {
        if (heuristicType == SurfaceCoveragewithAccuracyH) f = parent->f_value + ((1 / d) * c * a);   // :512
        else f = parent->f_value + ((1 / d) * c);                                                     // :487
      } else {
        if (heuristicType == SurfaceCoverageH) f = parent->f_value + c;                               // :780
        else if (heuristicType == SurfaceCoveragewithOrientationH) f = parent->f_value + normAngle * c;  // :807
        else f = parent->f_value + a * c * normAngle;                                                 // :840
      }
    }
    node->f_value = f;                                                        // :1088
  } else {
    // root (:1096-1247): coverage of its own views; f and g zeroed
    if (volumetric) node->coverage = ((double)node_vox / (double)info.n_cov_voxels) * 100;  // :1120
    else node->coverage = pct;                                                              // :1236
    extraCovPerViewpointAvg.push_back(node->coverage);
    extraCovSum += node->coverage;
    node->f_value = 0;
    node->g_value = 0;
  }
}

// :253-283
inline bool CoveragePathPlanningHeuristic::terminateConditionReached(Node* node) {
  const double deltaCoverage = coverageTarget - node->coverage;
  selectedPointsNum++;
  if (debug)
    std::printf("Total Coverage %%: %g  f = %g  viewpoints %d  delta %g\n", node->coverage, node->f_value, selectedPointsNum,
                deltaCoverage);
  return deltaCoverage <= coverageTolerance;
}

// :285-298: the segment between the two positions crosses no triangle of the mesh
inline bool CoveragePathPlanningHeuristic::isConnectionConditionSatisfied(geometry_msgs::Pose temp, geometry_msgs::Pose S) {
  const double a[3] = {temp.position.x, temp.position.y, temp.position.z};
  const double b[3] = {S.position.x, S.position.y, S.position.z};
  uint8_t free_ = 0;
  if (asscpp_mesh_segments_free(mesh_, a, b, 1, &free_) != ASSCPP_OK)
    throw std::runtime_error(std::string("asscpp_mesh_segments_free: ") + asscpp_mesh_last_error(mesh_));
  return free_ != 0;
}

// :317-364
inline bool CoveragePathPlanningHeuristic::isFilteringConditionSatisfied(
    geometry_msgs::Pose pose, geometry_msgs::PoseArray& correspondingSensorPoses, double minDist, double maxDist,
    pcl::PointCloud<pcl::PointXYZ>& globalCloud, std::vector<pcl::PointCloud<pcl::PointXYZ> >& accuracyClusters,
    double accuracyThreshhold) {
  // model-node collision based filtering (ray parity towards (100, 10, 56), :321-328) + distance based filtering (:331-335)
  const double pos[3] = {pose.position.x, pose.position.y, pose.position.z};
  const double through[3] = {100.0, 10.0, 56.0};
  uint8_t pass = 0;
  if (asscpp_mesh_filter(mesh_, pos, 1, through, minDist, maxDist, NULL, NULL, &pass) != ASSCPP_OK)
    throw std::runtime_error(std::string("asscpp_mesh_filter: ") + asscpp_mesh_last_error(mesh_));
  if (!pass) return false;
  // coverage based filtering (:337-360): every sensor in one launch -- visible counts, the accuracy of each sensor's
  // cloud, and the union of the clouds of more than 5 points into the device covered set
  const size_t ns = correspondingSensorPoses.poses.size();
  if (ns == 0) return false;
  std::vector<double> p(7 * ns);
  for (size_t s = 0; s < ns; ++s) OcclusionCulling::pack(correspondingSensorPoses.poses[s], &p[7 * s]);
  AsscppEvalOpts o;
  std::memset(&o, 0, sizeof(o));
  o.covered_set = -1; o.accum_set = global_set_; o.accum_min_visible = 5; o.group_size = 1;
  o.flags = ASSCPP_EVAL_ACCURACY;
  check(asscpp_eval_ex(ctx(), p.data(), (int64_t)ns, &o));
  std::vector<AsscppViewResult> res(ns);
  std::vector<AsscppViewAccuracy> acc(ns);
  check(asscpp_get_results(ctx(), res.data(), (int64_t)ns));
  check(asscpp_get_accuracy(ctx(), acc.data(), (int64_t)ns));
  launches_scored_ += 1;
  unsigned long long total = 0;
  for (size_t s = 0; s < ns; ++s) {
    total += res[s].visible;
    if (res[s].visible <= 5) continue;
    // calcAvgAccuracy(pts, pose) (occlusion_culling.cpp:459-475)
    const double accuracy = acc[s].err_sum_depth / (double)acc[s].visible;
    const bool cluster = accuracy >= accuracyThreshhold;
    if (appendGlobalCloud || cluster) {
      pcl::PointCloud<pcl::PointXYZ> pts = occlussionCulling->extractVisibleSurface(correspondingSensorPoses.poses[s]);
      if (appendGlobalCloud) globalCloud += pts;   // :346
      if (cluster) accuracyClusters.push_back(pts);  // :350
    }
  }
  return total > 5;  // :357
}

// :88-145: uncovered voxels of the model (cell centres, k j i order) and their share in percent
inline double CoveragePathPlanningHeuristic::pointCloudDiff(pcl::PointCloud<pcl::PointXYZ>::Ptr globalCloudPtr,
                                                            pcl::PointCloud<pcl::PointXYZ>::Ptr& pointCloudDiffPtr) {
  int set = global_set_;
  if (appendGlobalCloud) {
    // the caller's cloud is the truth (it may have been edited): voxelise it as :92-96 does
    set = scratch_set_;
    check(asscpp_covered_clear(ctx(), set));
    const std::vector<pcl::PointXYZ>& pts = globalCloudPtr->points;
    std::vector<float> xyz(3 * pts.size());
    for (size_t i = 0; i < pts.size(); ++i) { xyz[3 * i] = pts[i].x; xyz[3 * i + 1] = pts[i].y; xyz[3 * i + 2] = pts[i].z; }
    check(asscpp_covered_or_cloud(ctx(), set, xyz.data(), (int64_t)pts.size()));
  }
  AsscppModelInfo info;
  asscpp_model_info(ctx(), &info);
  std::vector<float> ctr(3 * (size_t)std::max<int64_t>(info.n_cov_voxels, 1));
  int64_t n = 0;
  // with the reference's unchecked linear-index lookup into the covered cloud's grid (:120-123, quirk Q11)
  check(asscpp_uncovered_voxels_ex(ctx(), set, ASSCPP_DIFF_PCL_ALIASING, NULL, ctr.data(), info.n_cov_voxels, &n));
  pointCloudDiffPtr->points.resize((size_t)n);
  for (int64_t i = 0; i < n; ++i) pointCloudDiffPtr->points[(size_t)i] = pcl::PointXYZ(ctr[3 * i], ctr[3 * i + 1], ctr[3 * i + 2]);
  return ((double)n / (double)info.n_cov_voxels) * 100;  // :143
}

// :147-190: Euclidean clustering (tolerance 0.8, sizes 5..50000) of the cloud pointCloudDiff returned -- a cloud of
// cell centres, one per voxel; clusters largest first, members in the cloud's order
inline void CoveragePathPlanningHeuristic::clusteringPointCloud(
    std::vector<pcl::PointCloud<pcl::PointXYZ> >& clustersPointCloudVec, pcl::PointCloud<pcl::PointXYZ>::Ptr pointCloudDiffPtr) {
  const std::vector<pcl::PointXYZ>& pts = pointCloudDiffPtr->points;
  if (pts.empty()) return;
  std::vector<float> xyz(3 * pts.size());
  for (size_t i = 0; i < pts.size(); ++i) { xyz[3 * i] = pts[i].x; xyz[3 * i + 1] = pts[i].y; xyz[3 * i + 2] = pts[i].z; }
  check(asscpp_covered_clear(ctx(), scratch_set_));
  check(asscpp_covered_or_cloud(ctx(), scratch_set_, xyz.data(), (int64_t)pts.size()));
  check(asscpp_covered_invert_voxels(ctx(), scratch_set_));  // cluster the voxels OF the cloud
  AsscppModelInfo info;
  asscpp_model_info(ctx(), &info);
  const int64_t nc = info.n_cov_voxels;
  std::vector<int32_t> labels((size_t)std::max<int64_t>(nc, 1)), ids((size_t)std::max<int64_t>(nc, 1));
  std::vector<AsscppCluster> clusters((size_t)std::max<int64_t>(nc, 1));
  std::vector<float> ctr(3 * (size_t)std::max<int64_t>(nc, 1));
  int64_t ncl = 0, nvox = 0;
  check(asscpp_uncovered_clusters(ctx(), scratch_set_, 0.8, 5, 50000, labels.data(), clusters.data(), nc, &ncl));
  check(asscpp_uncovered_voxels(ctx(), scratch_set_, ids.data(), ctr.data(), nc, &nvox));
  const size_t base = clustersPointCloudVec.size();
  clustersPointCloudVec.resize(base + (size_t)ncl);
  for (int64_t i = 0; i < nvox; ++i) {
    const int32_t l = labels[(size_t)ids[(size_t)i]];
    if (l < 0) continue;
    clustersPointCloudVec[base + (size_t)l].push_back(pcl::PointXYZ(ctr[3 * i], ctr[3 * i + 1], ctr[3 * i + 2]));
  }
}

// :218-246: box of the cluster on the 0.5 m lattice, grown by 5 m in x, y and 3 m in z
inline void CoveragePathPlanningHeuristic::findClusterBB(pcl::PointCloud<pcl::PointXYZ> clusterPoints,
                                                         geometry_msgs::Vector3& gridSize, geometry_msgs::Pose& gridStart) {
  const float leaf = (float)voxelResForConn, inv = 1.0f / leaf;
  int mn[3] = {INT_MAX, INT_MAX, INT_MAX}, mx[3] = {INT_MIN, INT_MIN, INT_MIN};
  for (size_t k = 0; k < clusterPoints.points.size(); ++k) {
    const float v[3] = {clusterPoints.points[k].x, clusterPoints.points[k].y, clusterPoints.points[k].z};
    for (int a = 0; a < 3; ++a) {
      const int c = (int)std::floor(v[a] * inv);
      mn[a] = std::min(mn[a], c);
      mx[a] = std::max(mx[a], c);
    }
  }
  float cmin[3], cmax[3];
  for (int a = 0; a < 3; ++a) {
    // getCentroidCoordinate of the grid's own min / max cells (voxel_grid_occlusion_estimation.h:143-157)
    const float b_min = (float)mn[a] * leaf;
    cmin[a] = (b_min + leaf * 0.5f) + (float)0 * leaf;
    cmax[a] = (b_min + leaf * 0.5f) + (float)(mx[a] - mn[a]) * leaf;
  }
  const double maximizeSizeXY = 5, maximizeSizeZ = 3;
  gridSize.x = std::abs(cmax[0] - cmin[0]) + maximizeSizeXY;
  gridSize.y = std::abs(cmax[1] - cmin[1]) + maximizeSizeXY;
  gridSize.z = std::abs(cmax[2] - cmin[2]) + maximizeSizeZ;
  gridStart.position.x = cmin[0] - double(maximizeSizeXY / 2);
  gridStart.position.y = cmin[1] - double(maximizeSizeXY / 2);
  gridStart.position.z = cmin[2];
}

}  // namespace SSPP
