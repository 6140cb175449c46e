// ref_facade.cpp -- C entry points over the REFERENCE's own sources, compiled where they lie under /root/reference
// by oracle/Makefile into oracle/_ref/libasscpp_ref.so (TEST INFRASTRUCTURE; never loaded by the product).
//
// What is the reference's own code here: OcclusionCulling, pcl::FrustumCullingTT, pcl::VoxelGridOcclusionEstimationT,
// SSPP::CoveragePathPlanningHeuristic, SSPP::PathPlanner / Astar / LList / SearchSpace / Node, Sensors -- unmodified
// .cpp files (cscpp/src, sspp/src).  What is NOT: their third-party dependencies (PCL, Eigen, tf, ROS, CGAL, octomap),
// which are absent from this image; they are replaced by the stand-in headers under compat/, restated from the
// published behaviour of those libraries (see the header comments there).  So this library pins the repo's oracle to
// the reference's control flow, loop order, quirks and expression structure; the float32 reduction orders inside
// Eigen's dot / norm and PCL's VoxelGrid binning remain stand-in choices (switchable, see compat/Eigen/Core).
//
// This file only adapts types: it calls the reference's public methods and turns point clouds back into index lists.
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include <cscpp/coverage_path_planning_heuristic.h>
#include <cscpp/occlusion_culling.h>
#include <sspp/pathplanner.h>
#include <sspp/sensors.h>

namespace {

ros::NodeHandle g_nh;

geometry_msgs::Pose to_pose(const double* p) {
  geometry_msgs::Pose q;
  q.position.x = p[0]; q.position.y = p[1]; q.position.z = p[2];
  q.orientation.x = p[3]; q.orientation.y = p[4]; q.orientation.z = p[5]; q.orientation.w = p[6];
  return q;
}

// `sub` is a subsequence of `all` in the same order (the reference copies points in ascending index order,
// frustum_culling.cpp:150-166, occlusion_culling.cpp:317-363): recover the indices by a merge walk on the bit
// patterns.  Points with identical coordinates share their voxel and their plane tests, so they are kept or dropped
// together and the greedy match is exact.
int64_t match_indices(const std::vector<pcl::PointXYZ>& all, const std::vector<pcl::PointXYZ>& sub, int32_t* out,
                      int64_t cap) {
  size_t i = 0;
  int64_t k = 0;
  uint32_t t1 = 0, t2 = 0;
  if (all.size() > 2) { std::memcpy(&t1, &all[1].data[3], 4); std::memcpy(&t2, &all[2].data[3], 4); }
  if (t1 == 1 && t2 == 2) {  // index-tagged cloud (ref_oc_create_points)
    for (size_t j = 0; j < sub.size(); ++j) {
      uint32_t t;
      std::memcpy(&t, &sub[j].data[3], 4);
      if (t >= all.size() || std::memcmp(all[t].data, sub[j].data, 12) != 0) return -1;
      if (k < cap && out) out[k] = (int32_t)t;
      ++k;
    }
    return k;
  }
  for (size_t j = 0; j < sub.size(); ++j) {
    while (i < all.size() && std::memcmp(all[i].data, sub[j].data, 12) != 0) ++i;
    if (i == all.size()) return -1;
    if (k < cap && out) out[k] = (int32_t)i;
    ++k;
    ++i;
  }
  return k;
}

uint64_t mix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

}  // namespace

extern "C" {

// OcclusionCulling(ros::NodeHandle&, std::string) -- occlusion_culling.cpp:26-79 (full path of the PCD file)
void* ref_oc_create_pcd(const char* pcd_path) { return new OcclusionCulling(g_nh, std::string(pcd_path)); }

// OcclusionCulling(ros::NodeHandle&, PointCloud::Ptr&) -- occlusion_culling.cpp:81-135; leaf > 0 overrides the
// public member voxelRes afterwards (the per-view grid of extractVisibleSurface reads it, :305)
void* ref_oc_create_points(const float* xyz, int64_t n, float leaf) {
  pcl::PointCloud<pcl::PointXYZ>::Ptr c(new pcl::PointCloud<pcl::PointXYZ>);
  c->points.resize((size_t)n);
  for (int64_t i = 0; i < n; ++i) {
    c->points[(size_t)i] = pcl::PointXYZ(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
    // the padding float of PointXYZ carries the point's index (bit pattern): the reference copies whole points and
    // computes with x, y, z only, so every cloud it returns can be read back as an index list without a search
    const uint32_t tag = (uint32_t)i;
    std::memcpy(&c->points[(size_t)i].data[3], &tag, 4);
  }
  OcclusionCulling* oc = new OcclusionCulling(g_nh, c);
  if (leaf > 0) oc->voxelRes = leaf;
  return oc;
}
void ref_oc_destroy(void* h) { delete static_cast<OcclusionCulling*>(h); }
int64_t ref_oc_num_points(void* h) { return (int64_t)static_cast<OcclusionCulling*>(h)->cloud->points.size(); }
float ref_oc_original_voxels(void* h) { return static_cast<OcclusionCulling*>(h)->OriginalVoxelsSize; }
void ref_oc_set_origin(void* h, const float o[3]) {  // cloud->sensor_origin_ (PCD VIEWPOINT), read at :338-341
  static_cast<OcclusionCulling*>(h)->cloud->sensor_origin_ = Eigen::Vector4f(o[0], o[1], o[2], 0.0f);
}

// extractVisibleSurface (occlusion_culling.cpp:265-373).  vis_idx / surv_idx: ascending original indices of the
// returned cloud / of FrustumCloud (either may be NULL); counts[0] = survivors, counts[1] = visible.
int ref_oc_extract(void* h, const double pose[7], int32_t* surv_idx, int32_t* vis_idx, int64_t cap, int64_t counts[2]) {
  OcclusionCulling* oc = static_cast<OcclusionCulling*>(h);
  pcl::PointCloud<pcl::PointXYZ> vis = oc->extractVisibleSurface(to_pose(pose));
  oc->occlusionFreeCloud->points.clear();  // grows for ever in the reference (:363); nothing reads it here
  counts[0] = match_indices(oc->cloudCopy->points, oc->FrustumCloud->points, surv_idx, cap);
  counts[1] = match_indices(oc->cloudCopy->points, vis.points, vis_idx, cap);
  return (counts[0] < 0 || counts[1] < 0) ? -1 : 0;
}

// calcCoveragePercent (occlusion_culling.cpp:375-444) of the model points with the given indices
float ref_oc_coverage_percent_idx(void* h, const int32_t* idx, int64_t n) {
  OcclusionCulling* oc = static_cast<OcclusionCulling*>(h);
  pcl::PointCloud<pcl::PointXYZ>::Ptr c(new pcl::PointCloud<pcl::PointXYZ>);
  for (int64_t i = 0; i < n; ++i) c->points.push_back(oc->cloudCopy->points[(size_t)idx[i]]);
  return oc->calcCoveragePercent(c);
}

// calcAvgAccuracy, both overloads (occlusion_culling.cpp:445-475), of the model points with the given indices
void ref_oc_avg_accuracy_idx(void* h, const double pose[7], const int32_t* idx, int64_t n, double out[2]) {
  OcclusionCulling* oc = static_cast<OcclusionCulling*>(h);
  pcl::PointCloud<pcl::PointXYZ> c;
  for (int64_t i = 0; i < n; ++i) c.points.push_back(oc->cloudCopy->points[(size_t)idx[i]]);
  out[0] = oc->calcAvgAccuracy(c, to_pose(pose));
  out[1] = oc->calcAvgAccuracy(c);
}

// pointCloudViewportTransform (occlusion_culling.cpp:538-618)
void ref_oc_viewport_transform(void* h, const double pose[7], const float* xyz, int64_t n, float* out) {
  OcclusionCulling* oc = static_cast<OcclusionCulling*>(h);
  pcl::PointCloud<pcl::PointXYZ> c;
  for (int64_t i = 0; i < n; ++i) c.points.push_back(pcl::PointXYZ(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]));
  pcl::PointCloud<pcl::PointXYZ> t = oc->pointCloudViewportTransform(c, to_pose(pose));
  for (int64_t i = 0; i < n; ++i) { out[3 * i] = t.points[(size_t)i].x; out[3 * i + 1] = t.points[(size_t)i].y; out[3 * i + 2] = t.points[(size_t)i].z; }
}

// Many viewpoints, one OcclusionCulling instance per host thread (the class is not re-entrant: member state at
// occlusion_culling.cpp:290,297,363,371).  stats: nviews x {survivors, visible, hash of the visible index set}.
int ref_oc_eval_batch(void** handles, int nthreads, const double* poses, int64_t nviews, int64_t* stats) {
  std::vector<std::thread> th;
  std::vector<int> rc((size_t)nthreads, 0);
  for (int t = 0; t < nthreads; ++t)
    th.push_back(std::thread([&, t]() {
      OcclusionCulling* oc = static_cast<OcclusionCulling*>(handles[t]);
      std::vector<int32_t> idx(oc->cloudCopy->points.size());
      for (int64_t v = t; v < nviews; v += nthreads) {
        int64_t counts[2];
        if (ref_oc_extract(oc, poses + 7 * v, nullptr, idx.data(), (int64_t)idx.size(), counts)) { rc[(size_t)t] = -1; return; }
        uint64_t hsh = 0;
        for (int64_t k = 0; k < counts[1]; ++k) hsh += mix64((uint64_t)(uint32_t)idx[(size_t)k]);
        stats[3 * v] = counts[0]; stats[3 * v + 1] = counts[1]; stats[3 * v + 2] = (int64_t)hsh;
      }
    }));
  for (size_t t = 0; t < th.size(); ++t) th[t].join();
  for (int t = 0; t < nthreads; ++t) if (rc[(size_t)t]) return -1;
  return 0;
}

// One instance per host thread, constructed concurrently (each constructor voxelises the whole model,
// occlusion_culling.cpp:101-131).  handles: count pointers out.
void ref_oc_create_many(const float* xyz, int64_t n, float leaf, int count, void** handles) {
  std::vector<std::thread> th;
  for (int t = 0; t < count; ++t) th.push_back(std::thread([&, t]() { handles[t] = ref_oc_create_points(xyz, n, leaf); }));
  for (size_t t = 0; t < th.size(); ++t) th[t].join();
}

// ref_oc_eval_batch + the gain of every view against a running covered set in bitset form (points by original index,
// coverage voxels by compact id; pt_cov = coverage-voxel id of every model point): the popcount form of
// heuristic.cpp:473-483, cheaper than the reference's own re-voxelisation of the accumulated cloud per candidate.
// stats: nviews x {survivors, visible, hash, gain_pts, gain_vox}.
int ref_oc_eval_batch_gain(void** handles, int nthreads, const double* poses, int64_t nviews, const uint64_t* cov_pts,
                           const uint64_t* cov_vox, const int32_t* pt_cov, int64_t n_cov, int64_t* stats) {
  std::vector<std::thread> th;
  std::vector<int> rc((size_t)nthreads, 0);
  for (int t = 0; t < nthreads; ++t)
    th.push_back(std::thread([&, t]() {
      OcclusionCulling* oc = static_cast<OcclusionCulling*>(handles[t]);
      std::vector<int32_t> idx(oc->cloudCopy->points.size());
      std::vector<uint64_t> seen((size_t)(n_cov + 63) / 64);
      for (int64_t v = t; v < nviews; v += nthreads) {
        int64_t counts[2];
        if (ref_oc_extract(oc, poses + 7 * v, nullptr, idx.data(), (int64_t)idx.size(), counts)) { rc[(size_t)t] = -1; return; }
        uint64_t hsh = 0;
        int64_t gp = 0, gv = 0;
        std::fill(seen.begin(), seen.end(), 0);
        for (int64_t k = 0; k < counts[1]; ++k) {
          const uint32_t i = (uint32_t)idx[(size_t)k];
          hsh += mix64((uint64_t)i);
          if (!cov_pts || !((cov_pts[i >> 6] >> (i & 63)) & 1)) ++gp;
          const uint32_t c = (uint32_t)pt_cov[i];
          if (!((seen[c >> 6] >> (c & 63)) & 1)) {
            seen[c >> 6] |= 1ull << (c & 63);
            if (!cov_vox || !((cov_vox[c >> 6] >> (c & 63)) & 1)) ++gv;
          }
        }
        int64_t* o = stats + 5 * v;
        o[0] = counts[0]; o[1] = counts[1]; o[2] = (int64_t)hsh; o[3] = gp; o[4] = gv;
      }
    }));
  for (size_t t = 0; t < th.size(); ++t) th[t].join();
  for (int t = 0; t < nthreads; ++t) if (rc[(size_t)t]) return -1;
  return 0;
}

// accept a viewpoint: OR its visible set into the covered bitsets (heuristic.cpp:344-346 / :473-480 in bitset form)
int ref_oc_accumulate(void* h, const double pose[7], uint64_t* cov_pts, uint64_t* cov_vox, const int32_t* pt_cov) {
  OcclusionCulling* oc = static_cast<OcclusionCulling*>(h);
  std::vector<int32_t> idx(oc->cloudCopy->points.size());
  int64_t counts[2];
  if (ref_oc_extract(oc, pose, nullptr, idx.data(), (int64_t)idx.size(), counts)) return -1;
  for (int64_t k = 0; k < counts[1]; ++k) {
    const uint32_t i = (uint32_t)idx[(size_t)k], c = (uint32_t)pt_cov[i];
    cov_pts[i >> 6] |= 1ull << (i & 63);
    cov_vox[c >> 6] |= 1ull << (c & 63);
  }
  return 0;
}

// Sensors::robot2sensorTransformation (sspp/src/sensors.cpp:47-90)
void ref_robot2sensor(const double robot[7], const double off[3], const double rpy[3], double out[7]) {
  Sensors s(58, 45, 0.255, 0.7, 6.0, 640, 480, Eigen::Vector3f((float)off[0], (float)off[1], (float)off[2]),
            Eigen::Vector3f((float)rpy[0], (float)rpy[1], (float)rpy[2]));
  geometry_msgs::Pose p = s.robot2sensorTransformation(to_pose(robot));
  out[0] = p.position.x; out[1] = p.position.y; out[2] = p.position.z;
  out[3] = p.orientation.x; out[4] = p.orientation.y; out[5] = p.orientation.z; out[6] = p.orientation.w;
}

// ---- the reference's planner, driven as its demo does (cscpp/src/coverage_heuristic_test.cpp:73-137, 153-181) ------
struct RefPlanner {
  SSPP::CoveragePathPlanningHeuristic* heuristic;
  SSPP::PathPlanner* planner;
  Robot* robot;
};

// heuristic_type: the enum of coverage_path_planning_heuristic.h:72
void* ref_planner_create(const char* obj_path, const char* pcd_path, int heuristic_type, double target, double tolerance,
                         double conn_radius) {
  RefPlanner* r = new RefPlanner();
  geometry_msgs::Point centre;
  centre.x = -0.3f; centre.y = 0.0f;
  r->robot = new Robot("Robot", 0.9, 0.5, 0.987, centre);                     // :85-91
  std::vector<Sensors> sensors;
  sensors.push_back(Sensors(58, 45, 0.255, 0.7, 6.0, 640, 480, Eigen::Vector3f(0, 0.022f, 0.065f), Eigen::Vector3f(0, -0.349f, 0)));  // :94
  sensors.push_back(Sensors(58, 45, 0.255, 0.7, 6.0, 640, 480, Eigen::Vector3f(0, 0.022f, -0.065f), Eigen::Vector3f(0, 0.349f, 0)));  // :95
  r->planner = new SSPP::PathPlanner(g_nh, r->robot, conn_radius, 0, sensors);  // :102 (progress display off)
  r->planner->setDebugDelay(0.0);
  r->heuristic = new SSPP::CoveragePathPlanningHeuristic(g_nh, obj_path, pcd_path, false, false, heuristic_type);  // :111
  r->heuristic->setCoverageTarget(target);
  r->heuristic->setCoverageTolerance(tolerance);
  r->planner->setHeuristicFucntion(r->heuristic);
  return r;
}
void ref_planner_destroy(void* h) {
  RefPlanner* r = static_cast<RefPlanner*>(h);
  delete r->planner; delete r->heuristic; delete r->robot; delete r;
}

// PathPlanner::dynamicNodesGenerationAndConnection (sspp/src/pathplanner.cpp:729-850), demo arguments at
// coverage_heuristic_test.cpp:75-80,137.  max_iterations < 0: run to the end.
void ref_planner_dynamic_sampling(void* h, const double grid_start[3], const double grid_size[3], double start_res,
                                  double res_decrement) {
  RefPlanner* r = static_cast<RefPlanner*>(h);
  geometry_msgs::Pose gs;
  gs.position.x = grid_start[0]; gs.position.y = grid_start[1]; gs.position.z = grid_start[2];
  geometry_msgs::Vector3 sz;
  sz.x = grid_size[0]; sz.y = grid_size[1]; sz.z = grid_size[2];
  r->planner->dynamicNodesGenerationAndConnection(gs, sz, start_res, res_decrement);
}

// One filtered grid (pathplanner.cpp:106-201 as called at :736): candidates -> isFilteringConditionSatisfied
void ref_planner_regular_grid(void* h, const double grid_start[3], const double grid_size[3], double res) {
  RefPlanner* r = static_cast<RefPlanner*>(h);
  geometry_msgs::Pose gs;
  gs.position.x = grid_start[0]; gs.position.y = grid_start[1]; gs.position.z = grid_start[2];
  geometry_msgs::Vector3 sz;
  sz.x = grid_size[0]; sz.y = grid_size[1]; sz.z = grid_size[2];
  r->planner->generateRegularGrid(gs, sz, (float)res, true, 45, true, true);
}

// load a search space from the three pose files (pathplanner.cpp:204-269) and connect it (:529-569)
void ref_planner_load_grid(void* h, const char* robot_file, const char* sensor0_file, const char* sensor1_file) {
  RefPlanner* r = static_cast<RefPlanner*>(h);
  r->planner->loadRegularGrid(robot_file, sensor0_file, sensor1_file);
  r->planner->connectNodes();
}

// the search space as the demo prints it (coverage_heuristic_test.cpp:153-181): rows {x y z yaw} + the robot and
// sensor poses; returns the number of nodes (rows may be NULL to ask for the count)
int64_t ref_planner_search_space(void* h, double* rows_xyzyaw, double* robot_poses, double* sensor_poses, int64_t cap) {
  RefPlanner* r = static_cast<RefPlanner*>(h);
  geometry_msgs::PoseArray robot;
  std::vector<geometry_msgs::PoseArray> sensors;
  r->planner->getRobotSensorPoses(robot, sensors);
  const int64_t n = (int64_t)robot.poses.size();
  for (int64_t j = 0; j < n && j < cap; ++j) {
    const geometry_msgs::Pose& p = robot.poses[(size_t)j];
    if (rows_xyzyaw) {
      tf::Quaternion qt(p.orientation.x, p.orientation.y, p.orientation.z, p.orientation.w);
      rows_xyzyaw[4 * j] = p.position.x; rows_xyzyaw[4 * j + 1] = p.position.y; rows_xyzyaw[4 * j + 2] = p.position.z;
      rows_xyzyaw[4 * j + 3] = tf::getYaw(qt);
    }
    if (robot_poses) {
      double* o = robot_poses + 7 * j;
      o[0] = p.position.x; o[1] = p.position.y; o[2] = p.position.z;
      o[3] = p.orientation.x; o[4] = p.orientation.y; o[5] = p.orientation.z; o[6] = p.orientation.w;
    }
    if (sensor_poses)
      for (size_t s = 0; s < sensors.size() && s < 2; ++s) {
        const geometry_msgs::Pose& q = sensors[s].poses[(size_t)j];
        double* o = sensor_poses + 14 * j + 7 * s;
        o[0] = q.position.x; o[1] = q.position.y; o[2] = q.position.z;
        o[3] = q.orientation.x; o[4] = q.orientation.y; o[5] = q.orientation.z; o[6] = q.orientation.w;
      }
  }
  return n;
}
int64_t ref_planner_num_connections(void* h) {
  return (int64_t)static_cast<RefPlanner*>(h)->planner->getConnections().size() / 2;
}

// CoveragePathPlanningHeuristic::pointCloudDiff (heuristic.cpp:88-145) of an arbitrary covered cloud
double ref_planner_point_cloud_diff(void* h, const float* xyz, int64_t n, float* centres, int64_t cap, int64_t* n_out) {
  RefPlanner* r = static_cast<RefPlanner*>(h);
  pcl::PointCloud<pcl::PointXYZ>::Ptr c(new pcl::PointCloud<pcl::PointXYZ>), d(new pcl::PointCloud<pcl::PointXYZ>);
  for (int64_t i = 0; i < n; ++i) c->points.push_back(pcl::PointXYZ(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]));
  const double pct = r->heuristic->pointCloudDiff(c, d);
  *n_out = (int64_t)d->points.size();
  for (int64_t i = 0; i < *n_out && i < cap; ++i) {
    centres[3 * i] = d->points[(size_t)i].x; centres[3 * i + 1] = d->points[(size_t)i].y; centres[3 * i + 2] = d->points[(size_t)i].z;
  }
  return pct;
}

// clusteringPointCloud (heuristic.cpp:147-190) + findClusterBB (:218-246): labels per input point (-1: in no cluster),
// boxes: per cluster {start x y z, size x y z}
int64_t ref_planner_clusters(void* h, const float* xyz, int64_t n, int32_t* labels, double* boxes, int64_t cap) {
  RefPlanner* r = static_cast<RefPlanner*>(h);
  pcl::PointCloud<pcl::PointXYZ>::Ptr c(new pcl::PointCloud<pcl::PointXYZ>);
  for (int64_t i = 0; i < n; ++i) c->points.push_back(pcl::PointXYZ(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]));
  std::vector<pcl::PointCloud<pcl::PointXYZ> > cl;
  r->heuristic->clusteringPointCloud(cl, c);
  for (int64_t i = 0; i < n; ++i) labels[i] = -1;
  for (size_t k = 0; k < cl.size(); ++k) {
    for (size_t j = 0; j < cl[k].points.size(); ++j)
      for (int64_t i = 0; i < n; ++i)
        if (std::memcmp(c->points[(size_t)i].data, cl[k].points[j].data, 12) == 0) { labels[i] = (int32_t)k; break; }
    if ((int64_t)k < cap) {
      geometry_msgs::Vector3 sz;
      geometry_msgs::Pose st;
      r->heuristic->findClusterBB(cl[k], sz, st);
      double* o = boxes + 6 * k;
      o[0] = st.position.x; o[1] = st.position.y; o[2] = st.position.z; o[3] = sz.x; o[4] = sz.y; o[5] = sz.z;
    }
  }
  return (int64_t)cl.size();
}

// setMaxMinSensorAccuracy over the current search space (coverage_heuristic_test.cpp:155-158)
void ref_planner_set_accuracy_range(void* h) {
  RefPlanner* r = static_cast<RefPlanner*>(h);
  geometry_msgs::PoseArray robot;
  std::vector<geometry_msgs::PoseArray> sensors;
  r->planner->getRobotSensorPoses(robot, sensors);
  r->heuristic->setMaxMinSensorAccuracy(sensors);
}

// PathPlanner::startSearch (coverage_heuristic_test.cpp:188): path rows {x y z qx qy qz qw coverage f distance}
int64_t ref_planner_search(void* h, const double start_xyz[3], double* rows, int64_t cap) {
  RefPlanner* r = static_cast<RefPlanner*>(h);
  Pose start(start_xyz[0], start_xyz[1], start_xyz[2], 0.0);
  SSPP::Node* path = r->planner->startSearch(start);
  int64_t n = 0;
  for (SSPP::Node* p = path; p != NULL; p = p->next, ++n) {
    if (n >= cap || !rows) continue;
    double* o = rows + 10 * n;
    o[0] = p->pose.p.position.x; o[1] = p->pose.p.position.y; o[2] = p->pose.p.position.z;
    o[3] = p->pose.p.orientation.x; o[4] = p->pose.p.orientation.y; o[5] = p->pose.p.orientation.z; o[6] = p->pose.p.orientation.w;
    o[7] = p->coverage; o[8] = p->f_value; o[9] = p->distance;
  }
  return n;
}

// InfoGainVolumetricH (heuristic.cpp:643-759, 974-1085, 1128-1225) on a chain of k nodes: node 0 is a root, node i's parent
// is node i-1 -- the REAL scored node, which owns the octree its block creates.  (Under the reference's own A* a child's
// parent is the copy made at astar.cpp:262, whose octree pointer Node::Node(Node*) never sets, node.cpp:43-68: the
// block is undefined there; on real parents it is the reference's code doing what it evidently means.)
// out: k x 6 {totalEntroby, coveredVolume, coverage, f_value, distance, coveredVoxelsNum}
int64_t ref_planner_infogain_chain(void* h, const double* robot_poses, const double* sensor_poses, int ns, int64_t k,
                                   double* out) {
  RefPlanner* r = static_cast<RefPlanner*>(h);
  std::vector<SSPP::Node*> nodes;
  for (int64_t i = 0; i < k; ++i) {
    SSPP::Node* n = new SSPP::Node();
    n->id = (int)i;
    n->depth = (int)i;
    n->parent = i ? nodes[(size_t)i - 1] : NULL;
    n->pose.p = to_pose(robot_poses + 7 * i);
    for (int s = 0; s < ns; ++s) {
      Pose sp;
      sp.p = to_pose(sensor_poses + 7 * (i * ns + s));
      n->senPoses.push_back(sp);
    }
    r->heuristic->calculateHeuristic(n);
    double* o = out + 6 * i;
    o[0] = n->totalEntroby; o[1] = n->coveredVolume; o[2] = n->coverage; o[3] = n->f_value; o[4] = n->distance;
    o[5] = (double)n->coveredVoxelsNum;
    nodes.push_back(n);
  }
  for (size_t i = 0; i < nodes.size(); ++i) { delete nodes[i]->octree; delete nodes[i]; }
  return k;
}

}  // extern "C"
