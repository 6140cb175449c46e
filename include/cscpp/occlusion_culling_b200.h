// occlusion_culling_b200.h -- drop-in shim of the reference's OcclusionCulling class
// (cscpp/include/cscpp/occlusion_culling.h:50-95) on top of the C ABI (include/asscpp_gpu.h).
//
// Same class name, same public data members the callers touch (cloud, cloudCopy, filtered_cloud,
// occlusionFreeCloud, FrustumCloud, freeCloud, voxelRes, OriginalVoxelsSize, maxAccuracyError, ...),
// same method signatures on the hot path:
//     pcl::PointCloud<pcl::PointXYZ> extractVisibleSurface(geometry_msgs::Pose)   (:82,  .cpp:265-373)
//     float calcCoveragePercent(pcl::PointCloud<pcl::PointXYZ>::Ptr)              (:84,  .cpp:375-444)
// plus non-breaking batched entry points used by the viewpoint batcher.  With real ROS/PCL headers on
// the include path the same source builds against them; otherwise the stand-ins under compat/ are used.
// All four constructors of the reference are there (occlusion_culling.h:77-80).
//
// There is no CPU fallback: construction throws std::runtime_error when no sm_100 GPU is usable.
#pragma once
#include <cstdlib>
#include <cstring>
#include <limits>
#include <stdexcept>
#include <string>
#include <vector>

#include <ros/ros.h>
#include <geometry_msgs/Point32.h>
#include <geometry_msgs/Pose.h>
#include <geometry_msgs/PoseArray.h>
#include <pcl/io/pcd_io.h>
#include <pcl/point_types.h>
#include <tf/tf.h>
#include <visualization_msgs/Marker.h>
#include <visualization_msgs/MarkerArray.h>

#include "../asscpp_gpu.h"

class OcclusionCulling {
 public:
  // attributes kept from the reference (occlusion_culling.h:54-74)
  std::string model;
  pcl::PointCloud<pcl::PointXYZ>::Ptr cloud;
  pcl::PointCloud<pcl::PointXYZ>::Ptr cloudCopy;
  pcl::PointCloud<pcl::PointXYZ>::Ptr filtered_cloud;
  pcl::PointCloud<pcl::PointXYZ>::Ptr occlusionFreeCloud;
  pcl::PointCloud<pcl::PointXYZ>::Ptr FrustumCloud;
  pcl::PointCloud<pcl::PointXYZ> freeCloud;
  float voxelRes, OriginalVoxelsSize, viewEntropy;
  double id;
  double maxAccuracyError, minAccuracyError;
  bool AccuracyMaxSet;
  std::string frame_id;
  // The reference appends every visible point of every call to occlusionFreeCloud forever
  // (occlusion_culling.cpp:363).  Kept on by default for fidelity; the batcher turns it off.
  bool accumulateOcclusionFreeCloud;

  ros::NodeHandle nh;
  ros::Publisher fov_pub;
  // occlusion_culling.cpp:26-79: modelName is the full path of the PCD file (coverage_path_planning_heuristic.cpp:32)
  OcclusionCulling(ros::NodeHandle& n, std::string modelName) : model(modelName), nh(n) {
    init_members();
    pcl::io::loadPCDFile<pcl::PointXYZ>(model, *cloud);
    finish_init(device_from_env());
  }
  // occlusion_culling.cpp:81-135
  OcclusionCulling(ros::NodeHandle& n, pcl::PointCloud<pcl::PointXYZ>::Ptr& cloudPtr) : nh(n) {
    init_members();
    cloud->points = cloudPtr->points;  // :94 copies only ->points: cloud->sensor_origin_ stays zero
    finish_init(device_from_env());
  }
  // occlusion_culling.cpp:190-: the reference's default constructor loads a hard-coded scene file of another package
  // ("scaled_desktop.pcd" under component_test); here it makes an object over an empty cloud
  OcclusionCulling() {
    init_members();
    finish_init(device_from_env());
  }
  explicit OcclusionCulling(std::string modelName, int device = 0) : model(modelName) {
    init_members();
    pcl::io::loadPCDFile(model, *cloud);  // occlusion_culling.cpp:39 (a missing file yields an empty cloud)
    finish_init(device);
  }
  explicit OcclusionCulling(pcl::PointCloud<pcl::PointXYZ>::Ptr& cloudPtr, int device = 0) {
    init_members();
    cloud->points = cloudPtr->points;  // occlusion_culling.cpp:94
    // only ->points is copied (occlusion_culling.cpp:94): cloud->sensor_origin_ stays zero, as in the reference
    finish_init(device);
  }
  ~OcclusionCulling() { asscpp_destroy(ctx_); }
  OcclusionCulling(const OcclusionCulling&) = delete;
  OcclusionCulling& operator=(const OcclusionCulling&) = delete;

  // occlusion_culling.cpp:265-373
  pcl::PointCloud<pcl::PointXYZ> extractVisibleSurface(geometry_msgs::Pose location) {
    std::vector<int32_t> idx;
    visibleIndices(location, idx);
    freeCloud.points.resize(idx.size());
    for (size_t k = 0; k < idx.size(); ++k) freeCloud.points[k] = cloudCopy->points[(size_t)idx[k]];
    freeCloud.width = (unsigned)idx.size();
    if (accumulateOcclusionFreeCloud)
      occlusionFreeCloud->points.insert(occlusionFreeCloud->points.end(), freeCloud.points.begin(),
                                        freeCloud.points.end());
    return freeCloud;
  }

  // occlusion_culling.cpp:375-444: any cloud (e.g. VoxelGrid centroids, heuristic.cpp:476-482)
  float calcCoveragePercent(pcl::PointCloud<pcl::PointXYZ>::Ptr cloud_filtered) {
    std::vector<float> xyz(3 * cloud_filtered->points.size());
    for (size_t i = 0; i < cloud_filtered->points.size(); ++i) {
      xyz[3 * i] = cloud_filtered->points[i].x;
      xyz[3 * i + 1] = cloud_filtered->points[i].y;
      xyz[3 * i + 2] = cloud_filtered->points[i].z;
    }
    float pct = 0;
    check(asscpp_coverage_percent_cloud(ctx_, xyz.data(), (int64_t)cloud_filtered->points.size(), &pct));
    return pct;
  }

  // occlusion_culling.cpp:445-458: raw z of the given cloud as depth (as the reference does)
  double calcAvgAccuracy(pcl::PointCloud<pcl::PointXYZ> pointCloud) {
    double errorSum = 0;
    for (size_t j = 0; j < pointCloud.points.size(); ++j) {
      double val = pointCloud.points[j].z;
      double pointError = 0.0000285 * val * val;
      errorSum += pointError;
    }
    return errorSum / pointCloud.points.size();
  }
  // occlusion_culling.cpp:459-475: depth = x after pointCloudViewportTransform
  double calcAvgAccuracy(pcl::PointCloud<pcl::PointXYZ> pointCloud, geometry_msgs::Pose cameraPose) {
    pcl::PointCloud<pcl::PointXYZ> t = pointCloudViewportTransform(pointCloud, cameraPose);
    double errorSum = 0;
    for (size_t j = 0; j < t.points.size(); ++j) {
      double val = t.points[j].x;
      double pointError = 0.0000285 * val * val;
      errorSum += pointError;
    }
    return errorSum / t.points.size();
  }
  // occlusion_culling.cpp:538-618 (host arithmetic of the C ABI, same as the device path)
  pcl::PointCloud<pcl::PointXYZ> pointCloudViewportTransform(pcl::PointCloud<pcl::PointXYZ> pointCloud,
                                                             geometry_msgs::Pose cameraPose) {
    double p[7];
    pack(cameraPose, p);
    std::vector<float> in(3 * pointCloud.points.size()), out(3 * pointCloud.points.size());
    for (size_t i = 0; i < pointCloud.points.size(); ++i) {
      in[3 * i] = pointCloud.points[i].x; in[3 * i + 1] = pointCloud.points[i].y; in[3 * i + 2] = pointCloud.points[i].z;
    }
    check(asscpp_viewport_transform(p, in.data(), (int64_t)pointCloud.points.size(), out.data()));
    pcl::PointCloud<pcl::PointXYZ> posArray;
    posArray.points.resize(pointCloud.points.size());
    for (size_t i = 0; i < pointCloud.points.size(); ++i) {
      posArray.points[i].x = out[3 * i]; posArray.points[i].y = out[3 * i + 1]; posArray.points[i].z = out[3 * i + 2];
    }
    return posArray;
  }
  // occlusion_culling.cpp:486-520: depth extrema over the whole search space in ONE batched launch; keeps the
  // reference's initial values (max starts at numeric_limits<double>::min(), the smallest positive double)
  void SSMaxMinAccuracy(std::vector<geometry_msgs::PoseArray> sensorsPoses) {
    std::vector<geometry_msgs::Pose> flat;
    for (size_t i = 0; i < sensorsPoses.size(); ++i)
      flat.insert(flat.end(), sensorsPoses[i].poses.begin(), sensorsPoses[i].poses.end());
    std::vector<AsscppViewResult> res;
    std::vector<AsscppViewAccuracy> acc;
    evalBatch(flat, res, -1, ASSCPP_EVAL_ACCURACY, &acc);
    double max = std::numeric_limits<double>::min(), min = std::numeric_limits<double>::max();
    for (size_t k = 0; k < acc.size(); ++k) {
      if (!acc[k].visible) continue;
      if ((double)acc[k].depth_max > max) max = acc[k].depth_max;
      if ((double)acc[k].depth_min < min) min = acc[k].depth_min;
    }
    maxAccuracyError = 0.0000285 * max * max;
    minAccuracyError = 0.0000285 * min * min;
    AccuracyMaxSet = true;
  }

  // occlusion_culling.cpp:521-531
  void transformPointMatVec(tf::Vector3 translation, tf::Matrix3x3 rotation, geometry_msgs::Point32 in,
                            geometry_msgs::Point32& out) {
    double x = rotation[0].x() * in.x + rotation[0].y() * in.y + rotation[0].z() * in.z + translation.x();
    double y = rotation[1].x() * in.x + rotation[1].y() * in.y + rotation[1].z() * in.z + translation.y();
    double z = rotation[2].x() * in.x + rotation[2].y() * in.y + rotation[2].z() * in.z + translation.z();
    out.x = x; out.y = y; out.z = z;
  }
  // occlusion_culling.cpp:244-252
  bool contains(pcl::PointCloud<pcl::PointXYZ> c, pcl::PointXYZ p) {
    for (size_t i = 0; i < c.points.size(); ++i)
      if (c.points[i].x == p.x && c.points[i].y == p.y && c.points[i].z == p.z) return true;
    return false;
  }
  // occlusion_culling.cpp:254-263: points of the model cloud that are not in c2
  pcl::PointCloud<pcl::PointXYZ> pointsDifference(pcl::PointCloud<pcl::PointXYZ> c2) {
    pcl::PointCloud<pcl::PointXYZ> inter;
    for (size_t i = 0; i < cloud->points.size(); ++i)
      if (!contains(c2, cloud->points[i])) inter.push_back(cloud->points[i]);
    return inter;
  }
  // occlusion_culling.cpp:622-718: RViz markers of the frustum.  Unavailable on this box (no graphics stack): the
  // methods exist so that callers compile and link; they publish nothing.
  void visualizeFOV(geometry_msgs::Pose) {}
  visualization_msgs::Marker drawLines(std::vector<geometry_msgs::Point> links, int id_, int[]) {
    visualization_msgs::Marker m;
    m.id = id_;
    m.points = links;
    return m;
  }

  // ---- additions (non-breaking) -----------------------------------------------------------------
  // ascending original indices of extractVisibleSurface's cloud
  void visibleIndices(const geometry_msgs::Pose& location, std::vector<int32_t>& idx) {
    double p[7];
    pack(location, p);
    idx.resize(cloudCopy->points.size());
    int64_t n = 0;
    check(asscpp_get_visible(ctx_, p, idx.data(), (int64_t)idx.size(), &n));
    idx.resize((size_t)n);
  }
  // one launch for a whole batch of viewpoints; covered_set = -1 scores against the empty set
  void evalBatch(const std::vector<geometry_msgs::Pose>& poses, std::vector<AsscppViewResult>& out,
                 int covered_set = -1, uint32_t flags = 0, std::vector<AsscppViewAccuracy>* accuracy = nullptr) {
    std::vector<double> p(7 * poses.size());
    for (size_t i = 0; i < poses.size(); ++i) pack(poses[i], &p[7 * i]);
    out.resize(poses.size());
    if (accuracy) flags |= ASSCPP_EVAL_ACCURACY;
    check(asscpp_eval_batch(ctx_, p.data(), (int64_t)poses.size(), covered_set, flags));
    check(asscpp_get_results(ctx_, out.data(), (int64_t)poses.size()));
    if (accuracy) {
      accuracy->resize(poses.size());
      check(asscpp_get_accuracy(ctx_, accuracy->data(), (int64_t)poses.size()));
    }
  }
  AsscppCtx* context() { return ctx_; }
  static int device_from_env() {  // the reference's constructors have no device argument: ASSCPP_DEVICE picks the GPU
    const char* e = std::getenv("ASSCPP_DEVICE");
    return e ? std::atoi(e) : 0;
  }
  static void pack(const geometry_msgs::Pose& q, double p[7]) {
    p[0] = q.position.x; p[1] = q.position.y; p[2] = q.position.z;
    p[3] = q.orientation.x; p[4] = q.orientation.y; p[5] = q.orientation.z; p[6] = q.orientation.w;
  }
  void check(int rc) const {
    if (rc != ASSCPP_OK) throw std::runtime_error(std::string("asscpp: ") + asscpp_last_error(ctx_));
  }

 private:
  AsscppCtx* ctx_ = nullptr;
  void init_members() {
    typedef pcl::PointCloud<pcl::PointXYZ> Cloud;
    cloud = Cloud::Ptr(new Cloud);
    cloudCopy = Cloud::Ptr(new Cloud);
    filtered_cloud = Cloud::Ptr(new Cloud);
    occlusionFreeCloud = Cloud::Ptr(new Cloud);
    FrustumCloud = Cloud::Ptr(new Cloud);
    voxelRes = 0.5f;  // occlusion_culling.cpp:41
    OriginalVoxelsSize = 0.0f;
    viewEntropy = 0.0f;
    id = 0.0;
    maxAccuracyError = minAccuracyError = 0.0;
    AccuracyMaxSet = false;
    frame_id = "world";
    accumulateOcclusionFreeCloud = true;
  }
  void finish_init(int device) {
    cloudCopy->points = cloud->points;  // occlusion_culling.cpp:40
    AsscppParams prm;
    asscpp_default_params(&prm);  // VFOV 45, HFOV 58, near 0.7, far 6.0 (occlusion_culling.cpp:73-76)
    prm.leaf = voxelRes;
    prm.cov_leaf = voxelRes;      // calcCoveragePercent voxelises at voxelRes (occlusion_culling.cpp:390)
    for (int k = 0; k < 3; ++k) prm.pcd_origin[k] = cloud->sensor_origin_[k];  // used at :338-341
    std::vector<float> xyz(3 * cloud->points.size());
    for (size_t i = 0; i < cloud->points.size(); ++i) {
      xyz[3 * i] = cloud->points[i].x; xyz[3 * i + 1] = cloud->points[i].y; xyz[3 * i + 2] = cloud->points[i].z;
    }
    int rc = asscpp_create(device, xyz.data(), (int64_t)cloud->points.size(), &prm, &ctx_);
    if (rc != ASSCPP_OK) throw std::runtime_error(std::string("asscpp_create: ") + asscpp_last_error(nullptr));
    AsscppModelInfo info;
    asscpp_model_info(ctx_, &info);
    OriginalVoxelsSize = (float)info.n_cov_voxels;  // occlusion_culling.cpp:50-65
  }
};
