// Stand-in for the PCL types the path's class surface uses (pcl::PointXYZ, pcl::PointCloud<T>), used ONLY when PCL
// is absent (this image has none).  Layout of PointXYZ matches PCL's (16-byte aligned x y z + padding).
#pragma once
#include <cmath>
#include <cstdint>
#include <memory>
#include <string>
#include <vector>
#include <Eigen/Core>
#include <cstdio>
#define pcl_isfinite(x) std::isfinite(x)
#define PCL_ERROR(...) std::fprintf(stderr, __VA_ARGS__)
#define PCL_WARN(...) std::fprintf(stderr, __VA_ARGS__)
#define PCL_INFO(...) std::fprintf(stderr, __VA_ARGS__)
namespace pcl {
struct PCLHeader { uint32_t seq = 0; uint64_t stamp = 0; std::string frame_id; };
struct alignas(16) PointXYZ {
  union {
    float data[4];
    struct { float x, y, z; };
  };
  PointXYZ() { data[0] = data[1] = data[2] = 0; data[3] = 1.0f; }
  PointXYZ(float x_, float y_, float z_) { data[0] = x_; data[1] = y_; data[2] = z_; data[3] = 1.0f; }
};
struct alignas(16) PointXYZRGB {
  float x = 0, y = 0, z = 0, pad = 1.0f;
  uint8_t b = 0, g = 0, r = 0, a = 255;
  PointXYZRGB() {}
  PointXYZRGB(uint8_t r_, uint8_t g_, uint8_t b_) : b(b_), g(g_), r(r_) {}
};
struct PointIndices { PCLHeader header; std::vector<int> indices; };
template <class PointT>
struct PointCloud {
  typedef std::shared_ptr<PointCloud<PointT> > Ptr;
  typedef std::shared_ptr<const PointCloud<PointT> > ConstPtr;
  typedef typename std::vector<PointT>::iterator iterator;
  typedef typename std::vector<PointT>::const_iterator const_iterator;
  PCLHeader header;
  std::vector<PointT> points;
  uint32_t width = 0, height = 1;
  bool is_dense = true;
  Eigen::Vector4f sensor_origin_;          // zero, as PCL initialises it
  Eigen::Quaternionf sensor_orientation_;  // identity
  size_t size() const { return points.size(); }
  bool empty() const { return points.empty(); }
  void clear() { points.clear(); width = 0; height = 0; }
  void resize(size_t n) { points.resize(n); if (width * height != n) { width = (uint32_t)n; height = 1; } }
  void push_back(const PointT& p) { points.push_back(p); width = (uint32_t)points.size(); height = 1; }
  PointT& at(size_t i) { return points.at(i); }
  const PointT& at(size_t i) const { return points.at(i); }
  PointT& operator[](size_t i) { return points[i]; }
  const PointT& operator[](size_t i) const { return points[i]; }
  iterator begin() { return points.begin(); }
  iterator end() { return points.end(); }
  const_iterator begin() const { return points.begin(); }
  const_iterator end() const { return points.end(); }
  Ptr makeShared() const { return Ptr(new PointCloud<PointT>(*this)); }
  PointCloud& operator+=(const PointCloud& o) {  // pcl::PointCloud::operator+= : concatenation
    points.insert(points.end(), o.points.begin(), o.points.end());
    width = (uint32_t)points.size();
    height = 1;
    if (!o.is_dense) is_dense = false;
    return *this;
  }
  PointCloud operator+(const PointCloud& o) const { PointCloud r(*this); r += o; return r; }
};
}  // namespace pcl
