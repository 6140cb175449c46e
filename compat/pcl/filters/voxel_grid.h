// Stand-in for pcl/filters/voxel_grid.h: pcl::VoxelGrid<PointT> with the binning of <PCL 1.7
// filters/impl/voxel_grid.hpp applyFilter>, restated from the published algorithm (PCL is not vendored by the
// reference and absent from this image):
//   min_b = floor(min_p * inverse_leaf), max_b likewise;  div_b = max_b - min_b + 1;  divb_mul = (1, div_x, div_x div_y);
//   per point: ijk = floor(p * inverse_leaf) - min_b, idx = ijk . divb_mul;  points sorted by idx;  one output point
//   (the float32 centroid, summed in sorted order) per occupied cell;  leaf_layout_[idx] = index of that output point;
//   getCentroidIndexAt checks only 0 <= idx < size (no per-axis range check).
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <limits>
#include <string>
#include <vector>
#include "../common/common.h"
#include "filter_indices.h"
namespace pcl {
template <class PointT> class VoxelGrid : public Filter<PointT> {
 protected:
  using Filter<PointT>::filter_name_;
  using Filter<PointT>::getClassName;
  using Filter<PointT>::input_;
  using Filter<PointT>::indices_;
  typedef typename Filter<PointT>::PointCloud PointCloud;

 public:
  VoxelGrid()
      : leaf_size_(Eigen::Vector4f::Zero()), inverse_leaf_size_(Eigen::Array4f::Zero()), downsample_all_data_(true),
        save_leaf_layout_(false), min_b_(Eigen::Vector4i::Zero()), max_b_(Eigen::Vector4i::Zero()),
        div_b_(Eigen::Vector4i::Zero()), divb_mul_(Eigen::Vector4i::Zero()), min_points_per_voxel_(0) {
    filter_name_ = "VoxelGrid";
  }
  virtual ~VoxelGrid() {}
  void setLeafSize(const Eigen::Vector4f& leaf) {
    leaf_size_ = leaf;
    if (leaf_size_[3] == 0) leaf_size_[3] = 1;
    for (int c = 0; c < 4; ++c) inverse_leaf_size_[c] = 1.0f / leaf_size_[c];  // Array4f::Ones() / leaf_size_.array()
  }
  void setLeafSize(float lx, float ly, float lz) { setLeafSize(Eigen::Vector4f(lx, ly, lz, 1.0f)); }
  Eigen::Vector3f getLeafSize() { return Eigen::Vector3f(leaf_size_[0], leaf_size_[1], leaf_size_[2]); }
  void setDownsampleAllData(bool d) { downsample_all_data_ = d; }
  void setMinimumPointsNumberPerVoxel(unsigned n) { min_points_per_voxel_ = n; }
  void setSaveLeafLayout(bool s) { save_leaf_layout_ = s; }
  bool getSaveLeafLayout() { return save_leaf_layout_; }
  Eigen::Vector3i getMinBoxCoordinates() { return Eigen::Vector3i(min_b_[0], min_b_[1], min_b_[2]); }
  Eigen::Vector3i getMaxBoxCoordinates() { return Eigen::Vector3i(max_b_[0], max_b_[1], max_b_[2]); }
  Eigen::Vector3i getNrDivisions() { return Eigen::Vector3i(div_b_[0], div_b_[1], div_b_[2]); }
  Eigen::Vector3i getDivisionMultiplier() { return Eigen::Vector3i(divb_mul_[0], divb_mul_[1], divb_mul_[2]); }
  int getCentroidIndex(const PointT& p) {
    return leaf_layout_.at((size_t)((Eigen::Vector4i((int)std::floor(p.x * inverse_leaf_size_[0]),
                                                     (int)std::floor(p.y * inverse_leaf_size_[1]),
                                                     (int)std::floor(p.z * inverse_leaf_size_[2]), 0) - min_b_).dot(divb_mul_)));
  }
  std::vector<int> getLeafLayout() { return leaf_layout_; }
  Eigen::Vector3i getGridCoordinates(float x, float y, float z) {
    return Eigen::Vector3i((int)std::floor(x * inverse_leaf_size_[0]), (int)std::floor(y * inverse_leaf_size_[1]),
                           (int)std::floor(z * inverse_leaf_size_[2]));
  }
  int getCentroidIndexAt(const Eigen::Vector3i& ijk) {
    int idx = (Eigen::Vector4i(ijk[0], ijk[1], ijk[2], 0) - min_b_).dot(divb_mul_);
    if (idx < 0 || idx >= (int)leaf_layout_.size()) return -1;
    return leaf_layout_[(size_t)idx];
  }

 protected:
  Eigen::Vector4f leaf_size_;
  Eigen::Array4f inverse_leaf_size_;
  bool downsample_all_data_, save_leaf_layout_;
  std::vector<int> leaf_layout_;
  Eigen::Vector4i min_b_, max_b_, div_b_, divb_mul_;
  unsigned min_points_per_voxel_;

  struct cloud_point_index_idx {
    unsigned idx, cloud_point_index;
    cloud_point_index_idx(unsigned i, unsigned c) : idx(i), cloud_point_index(c) {}
    bool operator<(const cloud_point_index_idx& p) const { return idx < p.idx; }
  };

  void applyFilter(PointCloud& output) {
    if (!input_) { output.width = output.height = 0; output.points.clear(); return; }
    output.height = 1;
    output.is_dense = true;
    Eigen::Vector4f min_p, max_p;
    getMinMax3D<PointT>(*input_, *indices_, min_p, max_p);
    const int64_t dx = (int64_t)((max_p[0] - min_p[0]) * inverse_leaf_size_[0]) + 1;
    const int64_t dy = (int64_t)((max_p[1] - min_p[1]) * inverse_leaf_size_[1]) + 1;
    const int64_t dz = (int64_t)((max_p[2] - min_p[2]) * inverse_leaf_size_[2]) + 1;
    if (dx * dy * dz > (int64_t)std::numeric_limits<int32_t>::max()) {
      std::fprintf(stderr, "[pcl::%s::applyFilter] Leaf size is too small for the input dataset. Integer indices would overflow.",
                   getClassName().c_str());
      output = *input_;
      return;
    }
    for (int c = 0; c < 3; ++c) {
      min_b_[c] = (int)std::floor(min_p[c] * inverse_leaf_size_[c]);
      max_b_[c] = (int)std::floor(max_p[c] * inverse_leaf_size_[c]);
    }
    div_b_ = max_b_ - min_b_ + Eigen::Vector4i::Ones();
    div_b_[3] = 0;
    divb_mul_ = Eigen::Vector4i(1, div_b_[0], div_b_[0] * div_b_[1], 0);
    std::vector<cloud_point_index_idx> index_vector;
    index_vector.reserve(indices_->size());
    for (std::vector<int>::const_iterator it = indices_->begin(); it != indices_->end(); ++it) {
      const PointT& p = input_->points[(size_t)*it];
      if (!input_->is_dense && (!pcl_isfinite(p.x) || !pcl_isfinite(p.y) || !pcl_isfinite(p.z))) continue;
      const int ijk0 = (int)(std::floor(p.x * inverse_leaf_size_[0]) - (float)min_b_[0]);
      const int ijk1 = (int)(std::floor(p.y * inverse_leaf_size_[1]) - (float)min_b_[1]);
      const int ijk2 = (int)(std::floor(p.z * inverse_leaf_size_[2]) - (float)min_b_[2]);
      const int idx = ijk0 * divb_mul_[0] + ijk1 * divb_mul_[1] + ijk2 * divb_mul_[2];
      index_vector.push_back(cloud_point_index_idx((unsigned)idx, (unsigned)*it));
    }
    // <PCL> uses std::sort (order inside a cell unspecified); a stable sort keeps the original order inside a cell,
    // which only fixes the float32 summation order of the centroid
    std::stable_sort(index_vector.begin(), index_vector.end());
    unsigned total = 0, index = 0;
    std::vector<std::pair<unsigned, unsigned> > first_and_last;
    first_and_last.reserve(index_vector.size());
    while (index < index_vector.size()) {
      unsigned i = index + 1;
      while (i < index_vector.size() && index_vector[i].idx == index_vector[index].idx) ++i;
      if (i - index >= min_points_per_voxel_) {
        ++total;
        first_and_last.push_back(std::make_pair(index, i));
      }
      index = i;
    }
    output.points.resize(total);
    if (save_leaf_layout_) {
      const uint32_t new_size = (uint32_t)(div_b_[0] * div_b_[1] * div_b_[2]);
      const uint32_t reinit = std::min<uint32_t>(new_size, (uint32_t)leaf_layout_.size());
      for (uint32_t i = 0; i < reinit; ++i) leaf_layout_[i] = -1;
      leaf_layout_.resize(new_size, -1);
    }
    index = 0;
    for (size_t cp = 0; cp < first_and_last.size(); ++cp) {
      const unsigned first = first_and_last[cp].first, last = first_and_last[cp].second;
      Eigen::Vector4f centroid = Eigen::Vector4f::Zero();
      for (unsigned i = first; i < last; ++i) {
        const PointT& p = input_->points[index_vector[i].cloud_point_index];
        centroid[0] += p.x; centroid[1] += p.y; centroid[2] += p.z;
      }
      if (save_leaf_layout_) leaf_layout_[index_vector[first].idx] = (int)index;
      centroid /= (float)(last - first);
      output.points[index].x = centroid[0];
      output.points[index].y = centroid[1];
      output.points[index].z = centroid[2];
      ++index;
    }
    output.width = (uint32_t)output.points.size();
  }
};
}  // namespace pcl
