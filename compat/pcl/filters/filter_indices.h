// Stand-in for pcl/filters/filter_indices.h (+ filter.h, pcl_base.h): the base classes the reference's two filter
// subclasses derive from (cscpp/include/cscpp/frustum_culling.h:73, voxel_grid_occlusion_estimation.h:52), with the
// behaviour of <PCL 1.7 Filter::filter>: default indices = every point, header and sensor pose copied input -> output
// before applyFilter.
#pragma once
#include <limits>
#include <string>
#include <vector>
#include <boost/shared_ptr.hpp>
#include "../common/io.h"
#include "../point_types.h"
namespace pcl {
typedef std::shared_ptr<std::vector<int> > IndicesPtr;
typedef std::shared_ptr<const std::vector<int> > IndicesConstPtr;
template <class PointT> class PCLBase {
 public:
  typedef pcl::PointCloud<PointT> PointCloud;
  typedef typename PointCloud::Ptr PointCloudPtr;
  typedef typename PointCloud::ConstPtr PointCloudConstPtr;
  PCLBase() : use_indices_(false), fake_indices_(false) {}
  virtual ~PCLBase() {}
  virtual void setInputCloud(const PointCloudConstPtr& cloud) { input_ = cloud; }
  PointCloudConstPtr const getInputCloud() const { return input_; }
  virtual void setIndices(const IndicesPtr& indices) { indices_ = indices; fake_indices_ = false; use_indices_ = true; }
  IndicesPtr const getIndices() { return indices_; }

 protected:
  PointCloudConstPtr input_;
  IndicesPtr indices_;
  bool use_indices_, fake_indices_;
  bool initCompute() {
    if (!input_) return false;
    if (!indices_) { fake_indices_ = true; indices_.reset(new std::vector<int>); }
    if (fake_indices_ && indices_->size() != input_->points.size()) {
      const size_t old = indices_->size();
      indices_->resize(input_->points.size());
      for (size_t i = old; i < indices_->size(); ++i) (*indices_)[i] = (int)i;
    }
    return true;
  }
  bool deinitCompute() { return true; }
};
template <class PointT> class Filter : public PCLBase<PointT> {
 public:
  typedef pcl::PointCloud<PointT> PointCloud;
  typedef typename PointCloud::Ptr PointCloudPtr;
  typedef typename PointCloud::ConstPtr PointCloudConstPtr;
  using PCLBase<PointT>::input_;
  using PCLBase<PointT>::indices_;
  Filter(bool extract_removed_indices = false)
      : removed_indices_(new std::vector<int>), extract_removed_indices_(extract_removed_indices) {}
  virtual ~Filter() {}
  IndicesConstPtr const getRemovedIndices() { return removed_indices_; }
  void filter(PointCloud& output) {
    if (!this->initCompute()) return;
    if (input_.get() == &output) {
      PointCloud tmp;
      applyFilter(tmp);
      tmp.header = input_->header;
      tmp.sensor_origin_ = input_->sensor_origin_;
      tmp.sensor_orientation_ = input_->sensor_orientation_;
      pcl::copyPointCloud(tmp, output);
    } else {
      output.header = input_->header;
      output.sensor_origin_ = input_->sensor_origin_;
      output.sensor_orientation_ = input_->sensor_orientation_;
      applyFilter(output);
    }
    this->deinitCompute();
  }

 protected:
  IndicesPtr removed_indices_;
  std::string filter_name_;
  bool extract_removed_indices_;
  virtual void applyFilter(PointCloud& output) = 0;
  const std::string& getClassName() const { return filter_name_; }
};
template <class PointT> class FilterIndices : public Filter<PointT> {
 public:
  typedef pcl::PointCloud<PointT> PointCloud;
  FilterIndices(bool extract_removed_indices = false)
      : Filter<PointT>(extract_removed_indices), negative_(false), keep_organized_(false),
        user_filter_value_(std::numeric_limits<float>::quiet_NaN()) {}
  virtual ~FilterIndices() {}
  using Filter<PointT>::filter;
  void filter(std::vector<int>& indices) {
    if (!this->initCompute()) return;
    applyFilter(indices);
    this->deinitCompute();
  }
  void setNegative(bool n) { negative_ = n; }
  bool getNegative() { return negative_; }
  void setKeepOrganized(bool k) { keep_organized_ = k; }
  void setUserFilterValue(float v) { user_filter_value_ = v; }

 protected:
  using Filter<PointT>::applyFilter;
  bool negative_, keep_organized_;
  float user_filter_value_;
  virtual void applyFilter(std::vector<int>& indices) = 0;
};
}  // namespace pcl
