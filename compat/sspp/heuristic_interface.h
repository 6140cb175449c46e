// Stand-in for sspp/include/sspp/heuristic_interface.h:35-56: the ten pure virtuals, same signatures (see utils.h in
// this directory).  tests/test_boundary.py compiles the plugin against BOTH this file and the reference's own header.
#ifndef HEURISTICT_H_
#define HEURISTICT_H_
#include <vector>
#include <geometry_msgs/Vector3.h>
#include "node.h"
#include "searchspacenode.h"
#include "ssppexception.h"
#include "utils.h"
namespace SSPP {
class Heuristic {
 public:
  Heuristic() {}
  virtual ~Heuristic() {}
  virtual void calculateHeuristic(Node*) = 0;
  virtual bool terminateConditionReached(Node*) = 0;
  virtual bool isConnectionConditionSatisfied(SearchSpaceNode*, SearchSpaceNode*) = 0;
  virtual bool isFilteringConditionSatisfied(geometry_msgs::Pose pose, geometry_msgs::PoseArray& correspondingSensorPoses,
                                             double minDist, double maxDist, pcl::PointCloud<pcl::PointXYZ>& globalCloud,
                                             std::vector<pcl::PointCloud<pcl::PointXYZ> >& accuracyClusters,
                                             double accuracyThreshhold) = 0;
  virtual void displayProgress(std::vector<Tree> tree) = 0;
  virtual double pointCloudDiff(pcl::PointCloud<pcl::PointXYZ>::Ptr globalCloudPtr,
                                pcl::PointCloud<pcl::PointXYZ>::Ptr& pointCloudDiffPtr) = 0;
  virtual void clusteringPointCloud(std::vector<pcl::PointCloud<pcl::PointXYZ> >& clustersPointCloudVec,
                                    pcl::PointCloud<pcl::PointXYZ>::Ptr pointCloudDiffPtr) = 0;
  virtual void findClusterOuterPoints(geometry_msgs::PoseArray waypoints, pcl::PointCloud<pcl::PointXYZ>& cloudHull) = 0;
  virtual void findClusterBB(pcl::PointCloud<pcl::PointXYZ> clusterPoints, geometry_msgs::Vector3& gridSize,
                             geometry_msgs::Pose& gridStart) = 0;
  virtual bool isCost() = 0;
};
}  // namespace SSPP
#endif
