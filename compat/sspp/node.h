// Stand-in for sspp/include/sspp/node.h:48-71 (see utils.h in this directory): the members a heuristic reads and
// writes, same names and types; Node(Node*) leaves the copy an EMPTY cloud_filtered, as sspp/src/node.cpp:43-68 does.
#ifndef NODE_H_
#define NODE_H_
#include <vector>
#include <octomap/octomap.h>
#include <pcl/point_types.h>
#include "utils.h"
namespace SSPP {
class Node {
 public:
  int id, depth, direction, coveredVoxelsNum;
  double nearest_obstacle, g_value, h_value, f_value, totalEntroby, distance, coverage, coveredVolume;
  octomap::OcTree* octree;
  Node *parent, *next, *prev;
  pcl::PointCloud<pcl::PointXYZ>::Ptr cloud_filtered;
  pcl::PointCloud<pcl::PointXYZ> cloud;
  pcl::PointCloud<pcl::PointXYZ> voxels;
  Pose pose;
  std::vector<Pose> senPoses;
  Node()
      : id(0), depth(0), direction(0), coveredVoxelsNum(0), nearest_obstacle(0), g_value(0), h_value(0), f_value(0),
        totalEntroby(0), distance(0), coverage(0), coveredVolume(0), octree(NULL), parent(NULL), next(NULL), prev(NULL),
        cloud_filtered(new pcl::PointCloud<pcl::PointXYZ>) {}
  explicit Node(Node* n)
      : id(n->id), depth(n->depth), direction(0), coveredVoxelsNum(n->coveredVoxelsNum),
        nearest_obstacle(n->nearest_obstacle), g_value(n->g_value), h_value(n->h_value), f_value(n->f_value),
        totalEntroby(n->totalEntroby), distance(n->distance), coverage(n->coverage), coveredVolume(n->coveredVolume),
        octree(NULL), parent(n->parent), next(n->next), prev(n->prev), cloud_filtered(new pcl::PointCloud<pcl::PointXYZ>),
        pose(n->pose), senPoses(n->senPoses) {}
};
}  // namespace SSPP
#endif
