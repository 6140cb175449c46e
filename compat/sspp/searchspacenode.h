// Stand-in for sspp/include/sspp/searchspacenode.h:34-46 (see utils.h in this directory).
#ifndef SEARCHSPACENODE_H_
#define SEARCHSPACENODE_H_
#include <vector>
#include <geometry_msgs/Pose.h>
#include <geometry_msgs/PoseArray.h>
namespace SSPP {
class SearchSpaceNode {
 public:
  geometry_msgs::Pose location;
  geometry_msgs::PoseArray sensorLocation;
  SearchSpaceNode* next;
  int type, id;
  std::vector<SearchSpaceNode*> children;
  SearchSpaceNode() : next(NULL), type(0), id(0) {}
};
}  // namespace SSPP
#endif
