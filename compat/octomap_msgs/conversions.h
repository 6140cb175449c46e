// Stand-in for octomap_msgs/conversions.h: serialisation for RViz, nothing to display here.
#pragma once
#include <octomap/octomap.h>
#include "Octomap.h"
namespace octomap_msgs {
template <class Tree> inline bool fullMapToMsg(const Tree&, Octomap&) { return false; }
template <class Tree> inline bool binaryMapToMsg(const Tree&, Octomap&) { return false; }
}  // namespace octomap_msgs
