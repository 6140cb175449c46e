// Stand-in for the part of octomap the reference uses (sspp/include/sspp/node.h:54-56;
// cscpp/src/coverage_path_planning_heuristic.cpp:63-73, 643-712, 974-1043, 1128-1195), used ONLY when octomap is
// absent (this image has none).  Restated from octomap's published behaviour:
//   * keys: k = int(floor(coordinate / resolution)) + 32768 per axis (16 levels), keyToCoord = (k - 32768 + 0.5) * res;
//   * computeUpdate: the occupied set is the set of keys of the end points (max range < 0: every point); the reference
//     never reads the free set, so the ray keys are not generated;
//   * occupancy is a float32 log-odds value per leaf; setNodeValue clamps to [clampingThresMin, clampingThresMax];
//     search returns the leaf or NULL.  The real tree prunes eight identical children into their parent: a search then
//     returns the parent, which holds the same value, so the flat map used here reads the same numbers.
//   * KeySet is an unordered set with octomap's KeyHash; its iteration order only decides the order of double sums.
#pragma once
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <unordered_map>
#include <unordered_set>
#include <vector>
namespace octomap {
inline float logodds(double probability) { return (float)std::log(probability / (1 - probability)); }
inline double probability(double logodds) { return 1. - (1. / (1. + std::exp(logodds))); }
struct point3d {
  float d[3];
  point3d() { d[0] = d[1] = d[2] = 0; }
  point3d(float x, float y, float z) { d[0] = x; d[1] = y; d[2] = z; }
  float x() const { return d[0]; } float y() const { return d[1]; } float z() const { return d[2]; }
  float operator()(int i) const { return d[i]; }
};
struct Pointcloud {
  std::vector<point3d> pts;
  void push_back(const point3d& p) { pts.push_back(p); }
  void push_back(float x, float y, float z) { pts.push_back(point3d(x, y, z)); }
  size_t size() const { return pts.size(); }
  void clear() { pts.clear(); }
};
struct OcTreeKey {
  uint16_t k[3];
  OcTreeKey() { k[0] = k[1] = k[2] = 0; }
  OcTreeKey(uint16_t a, uint16_t b, uint16_t c) { k[0] = a; k[1] = b; k[2] = c; }
  bool operator==(const OcTreeKey& o) const { return k[0] == o.k[0] && k[1] == o.k[1] && k[2] == o.k[2]; }
  bool operator!=(const OcTreeKey& o) const { return !(*this == o); }
  uint16_t operator[](int i) const { return k[i]; }
  struct KeyHash {
    size_t operator()(const OcTreeKey& key) const { return (size_t)key.k[0] + 1447 * (size_t)key.k[1] + 345637 * (size_t)key.k[2]; }
  };
};
typedef std::unordered_set<OcTreeKey, OcTreeKey::KeyHash> KeySet;
class OcTreeNode {
 public:
  OcTreeNode() : value(0) {}
  float getLogOdds() const { return value; }
  void setLogOdds(float l) { value = l; }
  double getOccupancy() const { return probability(value); }
  float value;
};
class OcTree {
 public:
  explicit OcTree(double res)
      : res_(res), factor_(1.0 / res), clamp_min_(logodds(0.1192)), clamp_max_(logodds(0.971)), prob_hit_(logodds(0.7)),
        prob_miss_(logodds(0.4)) {}
  double getResolution() const { return res_; }
  void setProbHit(double p) { prob_hit_ = logodds(p); }
  void setProbMiss(double p) { prob_miss_ = logodds(p); }
  void setClampingThresMin(double p) { clamp_min_ = logodds(p); }
  void setClampingThresMax(double p) { clamp_max_ = logodds(p); }
  bool coordToKeyChecked(const point3d& p, OcTreeKey& key) const {
    for (int i = 0; i < 3; ++i) {
      int scaled = ((int)std::floor(factor_ * (double)p(i))) + 32768;
      if (scaled < 0 || scaled >= 65536) return false;
      key.k[i] = (uint16_t)scaled;
    }
    return true;
  }
  point3d keyToCoord(const OcTreeKey& key) const {
    return point3d((float)((double((int)key.k[0] - 32768) + 0.5) * res_), (float)((double((int)key.k[1] - 32768) + 0.5) * res_),
                   (float)((double((int)key.k[2] - 32768) + 0.5) * res_));
  }
  void computeUpdate(const Pointcloud& scan, const point3d& origin, KeySet& free_cells, KeySet& occupied_cells,
                     double maxrange) {
    (void)free_cells;
    for (size_t i = 0; i < scan.pts.size(); ++i) {
      const point3d& p = scan.pts[i];
      if (maxrange >= 0.0) {
        double dx = p.x() - origin.x(), dy = p.y() - origin.y(), dz = p.z() - origin.z();
        if (std::sqrt(dx * dx + dy * dy + dz * dz) > maxrange) continue;
      }
      OcTreeKey key;
      if (coordToKeyChecked(p, key)) occupied_cells.insert(key);
    }
  }
  OcTreeNode* search(const OcTreeKey& key) {
    std::unordered_map<OcTreeKey, OcTreeNode, OcTreeKey::KeyHash>::iterator it = nodes_.find(key);
    return it == nodes_.end() ? NULL : &it->second;
  }
  OcTreeNode* setNodeValue(const OcTreeKey& key, float log_odds_value, bool = false) {
    log_odds_value = std::min(std::max(log_odds_value, clamp_min_), clamp_max_);
    OcTreeNode& n = nodes_[key];
    n.value = log_odds_value;
    return &n;
  }
  size_t size() const { return nodes_.size(); }

 private:
  double res_, factor_;
  float clamp_min_, clamp_max_, prob_hit_, prob_miss_;
  std::unordered_map<OcTreeKey, OcTreeNode, OcTreeKey::KeyHash> nodes_;
};
typedef OcTree ColorOcTree;
}  // namespace octomap
