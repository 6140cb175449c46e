// info_gain_b200.h -- InfoGainVolumetricH (cscpp/src/coverage_path_planning_heuristic.cpp:643-759, 974-1085, 1128-1225)
// in its evidently intended form (included by coverage_path_planning_heuristic_b200.h).
//
// The reference's block starts a child's tree as `new octomap::OcTree(*node->parent->octree)` (:649).  Under the
// reference's own A* the parent a child sees is the copy made at sspp/src/astar.cpp:262, and Node::Node(Node*)
// (sspp/src/node.cpp:43-68) copies neither `octree` nor initialises it: the reference dereferences an indeterminate
// pointer there, so its behaviour is undefined and there is no result to be identical to.  What the block evidently
// means -- the child inherits the tree of the node it was expanded from -- is offered here, opt-in
// (allowIntendedInfoGain()), and pinned against the reference's OWN block run on nodes whose parents are the real
// scored nodes, where that block is well defined (tests/test_ref_pin.py, tests/test_boundary.py).
//
// Division of work: the visible clouds of a node's sensors come from ONE launch (asscpp_eval_ex with
// ASSCPP_EVAL_VIEW_CELLS | ASSCPP_EVAL_ACCURACY); the kernel returns, per sensor, the set of coverage voxels the visible
// cloud touches.  At the reference's volumetricVoxelRes = 0.5 m that set IS the occupiedKeys of computeUpdate (:671): an
// octomap key is floor(coordinate / 0.5) + 32768 and the coverage voxel of a point is floor(x * 2.0f) -- the same
// cell, the scaling by two being exact in float32 -- and keyToCoord (cell + 0.5) * 0.5 is the centre
// asscpp_uncovered_voxels reports.  The per-key entropy arithmetic (a few dozen keys per sensor, double precision with
// the host's libm exactly as the reference writes it) stays on the host; the tree of a node is a flat array of float32
// log-odds per coverage voxel.  The only difference to the reference's numbers is the order of the double sums over
// a sensor's keys (the reference walks an unordered set): a few ulp of totalEntroby.
#pragma once

namespace SSPP {

// <octomap/octomap_utils.h>: logodds(p) = (float) log(p / (1 - p)), probability(l) = 1 - 1 / (1 + exp(l))
inline float CoveragePathPlanningHeuristic::oct_logodds(double p) { return (float)std::log(p / (1 - p)); }
inline double CoveragePathPlanningHeuristic::oct_probability(double l) { return 1. - (1. / (1. + std::exp(l))); }

inline void CoveragePathPlanningHeuristic::info_gain_init() {
  if (!info_centres_.empty()) return;
  AsscppModelInfo info;
  asscpp_model_info(ctx(), &info);
  const int64_t n = info.n_cov_voxels;
  // every coverage voxel of the model with its centre = keyToCoord of its octomap key: the uncovered voxels of an empty set
  std::vector<int32_t> ids((size_t)std::max<int64_t>(n, 1));
  info_centres_.assign((size_t)std::max<int64_t>(3 * n, 3), 0.0f);
  check(asscpp_covered_clear(ctx(), scratch_set_));
  int64_t got = 0;
  check(asscpp_uncovered_voxels(ctx(), scratch_set_, ids.data(), info_centres_.data(), n, &got));
  if (got != n) throw std::runtime_error("info gain: coverage voxel table incomplete");
  info_cells_ = n;
  const double res = volumetricVoxelRes;
  // :63-74: occupiedKeys of the full model, modelVolume, modelTotalEntroby (unknown = 0.5)
  modelVolume = (double)n * (res * res * res);
  modelTotalEntroby = (double)n * (-1 * (0.5) * (std::log(0.5) / std::log(2)));
}

inline void CoveragePathPlanningHeuristic::calculateInfoGain(Node* node) {
  info_gain_init();
  Node* parent = node->parent;
  const size_t ns = node->senPoses.size();
  const size_t n = (size_t)info_cells_;
  InfoTree tree;
  if (parent) {
    std::map<std::pair<int, const Node*>, InfoTree>::const_iterator it = info_trees_.find(std::make_pair(parent->id, (const Node*)parent->parent));
    if (it == info_trees_.end())
      throw std::runtime_error("info gain: the parent of this node was never scored by this heuristic (no tree to inherit)");
    tree = it->second;  // new octomap::OcTree(*node->parent->octree)  (:649)
    node->totalEntroby = parent->totalEntroby;        // :646-648
    node->coveredVoxelsNum = parent->coveredVoxelsNum;
    node->coveredVolume = parent->coveredVolume;
  } else {
    tree.logodds.assign(n, 0.0f);                      // new octomap::OcTree(volumetricVoxelRes)  (:1131)
    tree.known.assign(n, 0);
    node->totalEntroby = modelTotalEntroby;            // :1136
  }
  // setClampingThresMin(0.5) / Max(0.9999999) (:652-653): log-odds bounds applied by setNodeValue
  const float clamp_min = oct_logodds(0.5), clamp_max = oct_logodds(0.9999999);
  const double max_err = occlussionCulling->maxAccuracyError;
  const double res = volumetricVoxelRes;

  const int64_t words = asscpp_view_cells_words(ctx());
  // the visible clouds of this node's sensors: from the cache of the expansion's batch, or one launch -- for this
  // node alone, or (batchSiblingsFrom) for it and every not yet scored sibling in Astar::childList
  InfoViews mine;
  std::map<const Node*, InfoViews>::iterator hit = info_cache_.find(node);
  if (hit != info_cache_.end() && info_cache_parent_ == parent) {
    mine.rows.swap(hit->second.rows); mine.res.swap(hit->second.res); mine.acc.swap(hit->second.acc);
    info_cache_.erase(hit);
  } else {
    std::vector<Node*> group(1, node);
    if (sibling_source_ && parent)
      for (Node* sib = *sibling_source_; sib != NULL; sib = sib->next)
        if (sib != node && sib->parent == parent && sib->senPoses.size() == ns) group.push_back(sib);
    info_cache_.clear();
    info_cache_parent_ = parent;
    const size_t nv = ns * group.size();
    if (nv > 4096) group.resize(std::max<size_t>(1, 4096 / std::max<size_t>(ns, 1)));  // the limit of ASSCPP_EVAL_VIEW_CELLS
    const size_t nviews = ns * group.size();
    std::vector<uint32_t> rows((size_t)words * std::max<size_t>(nviews, 1));
    std::vector<AsscppViewResult> vres(std::max<size_t>(nviews, 1));
    std::vector<AsscppViewAccuracy> vacc(std::max<size_t>(nviews, 1));
    if (nviews) {
      std::vector<double> p(7 * nviews);
      for (size_t i = 0; i < group.size(); ++i)
        for (size_t s = 0; s < ns; ++s) OcclusionCulling::pack(group[i]->senPoses[s].p, &p[7 * (i * ns + s)]);
      AsscppEvalOpts o;
      std::memset(&o, 0, sizeof(o));
      o.covered_set = -1; o.accum_set = -1; o.group_size = 1;
      o.flags = (uint32_t)(ASSCPP_EVAL_VIEW_CELLS | ASSCPP_EVAL_ACCURACY);
      check(asscpp_eval_ex(ctx(), p.data(), (int64_t)nviews, &o));
      check(asscpp_get_results(ctx(), vres.data(), (int64_t)nviews));
      check(asscpp_get_accuracy(ctx(), vacc.data(), (int64_t)nviews));
      check(asscpp_get_view_cells(ctx(), rows.data(), (int64_t)nviews));
      launches_scored_ += 1;
    }
    for (size_t i = 0; i < group.size(); ++i) {
      InfoViews& dst = i == 0 ? mine : info_cache_[group[i]];
      dst.rows.assign(rows.begin() + (size_t)words * ns * i, rows.begin() + (size_t)words * ns * (i + 1));
      dst.res.assign(vres.begin() + ns * i, vres.begin() + ns * (i + 1));
      dst.acc.assign(vacc.begin() + ns * i, vacc.begin() + ns * (i + 1));
    }
  }
  const std::vector<uint32_t>& rows = mine.rows;
  const std::vector<AsscppViewResult>& vres = mine.res;
  const std::vector<AsscppViewAccuracy>& vacc = mine.acc;
  nodes_scored_ += 1;

  std::vector<size_t> ids;
  std::vector<float> centres, moved;
  for (size_t s = 0; s < ns; ++s) {
    double pose[7];
    OcclusionCulling::pack(node->senPoses[s].p, pose);
    const uint32_t* row = &rows[(size_t)words * s];
    // the occupied keys of this sensor's visible cloud (:674, :1004, :1155), their centres (keyToCoord, :676) and the
    // depth of each centre: x of pointCloudViewportTransform (:681-682)
    ids.clear(); centres.clear();
    for (size_t id = 0; id < n; ++id)
      if ((row[id >> 5] >> (id & 31)) & 1u) {
        ids.push_back(id);
        centres.push_back(info_centres_[3 * id]); centres.push_back(info_centres_[3 * id + 1]); centres.push_back(info_centres_[3 * id + 2]);
      }
    moved.assign(centres.size() + 3, 0.0f);
    if (!ids.empty()) check(asscpp_viewport_transform(pose, centres.data(), (int64_t)ids.size(), moved.data()));
    for (size_t k = 0; k < ids.size(); ++k) {
      const size_t id = ids[k];
      double dist = moved[3 * k];
      if (dist > maxDepth) dist = maxDepth;                                                       // :684-685
      const double normAcc = (max_err - (0.0000285 * dist * dist * 0.5)) / max_err;               // :686
      double postProb, before;
      if (tree.known[id]) {                                                                       // :690-703
        const double prob = oct_probability(tree.logodds[id]);
        before = (-1 * prob * (std::log(prob) / std::log(2)));
        postProb = (normAcc * prob) / ((normAcc * prob) + ((1 - normAcc) * (1 - prob)));
      } else {                                                                                    // :704-718
        node->coveredVoxelsNum++;
        before = (-1 * (0.5) * (std::log(0.5) / std::log(2)));
        postProb = (normAcc * 0.5) / ((normAcc * 0.5) + ((1 - normAcc) * (1 - 0.5)));
        node->coveredVolume += (res * res * res);
      }
      node->totalEntroby -= before;
      const double logO = oct_logodds(postProb);
      tree.logodds[id] = std::min(std::max((float)logO, clamp_min), clamp_max);  // setNodeValue clamps
      tree.known[id] = 1;
      node->totalEntroby += (-1 * postProb * (std::log(postProb) / std::log(2)));
    }
    if (vres[s].visible != 0) {                                                                   // :723-730
      const double avgAcc = vacc[s].err_sum_depth / (double)vres[s].visible;
      const double a = (max_err - avgAcc) / max_err;
      accuracyPerViewpointAvg.push_back(a);
      accuracySum += avgAcc;
    }
  }
  node->coverage = ((double)node->coveredVolume / modelVolume) * 100;                             // :735, :1218
  if (parent) {
    node->h_value = 0;
    node->g_value = 0;
    const double d = Dist(node->pose.p, parent->pose.p);                                          // :448
    node->distance = parent->distance + d;                                                       // :450
    const double c = node->coverage - parent->coverage;                                           // :737
    extraCovPerViewpointAvg.push_back(c);
    extraCovSum += c;
    const double localE = node->totalEntroby - parent->totalEntroby;                              // :754
    node->f_value = parent->f_value + localE * std::exp(d * -0.2);                                // :756, :1083
  } else {
    extraCovPerViewpointAvg.push_back(node->coverage);                                            // :1219-1221
    extraCovSum += node->coverage;
    node->f_value = node->totalEntroby;                                                           // :1223
  }
  info_trees_[std::make_pair(node->id, (const Node*)node->parent)] = tree;
}

}  // namespace SSPP
