// Exercises the C++ class surface the reference's callers use (OcclusionCulling::extractVisibleSurface /
// calcCoveragePercent, the SurfaceCoverageH node scoring and the ASSCPP acceptance rule) through the
// shim headers, and prints results for tests/test_cpp_shim.py to compare with the oracle.
//   test_shim <model.pcd> <poses.txt (x y z qx qy qz qw per line, two consecutive lines = one waypoint)>
#include <cstdio>
#include <fstream>
#include <sstream>

#include <cscpp/occlusion_culling_b200.h>
#include <sspp/coverage_heuristic_b200.h>

static unsigned long long mix64(unsigned long long x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

int main(int argc, char** argv) {
  if (argc < 3) return 2;
  OcclusionCulling oc{std::string(argv[1])};
  std::vector<geometry_msgs::Pose> poses;
  std::ifstream f(argv[2]);
  std::string line;
  while (std::getline(f, line)) {
    std::istringstream ss(line);
    geometry_msgs::Pose p;
    if (ss >> p.position.x >> p.position.y >> p.position.z >> p.orientation.x >> p.orientation.y >> p.orientation.z >>
        p.orientation.w)
      poses.push_back(p);
  }
  std::printf("model %zu points OriginalVoxelsSize %.0f\n", oc.cloud->points.size(), oc.OriginalVoxelsSize);
  // 1. extractVisibleSurface + calcCoveragePercent, one call per pose like the reference's loops
  for (size_t k = 0; k < poses.size(); ++k) {
    pcl::PointCloud<pcl::PointXYZ> vis = oc.extractVisibleSurface(poses[k]);
    std::vector<int32_t> idx;
    oc.visibleIndices(poses[k], idx);
    unsigned long long h = 0;
    bool same = idx.size() == vis.points.size();
    for (size_t i = 0; i < idx.size(); ++i) {
      h += mix64((unsigned long long)idx[i]);
      const pcl::PointXYZ& a = vis.points[i];
      const pcl::PointXYZ& b = oc.cloudCopy->points[(size_t)idx[i]];
      same = same && a.x == b.x && a.y == b.y && a.z == b.z;
    }
    pcl::PointCloud<pcl::PointXYZ>::Ptr vp(new pcl::PointCloud<pcl::PointXYZ>(vis));
    float pct = oc.calcCoveragePercent(vp);
    std::printf("view %zu visible %zu hash %llu cloud_matches %d coverage %.9g\n", k, vis.points.size(), h, (int)same,
                (double)pct);
  }
  // 2. SurfaceCoverageH greedy walk: waypoints = consecutive pose pairs (two sensors each)
  std::vector<SSPP::NodeB200> nodes(poses.size() / 2);
  for (size_t i = 0; i < nodes.size(); ++i) {
    nodes[i].id = (int)i;
    nodes[i].pose = poses[2 * i];
    nodes[i].senPoses.push_back(poses[2 * i]);
    nodes[i].senPoses.push_back(poses[2 * i + 1]);
  }
  SSPP::CoverageHeuristicB200 heur(&oc, 100.0, 1.0, SSPP::SurfaceCoverageH, true);
  std::vector<SSPP::NodeB200> path;
  path.reserve(16);
  SSPP::NodeB200 root = nodes[0];
  heur.calculateHeuristic(&root);
  path.push_back(root);
  std::printf("root coverage %.17g f %.17g\n", root.coverage, root.f_value);
  for (int step = 0; step < 4; ++step) {
    SSPP::NodeB200* parent = &path.back();
    std::vector<SSPP::NodeB200> kids(nodes.begin() + 1, nodes.end());
    std::vector<SSPP::NodeB200*> ptrs;
    for (size_t i = 0; i < kids.size(); ++i) { kids[i].parent = parent; ptrs.push_back(&kids[i]); }
    heur.batcher().prefetch(ptrs.data(), (int)ptrs.size(), parent->covered_set);
    int best = -1;
    for (size_t i = 0; i < kids.size(); ++i) {
      heur.calculateHeuristic(&kids[i]);  // astar.cpp:318, one child at a time
      if (best < 0 || kids[i].f_value > kids[(size_t)best].f_value) best = (int)i;
    }
    std::printf("step %d best %d coverage %.17g f %.17g\n", step, kids[(size_t)best].id, kids[(size_t)best].coverage,
                kids[(size_t)best].f_value);
    heur.ensureSet(&kids[(size_t)best]);  // the chosen node is the next one expanded (astar.cpp:262)
    path.push_back(kids[(size_t)best]);
    path.back().parent = nullptr;  // the walk keeps only coverage / f / set of the chosen node
    path.back().distance = 0;
  }
  std::printf("walk launches %d\n", heur.batcher().launches());
  {
    // 2b. the DEFAULT mode (accumulate = false, quirk Q10): one expansion from waypoint 0, children scored against the
    // empty set through the batcher; coverage = the child's own, c = coverage - parent.coverage
    SSPP::CoverageHeuristicB200 hq(&oc, 100.0, 1.0);
    SSPP::NodeB200 r0 = nodes[0];
    hq.calculateHeuristic(&r0);
    std::vector<SSPP::NodeB200> kids(nodes.begin() + 1, nodes.end());
    std::vector<SSPP::NodeB200*> ptrs;
    for (size_t i = 0; i < kids.size(); ++i) { kids[i].parent = &r0; ptrs.push_back(&kids[i]); }
    hq.batcher().prefetch(ptrs.data(), (int)ptrs.size(), hq.scoringSet(&r0));
    for (size_t i = 0; i < kids.size(); ++i) {
      hq.calculateHeuristic(&kids[i]);
      std::printf("q10 %d coverage %.17g f %.17g\n", kids[i].id, kids[i].coverage, kids[i].f_value);
    }
    std::printf("q10 launches %d root %.17g\n", hq.batcher().launches(), r0.coverage);
  }
  // 3. ASSCPP acceptance rule + uncovered percentage
  int global_set = -1;
  oc.check(asscpp_covered_new(oc.context(), &global_set));
  std::vector<std::vector<geometry_msgs::Pose> > cand;
  for (size_t i = 0; i < nodes.size(); ++i) cand.push_back(nodes[i].senPoses);
  std::vector<char> accepted;
  heur.filterCandidates(cand, global_set, accepted);
  int nacc = 0;
  for (size_t i = 0; i < accepted.size(); ++i) nacc += accepted[i];
  std::printf("accepted %d of %zu uncovered %.17g\n", nacc, accepted.size(), heur.uncoveredPercent(global_set));
  {
    std::vector<AsscppCluster> cl;
    heur.clusterUncovered(global_set, cl);
    std::printf("clusters %zu", cl.size());
    for (size_t k = 0; k < cl.size() && k < 3; ++k)
      std::printf(" | %d %.17g %.17g %.17g %.17g %.17g %.17g", cl[k].size, cl[k].grid_start[0], cl[k].grid_start[1],
                  cl[k].grid_start[2], cl[k].grid_size[0], cl[k].grid_size[1], cl[k].grid_size[2]);
    std::printf("\n");
  }
  // 4. accuracy terms: the two calcAvgAccuracy overloads per view and SSMaxMinAccuracy over all poses
  oc.accumulateOcclusionFreeCloud = false;
  for (size_t k = 0; k < poses.size(); ++k) {
    pcl::PointCloud<pcl::PointXYZ> vis = oc.extractVisibleSurface(poses[k]);
    if (vis.points.empty()) continue;
    std::printf("accuracy %zu pose %.17g z %.17g\n", k, oc.calcAvgAccuracy(vis, poses[k]), oc.calcAvgAccuracy(vis));
  }
  geometry_msgs::PoseArray pa;
  pa.poses = poses;
  oc.SSMaxMinAccuracy(std::vector<geometry_msgs::PoseArray>(1, pa));
  std::printf("ssmaxmin max %.17g min %.17g set %d\n", oc.maxAccuracyError, oc.minAccuracyError, (int)oc.AccuracyMaxSet);
  // 5. SurfaceCoveragewithAccuracyH: one expansion from waypoint 0 (heuristic.cpp:490-513)
  {
    SSPP::CoverageHeuristicB200 hacc(&oc, 100.0, 1.0, SSPP::SurfaceCoveragewithAccuracyH, true);
    hacc.setMaxMinSensorAccuracy(std::vector<geometry_msgs::PoseArray>(1, pa));
    SSPP::NodeB200 r0 = nodes[0];
    hacc.calculateHeuristic(&r0);
    std::printf("acc_root coverage %.17g f %.17g maxDepth %.17g\n", r0.coverage, r0.f_value, hacc.maxDepth);
    for (size_t i = 1; i < nodes.size(); ++i) {
      SSPP::NodeB200 kid = nodes[i];
      kid.parent = &r0;
      hacc.calculateHeuristic(&kid);
      std::printf("acc_child %zu coverage %.17g f %.17g\n", i, kid.coverage, kid.f_value);
    }
  }
  // 7. VolumetricCoverageH: one expansion from waypoint 0 (heuristic.cpp:581-641)
  {
    SSPP::CoverageHeuristicB200 hv(&oc, 100.0, 1.0, SSPP::VolumetricCoverageH, true);
    hv.setMaxMinSensorAccuracy(std::vector<geometry_msgs::PoseArray>(1, pa));
    SSPP::NodeB200 r0 = nodes[0];
    hv.calculateHeuristic(&r0);
    std::printf("vol_root coverage %.17g f %.17g\n", r0.coverage, r0.f_value);
    for (size_t i = 1; i < nodes.size(); ++i) {
      SSPP::NodeB200 kid = nodes[i];
      kid.parent = &r0;
      hv.calculateHeuristic(&kid);
      std::printf("vol_child %zu coverage %.17g f %.17g\n", i, kid.coverage, kid.f_value);
    }
  }
  // 6. mesh sample filters (heuristic.cpp:300-335): <mesh.txt> = "v x y z" / "f a b c" lines (0-based)
  if (argc > 3) {
    std::vector<float> vtx;
    std::vector<int32_t> tri;
    std::ifstream mf(argv[3]);
    while (std::getline(mf, line)) {
      std::istringstream ss(line);
      char tag;
      ss >> tag;
      if (tag == 'v') { float x, y, z; ss >> x >> y >> z; vtx.push_back(x); vtx.push_back(y); vtx.push_back(z); }
      if (tag == 'f') { int a, b, c; ss >> a >> b >> c; tri.push_back(a); tri.push_back(b); tri.push_back(c); }
    }
    SSPP::CoverageHeuristicB200 hm(&oc, 100.0, 1.0);
    hm.setMesh(vtx, tri);
    std::vector<char> pass, conn;
    hm.filterPositions(poses, 1.0, 4.0, pass);
    std::vector<geometry_msgs::Pose> nxt(poses.begin() + 2, poses.end());
    nxt.push_back(poses[0]); nxt.push_back(poses[1]);
    hm.connectionsFree(poses, nxt, conn);
    std::printf("mesh pass");
    for (size_t k = 0; k < pass.size(); ++k) std::printf(" %d", (int)pass[k]);
    std::printf("\nmesh free");
    for (size_t k = 0; k < conn.size(); ++k) std::printf(" %d", (int)conn[k]);
    std::printf("\nmesh single %d\n", (int)hm.isConnectionConditionSatisfied(poses[0], nxt[0]));
  }
  return 0;
}
