// ref_sspp_driver.cpp -- the REFERENCE'S OWN planners (sspp/src/astar.cpp, pathplanner.cpp, llist.cpp, node.cpp,
// searchspace.cpp, sensors.cpp ...: the objects oracle/Makefile compiles from /root/reference) linked against THIS
// repository's drop-in plugin class SSPP::CoveragePathPlanningHeuristic (include/cscpp/coverage_path_planning_heuristic_b200.h)
// and run on the GPU.  Nothing of the reference is modified: Astar::astarSearch calls heuristic->calculateHeuristic one
// child at a time (astar.cpp:318), PathPlanner::generateRegularGrid calls isFilteringConditionSatisfied one candidate at
// a time (pathplanner.cpp:160), exactly as with the reference's own heuristic.  The wiring mirrors the demo
// (cscpp/src/coverage_heuristic_test.cpp:73-137, 153-188).
//
// Built by `make -C oracle driver` into oracle/_ref/ref_sspp_driver (the reference is only present in the build
// container; the binary travels to the GPU box).  Output: lines tests/test_boundary.py parses.
//
//   ref_sspp_driver grid    <pcd> <obj> sx sy sz  gx gy gz  res          one filtered sampling stage
//   ref_sspp_driver dynamic <pcd> <obj> sx sy sz  gx gy gz  res dec      the whole adaptive sampling loop
//   ref_sspp_driver astar   <pcd> <obj> <robot.txt> <cam0.txt> <cam1.txt> htype target tol batch   A* over a loaded space
//   ref_sspp_driver infogain <pcd> <obj> <robot.txt> <cam0.txt> <cam1.txt> i0 i1 i2 ...   InfoGainVolumetricH (intended
//                            form) along the chain of search-space rows i0 -> i1 -> ...: every child's parent is a COPY of
//                            the scored node (Node::Node(Node*), as astar.cpp:262 makes it), not the node itself
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include <cscpp/coverage_path_planning_heuristic_b200.h>
#include <sspp/pathplanner.h>
#include <sspp/sensors.h>

using namespace SSPP;

static void print_space(PathPlanner* planner) {
  geometry_msgs::PoseArray robot;
  std::vector<geometry_msgs::PoseArray> sensors;
  planner->getRobotSensorPoses(robot, sensors);
  std::printf("\nSPACE %zu %zu\n", robot.poses.size(), planner->getConnections().size() / 2);
  for (size_t j = 0; j < robot.poses.size(); ++j) {
    const geometry_msgs::Pose& p = robot.poses[j];
    tf::Quaternion qt(p.orientation.x, p.orientation.y, p.orientation.z, p.orientation.w);
    std::printf("ROW %.17g %.17g %.17g %.17g\n", p.position.x, p.position.y, p.position.z, tf::getYaw(qt));
  }
}

int main(int argc, char** argv) {
  if (argc < 4) { std::fprintf(stderr, "usage: see the header of tests/cpp/ref_sspp_driver.cpp\n"); return 2; }
  const std::string mode = argv[1], pcd = argv[2], obj = argv[3];
  ros::NodeHandle nh;
  geometry_msgs::Point robotCenter;
  robotCenter.x = -0.3f; robotCenter.y = 0.0f;
  Robot* robot = new Robot("Robot", 0.9, 0.5, 0.987, robotCenter);                            // :85-91
  std::vector<Sensors> sensors;
  sensors.push_back(Sensors(58, 45, 0.255, 0.7, 6.0, 640, 480, Eigen::Vector3f(0, 0.022f, 0.065f), Eigen::Vector3f(0, -0.349f, 0)));
  sensors.push_back(Sensors(58, 45, 0.255, 0.7, 6.0, 640, 480, Eigen::Vector3f(0, 0.022f, -0.065f), Eigen::Vector3f(0, 0.349f, 0)));
  PathPlanner* planner = new PathPlanner(nh, robot, 2.5, 0, sensors);                           // :102, regGridConRad 2.5
  planner->setDebugDelay(0.0);
  try {
    if (mode == "grid" || mode == "dynamic") {
      if (argc < 11) return 2;
      CoveragePathPlanningHeuristic heuristic(nh, obj, pcd, false, false, SurfaceCoverageH);
      heuristic.setCoverageTarget(5.0);
      heuristic.setCoverageTolerance(1.0);
      planner->setHeuristicFucntion(&heuristic);                                               // :114
      geometry_msgs::Pose gs;
      gs.position.x = std::atof(argv[4]); gs.position.y = std::atof(argv[5]); gs.position.z = std::atof(argv[6]);
      geometry_msgs::Vector3 sz;
      sz.x = std::atof(argv[7]); sz.y = std::atof(argv[8]); sz.z = std::atof(argv[9]);
      const double res = std::atof(argv[10]);
      if (mode == "grid") planner->generateRegularGrid(gs, sz, (float)res, true, 45, true, true);  // pathplanner.cpp:736
      else planner->dynamicNodesGenerationAndConnection(gs, sz, res, std::atof(argv[11]));          // :137
      print_space(planner);
      std::printf("LAUNCHES %ld\n", heuristic.launchesScored());
    } else if (mode == "astar") {
      if (argc < 11) return 2;
      const int htype = std::atoi(argv[7]);
      CoveragePathPlanningHeuristic heuristic(nh, obj, pcd, false, false, htype);
      heuristic.setCoverageTarget(std::atof(argv[8]));
      heuristic.setCoverageTolerance(std::atof(argv[9]));
      planner->setHeuristicFucntion(&heuristic);
      if (std::atoi(argv[10])) heuristic.batchSiblingsFrom(&planner->childList);  // the one added line of the batcher
      planner->loadRegularGrid(argv[4], argv[5], argv[6]);                        // coverage_heuristic_test.cpp:126
      planner->connectNodes();                                                    // :127
      if (htype == InfoGainVolumetricH) heuristic.allowIntendedInfoGain();  // the reference's own block is undefined here
      if (htype == SurfaceCoveragewithAccuracyH || htype == VolumetricCoverageH || htype == InfoGainVolumetricH) {
        geometry_msgs::PoseArray rp;
        std::vector<geometry_msgs::PoseArray> sp;
        planner->getRobotSensorPoses(rp, sp);
        heuristic.setMaxMinSensorAccuracy(sp);                                    // :158
        std::printf("\nACCURACY %.17g %.17g\n", heuristic.occlusionCulling()->maxAccuracyError,
                    heuristic.occlusionCulling()->minAccuracyError);
      }
      Pose start(3.0, -34.5, 9, DTOR(0.0));                                       // :83
      Node* path = planner->startSearch(start);                                   // :188
      size_t n = 0;
      for (Node* p = path; p != NULL; p = p->next) ++n;
      std::printf("\nPATH %zu %zu\n", n, planner->getConnections().size() / 2);
      for (Node* p = path; p != NULL; p = p->next)
        std::printf("NODE %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g\n", p->pose.p.position.x,
                    p->pose.p.position.y, p->pose.p.position.z, p->pose.p.orientation.x, p->pose.p.orientation.y,
                    p->pose.p.orientation.z, p->pose.p.orientation.w, p->coverage, p->f_value, p->distance);
      std::printf("LAUNCHES %ld NODES %ld\n", heuristic.launchesScored(), heuristic.nodesScored());
    } else if (mode == "infogain") {
      if (argc < 9) return 2;
      CoveragePathPlanningHeuristic heuristic(nh, obj, pcd, false, false, InfoGainVolumetricH);
      {  // the type is refused, with the reason, until the intended form is asked for
        Node probe;
        try { heuristic.calculateHeuristic(&probe); std::printf("\nREFUSED no\n"); }
        catch (const std::exception& e) { std::printf("\nREFUSED %s\n", e.what()); }
      }
      heuristic.allowIntendedInfoGain();
      planner->setHeuristicFucntion(&heuristic);
      planner->loadRegularGrid(argv[4], argv[5], argv[6]);
      geometry_msgs::PoseArray rp;
      std::vector<geometry_msgs::PoseArray> sp;
      planner->getRobotSensorPoses(rp, sp);
      heuristic.setMaxMinSensorAccuracy(sp);
      std::printf("\nACCURACY %.17g %.17g\n", heuristic.occlusionCulling()->maxAccuracyError,
                  heuristic.occlusionCulling()->minAccuracyError);
      // the chain's poses by FILE row (the planner keeps its search space newest first)
      std::vector<std::vector<geometry_msgs::Pose> > file_rows(3);
      for (int f = 0; f < 3; ++f) {
        FILE* fp = std::fopen(argv[4 + f], "r");
        if (!fp) return 2;
        double v[7];
        while (std::fscanf(fp, "%lf %lf %lf %lf %lf %lf %lf", &v[0], &v[1], &v[2], &v[3], &v[4], &v[5], &v[6]) == 7) {
          geometry_msgs::Pose q;
          q.position.x = v[0]; q.position.y = v[1]; q.position.z = v[2];
          q.orientation.x = v[3]; q.orientation.y = v[4]; q.orientation.z = v[5]; q.orientation.w = v[6];
          file_rows[f].push_back(q);
        }
        std::fclose(fp);
      }
      Node* prev_copy = NULL;
      for (int a = 7; a < argc; ++a) {
        const size_t row = (size_t)std::atoi(argv[a]);
        Node* n = new Node();
        n->id = a - 7;
        n->depth = a - 7;
        n->parent = prev_copy;
        n->pose.p = file_rows[0][row];
        for (size_t s = 1; s < 3; ++s) { Pose q; q.p = file_rows[s][row]; n->senPoses.push_back(q); }
        heuristic.calculateHeuristic(n);
        std::printf("IG %.17g %.17g %.17g %.17g %.17g %d\n", n->totalEntroby, n->coveredVolume, n->coverage, n->f_value,
                    n->distance, n->coveredVoxelsNum);
        prev_copy = new Node(n);  // what the children of the next expansion see as their parent (astar.cpp:262)
      }
      std::printf("LAUNCHES %ld NODES %ld\n", heuristic.launchesScored(), heuristic.nodesScored());
    } else {
      return 2;
    }
  } catch (const std::exception& e) {
    std::fprintf(stderr, "ref_sspp_driver: %s\n", e.what());
    return 1;
  }
  return 0;
}
