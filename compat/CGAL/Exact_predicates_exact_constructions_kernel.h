// Stand-in for CGAL's exact kernel: sspp/include/sspp/node.h:29-40 only names Triangle_3<exactKernel> for a member
// (Node::surfaceTriangles) that the coverage path never fills.
#pragma once
namespace CGAL {
struct Exact_predicates_exact_constructions_kernel {};
}  // namespace CGAL
