// Stand-in: see Exact_predicates_exact_constructions_kernel.h.
#pragma once
namespace CGAL {
template <class K> struct Triangle_3 {};
}  // namespace CGAL
