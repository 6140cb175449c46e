// Stand-in for boost/shared_ptr.hpp: the reference only names the type (PCL 1.7 Ptr typedefs).
#pragma once
#include <memory>
namespace boost {
template <class T> using shared_ptr = std::shared_ptr<T>;
template <class T, class... A> inline std::shared_ptr<T> make_shared(A&&... a) { return std::make_shared<T>(static_cast<A&&>(a)...); }
}  // namespace boost
