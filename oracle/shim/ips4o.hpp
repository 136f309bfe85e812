// ORACLE SHIM (test infrastructure).  ips4o::sort / ips4o::parallel::sort (call sites
// approximate/common.cpp:39, sample_index.hpp:324,350, construction.cpp:138) mapped to the
// standard (parallel-mode) sorts.  Neither is stable; callers use strict weak orders whose ties
// are only between interchangeable elements on the 3-aprx path.
#pragma once
#include <algorithm>
#include <parallel/algorithm>
namespace ips4o {
template <typename It> void sort(It b, It e) { std::sort(b, e); }
template <typename It, typename C> void sort(It b, It e, C c) { std::sort(b, e, c); }
namespace parallel {
template <typename It> void sort(It b, It e) { __gnu_parallel::sort(b, e); }
template <typename It, typename C> void sort(It b, It e, C c) { __gnu_parallel::sort(b, e, c); }
}  // namespace parallel
}  // namespace ips4o
