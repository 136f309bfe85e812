// ORACLE SHIM (test infrastructure).  ips4o::sort / ips4o::parallel::sort (call sites
// approximate/common.cpp:39, sample_index.hpp:324,350, construction.cpp:138) mapped to the
// standard sorts.  The sequential one is mapped to std::stable_sort: greedy_phrase_selection
// (approximate/common.cpp:39-41) sorts phrases by (beg asc, end desc), and phrases with equal (beg, end)
// but different src are ties whose order decides which src is emitted.  ips4o leaves that order
// unspecified (parity unpinned there); the canonical choice here is insertion order (LNF phrases
// before LPF phrases, each in text order), which a stable device radix sort reproduces.
#pragma once
#include <algorithm>
#include <parallel/algorithm>
namespace ips4o {
template <typename It> void sort(It b, It e) { std::sort(b, e); }
template <typename It, typename C> void sort(It b, It e, C c) { std::stable_sort(b, e, c); }
namespace parallel {
template <typename It> void sort(It b, It e) { __gnu_parallel::sort(b, e); }
template <typename It, typename C> void sort(It b, It e, C c) { __gnu_parallel::sort(b, e, c); }
}  // namespace parallel
}  // namespace ips4o
