// ORACLE SHIM (test infrastructure).  Range-minimum over the sparse LCP array; contract from
// lce_classic_for_sss.hpp:133,153-155: rmq_shifted(a,b) = arg-min of lcp over (min(a,b), max(a,b)].
#pragma once
#include <cstddef>
#include <cstdint>
#include <vector>
#include "../../core/sss_core.hpp"
namespace lce::rmq {
template <typename t_value>
class rmq_n {
 public:
  rmq_n() = default;
  rmq_n(std::vector<t_value> const& v) { m_rmq.build(v.data(), v.size()); }
  size_t rmq_shifted(size_t a, size_t b) const {
    size_t lo = (a < b ? a : b) + 1, hi = a < b ? b : a;
    return m_rmq.argmin(lo, hi);
  }
 private:
  oracle::block_rmq<t_value> m_rmq;
};
}  // namespace lce::rmq
