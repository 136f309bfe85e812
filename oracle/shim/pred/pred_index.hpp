// ORACLE SHIM (test infrastructure).  Successor index over the sorted synchronizing set; contract
// from lce_sss.hpp:68-69,132-134,159-160: successor(x) -> {exists, pos = index into S}.
#pragma once
#include <cstddef>
#include <cstdint>
#include <vector>
namespace lce::pred {
struct result { bool exists; size_t pos; };
template <typename t_value, size_t t_shift, typename t_index>
class pred_index {
 public:
  pred_index() = default;
  pred_index(std::vector<t_value> const& s) : m_s(s.data()), m_n(s.size()) {
    size_t nb = s.empty() ? 1 : (size_t(s.back()) >> t_shift) + 2;
    m_first.assign(nb + 1, (t_index)s.size());
    size_t idx = 0;
    for (size_t b = 0; b <= nb; b++) {
      while (idx < s.size() && (size_t(s[idx]) >> t_shift) < b) idx++;
      m_first[b] = (t_index)idx;
    }
  }
  result successor(size_t x) const {
    size_t b = x >> t_shift;
    if (b + 1 >= m_first.size()) return {false, m_n};
    size_t idx = m_first[b];
    while (idx < m_n && m_s[idx] < x) idx++;
    return {idx < m_n, idx};
  }
 private:
  t_value const* m_s = nullptr;  // the vector's buffer (survives a move of the owning lce_sss)
  size_t m_n = 0;
  std::vector<t_index> m_first;
};
}  // namespace lce::pred
