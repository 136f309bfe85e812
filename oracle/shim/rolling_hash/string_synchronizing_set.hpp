// ORACLE SHIM (test infrastructure).  lce::rolling_hash::sss<idx,tau>: the tau-synchronizing set.
// Shape required by the call sites lce_sss.hpp:53,58-59,81,124,209,213,234.
// The construction itself is oracle/core/sss_core.hpp (canonical definition shared with the
// B200 kernels; PARITY UNPINNED against the missing upstream lce, see that header).
#pragma once
#include <cstddef>
#include <cstdint>
#include <vector>
#include "../../core/sss_core.hpp"
namespace oracle {
// hash parameters of the SSS, pinned by the driver before a run (the reference draws them at random)
inline uint32_t g_sss_base = DEFAULT_SSS_BASE;
inline uint32_t g_sss_qmax = DEFAULT_QMAX;
}  // namespace oracle
namespace lce::rolling_hash {
template <typename t_index = uint32_t, uint64_t t_tau = 1024>
class sss {
 public:
  sss() = default;
  template <typename t_char>
  sss(t_char const* text_, size_t size, bool /*store_fps*/) {
    static_assert(sizeof(t_char) == 1);
    uint8_t const* text = reinterpret_cast<uint8_t const*>(text_);
    oracle::sss_result r = oracle::build_sss(text, size, t_tau, oracle::g_sss_base, oracle::g_sss_qmax);
    m_sss.assign(r.S.begin(), r.S.end());
    m_num_q = r.num_q;
  }
  std::vector<t_index> const& get_sss() const { return m_sss; }
  size_t size() const { return m_sss.size(); }
  size_t num_runs() const { return m_num_q; }
  bool has_runs() const { return m_num_q != 0; }
 private:
  std::vector<t_index> m_sss;
  size_t m_num_q = 0;
};
}  // namespace lce::rolling_hash
