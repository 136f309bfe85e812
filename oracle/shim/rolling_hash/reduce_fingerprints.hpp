// ORACLE SHIM (test infrastructure).  reduce_fps_3tau_lexicographic (call site lce_sss.hpp:82-83):
// one integer per synchronizing position whose suffix order equals the lexicographic order of the
// text suffixes.  Restated in oracle/core/sss_core.hpp::reduce_contexts.
#pragma once
#include <cstddef>
#include <cstdint>
#include <vector>
#include "rolling_hash/string_synchronizing_set.hpp"
namespace lce::rolling_hash {
template <typename t_index, uint64_t t_tau>
std::vector<t_index> reduce_fps_3tau_lexicographic(uint8_t const* text, size_t size,
                                                   sss<t_index, t_tau> const& sync_set) {
  std::vector<uint64_t> S(sync_set.get_sss().begin(), sync_set.get_sss().end());
  std::vector<uint32_t> red = oracle::reduce_contexts(text, size, t_tau, oracle::g_sss_qmax, S);
  return std::vector<t_index>(red.begin(), red.end());
}
}  // namespace lce::rolling_hash
