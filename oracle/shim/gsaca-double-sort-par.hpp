// ORACLE SHIM (test infrastructure).  gsaca_for_lce(text, sa, n): suffix array of a byte or
// integer string (call sites lce_classic_for_sss.hpp:52, lce_classic.hpp:51).  Prefix doubling;
// requires what the caller guarantees here: the reduced string ends in a unique symbol.
#pragma once
#include <cstddef>
#include <cstdint>
#include "../core/sss_core.hpp"
#include "ips4o.hpp"  // upstream lce pulls ips4o in transitively; the reference relies on that
template <typename sym_t>
inline void gsaca_for_lce(sym_t const* text, uint32_t* sa, size_t n) {
  oracle::suffix_sort_ints(text, sa, n);
}
