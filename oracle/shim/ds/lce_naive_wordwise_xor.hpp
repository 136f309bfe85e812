// ORACLE SHIM (test infrastructure).  Word-wise naive LCE; contract from the call sites
// lce_sss.hpp:141,152 (lce_lr(text,end,l,r): bounded by end-r), lce_classic_for_sss.hpp:104-110
// (lce_uneq(text,size,i,j)), sample_index.hpp:43 / tests/test_sample_index.cpp:33,123 (ctor from a
// container, member lce(i,j)).
#pragma once
#include <cstddef>
#include <cstdint>
#include "../../core/sss_core.hpp"
namespace lce::ds {
template <typename t_char_type = uint8_t>
class lce_naive_wordwise_xor {
 public:
  typedef t_char_type char_type;
  lce_naive_wordwise_xor() : m_text(nullptr), m_size(0) {}
  lce_naive_wordwise_xor(char_type const* text, size_t size) : m_text(text), m_size(size) {}
  template <typename C>
  lce_naive_wordwise_xor(C const& c) : m_text(c.data()), m_size(c.size()) {}
  size_t lce(size_t i, size_t j) const {
    if (i == j) return m_size - i;
    return lce_uneq(m_text, m_size, i, j);
  }
  size_t lce_uneq(size_t i, size_t j) const { return lce_uneq(m_text, m_size, i, j); }
  size_t lce_lr(size_t l, size_t r) const { return lce_lr(m_text, m_size, l, r); }
  static size_t lce_uneq(char_type const* text, size_t size, size_t i, size_t j) {
    return lce_lr(text, size, i < j ? i : j, i < j ? j : i);
  }
  static size_t lce_lr(char_type const* text, size_t end, size_t l, size_t r) {
    return oracle::naive_lce_lr(reinterpret_cast<const uint8_t*>(text), end, l, r);
  }
  char_type operator[](size_t i) const { return m_text[i]; }
  size_t size() const { return m_size; }
 private:
  char_type const* m_text;
  size_t m_size;
};
}  // namespace lce::ds
