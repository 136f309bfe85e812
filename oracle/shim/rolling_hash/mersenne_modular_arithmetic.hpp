// ORACLE SHIM (test infrastructure).  Restates the helpers rolling_hash.hpp:18,34,77,135-152 needs
// from the missing lce submodule: arithmetic modulo a Mersenne prime 2^e-1 on unsigned __int128.
#pragma once
#include <cstdint>
namespace lce::rolling_hash::mersenne {
__extension__ typedef unsigned __int128 uint128_t;
constexpr unsigned bit_width(uint128_t x) {
    unsigned w = 0;
    while (x) { w++; x >>= 1; }
    return w;
}
// canonical residue of x modulo P = 2^e - 1 (x may use all 128 bits)
template <typename T, T P>
constexpr T mod(T x) {
    constexpr unsigned e = bit_width(P);
    x = (x & P) + (x >> e);
    x = (x & P) + (x >> e);
    return x >= P ? x - P : x;
}
template <typename T, T P>
constexpr T add_mod(T a, T b) { return mod<T, P>(a + b); }
template <typename T, T P>
constexpr T additive_inverse_mod(T a) { a = mod<T, P>(a); return a == 0 ? 0 : P - a; }
}  // namespace lce::rolling_hash::mersenne
