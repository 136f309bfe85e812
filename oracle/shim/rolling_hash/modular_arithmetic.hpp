// ORACLE SHIM (test infrastructure).  pow_mod as used at rolling_hash.hpp:135-136.
#pragma once
#include <cstdint>
#include "rolling_hash/mersenne_modular_arithmetic.hpp"
namespace lce::rolling_hash::modular {
__extension__ typedef unsigned __int128 uint128_t;
// b^e mod p for p < 2^107, b < p: products need up to 214 bits, so multiply by shift-and-add.
template <typename T>
constexpr T mul_mod(T a, T b, T p) {
    T r = 0;
    a %= p;
    while (b) {
        if (b & 1) { r += a; if (r >= p) r -= p; }
        a += a; if (a >= p) a -= p;
        b >>= 1;
    }
    return r;
}
template <typename T>
constexpr T pow_mod(T b, T e, T p) {
    T r = 1 % p;
    b %= p;
    while (e) {
        if (e & 1) r = mul_mod<T>(r, b, p);
        b = mul_mod<T>(b, b, p);
        e >>= 1;
    }
    return r;
}
}  // namespace lce::rolling_hash::modular
