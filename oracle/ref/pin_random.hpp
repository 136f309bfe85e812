// ORACLE (test infrastructure).  Force-included (-include) before the reference headers so that
// every `std::random_device` the reference seeds a generator from returns a fixed seed.  The
// reference draws its rolling-hash bases that way (rolling_hash.hpp:127-130,
// rabin_karp_substring.hpp:89-96); bit-exact comparison needs them pinned.
#pragma once
#include <algorithm>
#include <array>
#include <chrono>
#include <filesystem>
#include <fstream>
#include <functional>
#include <iostream>
#include <parallel/algorithm>
#include <random>
#include <string>
#include <vector>
#ifndef ORACLE_PINNED_SEED
#define ORACLE_PINNED_SEED 0x5eed1234u
#endif
namespace std {
struct oracle_pinned_random_device {
    using result_type = unsigned int;
    oracle_pinned_random_device() {}
    result_type operator()() { return ORACLE_PINNED_SEED; }
    static constexpr result_type min() { return 0; }
    static constexpr result_type max() { return 0xFFFFFFFFu; }
    double entropy() const { return 0.0; }
};
}  // namespace std
#define random_device oracle_pinned_random_device
