// Same program as the reference's examples/lz77_sss_example.cpp:28-53, built against the B200 header
// (factorize_exact is not built yet, so the approximation-ratio part is left out).
//   g++ -std=c++20 -Iinclude -Ilz77_sss_b200/include examples/lz77_sss_example.cpp -Llz77_sss_b200 -llz77sss_b200
#include <lz77_sss/lz77_sss.hpp>
#include <lz77_sss/data_structures/rabin_karp_substring.hpp>

#include <iostream>
#include <random>
#include <string>
#include <vector>

int main(int argc, char** argv) {
    // a repetitive random string (the reference uses random_repetitive_string(4'000, 1'000'000))
    std::mt19937_64 gen(argc > 1 ? std::stoull(argv[1]) : 42);
    std::string input;
    while (input.size() < 1'000'000) {
        if (input.size() < 4000 || gen() % 4 == 0) {
            input.push_back((char)('a' + gen() % 26));
        } else {
            size_t len = 1 + gen() % 5000, src = gen() % input.size();
            for (size_t i = 0; i < len; i++) input.push_back(input[src + i]);
        }
    }
    std::cout << "input length: " << input.size() << std::endl;

    std::vector<lz77_sss<>::factor> factorization;
    lz77_sss<>::factorize_approximate<>(input.data(), (uint32_t)input.size(), [&](auto f) { factorization.emplace_back(f); });
    std::cout << "num. of factors: " << factorization.size() << std::endl;
    std::cout << "input length / num. of factors: " << input.size() / (double)factorization.size() << std::endl;

    std::string decoded(input.size(), '\0');
    lz77_sss<>::decode(factorization.begin(), decoded.data(), (uint32_t)input.size());
    std::cout << (decoded == input ? "decode ok" : "DECODE MISMATCH") << std::endl;

    // the fingerprint structure of the exact path (the checks of the reference's tests/test_rabin_karp_substring.cpp:53,74,91)
    const uint32_t window = 300;
    rabin_karp_substring<31, uint32_t> rks(input.data(), (uint32_t)input.size(), 16, window);
    bool rks_ok = true;
    for (int it = 0; it < 200; it++) {
        const uint32_t pos = gen() % (input.size() - 3000), l1 = 1 + gen() % 1000, l2 = 1 + gen() % 1000;
        rks_ok &= rks.concat(rks.substring(pos, l1), rks.substring(pos + l1, l2), l2) == rks.substring(pos, l1 + l2);
        rks_ok &= rks.substring(pos, l1) == rks.substring_naive(pos, l1);
        uint32_t fp = rks.substring(pos, window);
        for (uint32_t d = 0; d < 50; d++) fp = rks.roll(fp, input[pos + d], input[pos + d + window]);
        rks_ok &= fp == rks.substring(pos + 50, window);
    }
    auto batch = rks.substring_batch({0, 17, 999}, {100, 2000, 31});
    rks_ok &= batch[0] == rks.substring(0, 100) && batch[1] == rks.substring(17, 2000) && batch[2] == rks.substring(999, 31);
    std::cout << (rks_ok ? "rabin_karp_substring ok" : "RABIN KARP MISMATCH") << std::endl;
    return decoded == input && rks_ok ? 0 : 1;
}
