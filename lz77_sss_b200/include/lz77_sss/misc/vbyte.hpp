// vbyte.hpp — variable-byte integers of ssszip's gapped factor format (row f3 of SURVEY.md §8).
//
// Same function templates as /root/reference/include/lz77_sss/misc/vbyte.hpp:64-101 (encode_vbyte / decode_vbyte, from
// pdinklag/alz) and the same bytes: 7 value bits per byte, least significant group first, the top bit of a byte says
// that another byte follows.  cli/ssszip.cpp:120-180 writes the skip_phrases stream of factorize_approximate with it.
#pragma once
#include <cstdint>
#include <istream>
#include <ostream>

// writes x, returns the number of bytes written
template <typename pos_t>
inline pos_t encode_vbyte(std::ostream& out, pos_t x) {
    pos_t bytes = 0;
    for (;;) {
        const uint8_t group = (uint8_t)(x & 0x7F);
        x >>= 7;
        out.put((char)(x ? group | 0x80 : group));
        ++bytes;
        if (!x) return bytes;
    }
}

// reads one value; a stream that ends inside a value yields the groups read so far (like the reference)
template <typename pos_t>
inline pos_t decode_vbyte(std::istream& in) {
    pos_t x = 0;
    for (unsigned shift = 0;; shift += 7) {
        const int c = in.get();
        if (!in.good()) return x;
        x |= (pos_t)(c & 0x7F) << shift;
        if (!(c & 0x80)) return x;
    }
}
