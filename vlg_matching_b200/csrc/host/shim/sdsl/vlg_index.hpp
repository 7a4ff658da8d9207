// shim/sdsl/vlg_index.hpp — forwarding header: lets a program written against the reference's library surface
// (#include <sdsl/vlg_index.hpp>, namespace sdsl) compile UNMODIFIED against the B200 matcher.
//   g++ -std=c++11 -I vlg_matching_b200/csrc/host/shim /root/reference/examples/vlg_matching.cpp -lvlgb200
// is what csrc/host/Makefile does when the reference is mounted.  Only the names that
// examples/vlg_matching.cpp uses exist: vlg_index<>, construct, construct_im, locate, count, write_structure,
// HTML_FORMAT, int_alphabet_tag and int_vector<> as an initializer-list container (int_vector.hpp:286-288).
#pragma once
#include <initializer_list>

#include "../../vlg_b200.hpp"

namespace vlg_b200 {

template <uint8_t t_width = 0>
struct int_vector : std::vector<uint64_t> {
    int_vector() {}
    int_vector(std::initializer_list<uint64_t> il) : std::vector<uint64_t>(il) {}
    explicit int_vector(size_t n, uint64_t v = 0) : std::vector<uint64_t>(n, v) {}
};

}  // namespace vlg_b200

namespace sdsl = vlg_b200;
