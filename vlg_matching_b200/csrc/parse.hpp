// parse.hpp — query-string parsers of the two reference surfaces (host only, no CUDA).
//
//   library   syntax  P1.{a,b}?P2...   gapped_pattern_query ctor, include/sdsl/vlg_index.hpp:54-105
//   benchmark syntax  P1.{a,b}P2...    gapped_pattern ctor, benchmark/gapped-matching/include/utils.hpp:19-71
//
// Behaviour kept: the string is cut at every ".{"; the bytes in between are literal subpattern bytes
// (no escapes); the gap is "{a,b}" with a and b read the way std::stoull reads them (leading blanks,
// optional sign, trailing garbage ignored); a > b is an error; the library syntax demands '?' right
// after '}' and checks it AFTER the min/max check (vlg_index.hpp:92-99).
#pragma once
#include <cctype>
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/vlgb200.h"

namespace vlgb {

struct ParsedQuery {
    std::vector<std::string> subs;
    std::vector<uint64_t> gap_min, gap_max;  // one per gap (subs.size() - 1)
};

inline bool parse_stoull(const char* s, size_t len, uint64_t* out)
{
    size_t i = 0;
    while (i < len && isspace((unsigned char)s[i])) i++;
    bool neg = false;
    if (i < len && (s[i] == '+' || s[i] == '-')) {
        neg = s[i] == '-';
        i++;
    }
    if (i >= len || !isdigit((unsigned char)s[i])) return false;  // std::invalid_argument
    uint64_t v = 0;
    for (; i < len && isdigit((unsigned char)s[i]); i++) {
        uint64_t d = (uint64_t)(s[i] - '0');
        if (v > (UINT64_MAX - d) / 10) return false;  // std::out_of_range
        v = v * 10 + d;
    }
    *out = neg ? (uint64_t)(0 - v) : v;
    return true;
}

// Returns VLGB_OK or VLGB_ERR_NUMBER / _MINMAX / _LAZY.
inline int parse_query(const char* s, size_t len, int syntax, ParsedQuery& q)
{
    q.subs.clear();
    q.gap_min.clear();
    q.gap_max.clear();
    size_t start = 0;
    for (;;) {
        size_t gap = std::string::npos;
        for (size_t i = start; i + 1 < len; i++)
            if (s[i] == '.' && s[i + 1] == '{') {
                gap = i;
                break;
            }
        if (gap == std::string::npos) break;
        q.subs.emplace_back(s + start, gap - start);
        size_t close = gap;
        while (close < len && s[close] != '}') close++;
        if (close == len) return VLGB_ERR_NUMBER;  // unterminated gap: fenced as a number error
        size_t comma = gap;
        while (comma < close && s[comma] != ',') comma++;
        if (comma == close) return VLGB_ERR_NUMBER;  // no ',' -> the reference's stoull throws
        uint64_t a, b;
        if (!parse_stoull(s + gap + 2, comma - (gap + 2), &a)) return VLGB_ERR_NUMBER;
        if (!parse_stoull(s + comma + 1, close - (comma + 1), &b)) return VLGB_ERR_NUMBER;
        if (a > b) return VLGB_ERR_MINMAX;
        q.gap_min.push_back(a);
        q.gap_max.push_back(b);
        if (syntax == VLGB_SYNTAX_LIBRARY) {
            if (close + 1 >= len || s[close + 1] != '?') return VLGB_ERR_LAZY;
            start = close + 2;
        } else {
            start = close + 1;
        }
    }
    q.subs.emplace_back(s + start, len - start);
    return VLGB_OK;
}

// A one-query (or n-query) batch that owns its arrays; `pub` is what crosses the ABI.
struct OwnedBatch {
    vlgb_batch pub;
    std::vector<uint32_t> sub_off;
    std::vector<uint64_t> pat_off;
    std::vector<uint8_t> bytes;
    std::vector<uint64_t> gmin, gmax;
    void add(const ParsedQuery& q)
    {
        if (sub_off.empty()) sub_off.push_back(0);
        if (pat_off.empty()) pat_off.push_back(0);
        for (size_t i = 0; i < q.subs.size(); i++) {
            bytes.insert(bytes.end(), q.subs[i].begin(), q.subs[i].end());
            pat_off.push_back(bytes.size());
            gmin.push_back(i < q.gap_min.size() ? q.gap_min[i] : 0);
            gmax.push_back(i < q.gap_max.size() ? q.gap_max[i] : 0);
        }
        sub_off.push_back((uint32_t)(pat_off.size() - 1));
    }
    void seal()
    {
        if (bytes.empty()) bytes.push_back(0);
        pub.n_queries = (uint32_t)(sub_off.size() - 1);
        pub.sub_off = sub_off.data();
        pub.pat_off = pat_off.data();
        pub.pat_bytes = bytes.data();
        pub.gap_min = gmin.data();
        pub.gap_max = gmax.data();
    }
};

}  // namespace vlgb
