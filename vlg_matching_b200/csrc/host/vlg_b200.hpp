// vlg_b200.hpp — the reference's matcher API on top of the C ABI of libvlgb200.so (include/vlgb200.h).
//
// Library surface (what examples/vlg_matching.cpp uses, /root/reference/include/sdsl/vlg_index.hpp):
//     vlg_b200::vlg_index<> idx;                         vlg_index<alphabet_tag>           vlg_index.hpp:116-199
//     construct(idx, file, 1); construct_im(idx, s, 1);  construct / construct_im          vlg_index.hpp:375-392, construct.hpp:66-73
//     for (auto it = locate(idx, q).begin(); ...)        locate -> container<vlg_iterator> vlg_index.hpp:395-401
//         *it, it[i], it.size()                          vlg_iterator                      vlg_index.hpp:331-346
//     count(idx, q)                                      count                             vlg_index.hpp:404-411
//     write_structure<HTML_FORMAT>(idx, out)             structure_tree report             examples/vlg_matching.cpp:36-37
// Error behaviour: std::runtime_error for a bad gap description (same texts as vlg_index.hpp:92-99),
// std::logic_error for a zero symbol in the text (construct.hpp:35-45).  Everything is computed by the
// CUDA kernels behind the ABI; this header only packs queries and unpacks results.  Header-only, C++11.
#pragma once
#include <unistd.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <iterator>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <utility>
#include <vector>

#include "../../../include/vlgb200.h"

namespace vlg_b200 {

struct byte_alphabet_tag { static const uint8_t WIDTH = 8; };
struct int_alphabet_tag { static const uint8_t WIDTH = 0; };
struct HTML_FORMAT {};
struct JSON_FORMAT {};

namespace detail {

inline void throw_status(int rc)
{
    const std::string msg = vlgb_last_error();
    switch (rc) {
        case VLGB_ERR_LAZY:
        case VLGB_ERR_MINMAX:
        case VLGB_ERR_NUMBER: throw std::runtime_error(msg);  // vlg_index.hpp:92-99 (stoull: invalid_argument)
        case VLGB_ERR_ZERO_SYMBOL: throw std::logic_error(msg);  // construct.hpp:35-45
        case VLGB_ERR_NOMEM: throw std::bad_alloc();
        default: throw std::runtime_error("vlgb200 (" + std::to_string(rc) + "): " + msg);
    }
}
inline void check(int rc)
{
    if (rc != VLGB_OK) throw_status(rc);
}

struct index_handle {
    vlgb_index* h;
    explicit index_handle(vlgb_index* p) : h(p) {}
    ~index_handle() { vlgb_index_free(h); }
    index_handle(const index_handle&) = delete;
    index_handle& operator=(const index_handle&) = delete;
};
struct result_handle {
    vlgb_result* h;
    explicit result_handle(vlgb_result* p) : h(p) {}
    ~result_handle() { vlgb_result_free(h); }
    result_handle(const result_handle&) = delete;
    result_handle& operator=(const result_handle&) = delete;
};

// CSR arrays of a batch (what vlgb_batch points to)
struct packed_batch {
    std::vector<uint32_t> sub_off{0};
    std::vector<uint64_t> pat_off{0};
    std::vector<uint8_t> bytes;
    std::vector<uint64_t> gmin, gmax;
    void add(const std::vector<std::string>& subs, const std::vector<std::pair<uint64_t, uint64_t>>& raw_gaps)
    {
        for (size_t i = 0; i < subs.size(); i++) {
            bytes.insert(bytes.end(), subs[i].begin(), subs[i].end());
            pat_off.push_back(bytes.size());
            gmin.push_back(i < raw_gaps.size() ? raw_gaps[i].first : 0);
            gmax.push_back(i < raw_gaps.size() ? raw_gaps[i].second : 0);
        }
        sub_off.push_back((uint32_t)(pat_off.size() - 1));
    }
    vlgb_batch view()
    {
        if (bytes.empty()) bytes.push_back(0);
        if (gmin.empty()) { gmin.push_back(0); gmax.push_back(0); }
        vlgb_batch b;
        b.n_queries = (uint32_t)(sub_off.size() - 1);
        b.sub_off = sub_off.data();
        b.pat_off = pat_off.data();
        b.pat_bytes = bytes.data();
        b.gap_min = gmin.data();
        b.gap_max = gmax.data();
        return b;
    }
};

// split a query string with the library's own parser (vlgb_parse_query): literal subpattern strings + raw gaps
inline void split_query(const std::string& raw, int syntax, std::vector<std::string>& subs,
                        std::vector<std::pair<uint64_t, uint64_t>>& raw_gaps)
{
    vlgb_batch* b = nullptr;
    check(vlgb_parse_query(raw.data(), raw.size(), syntax, &b));
    const uint32_t k = b->sub_off[1];
    subs.clear();
    raw_gaps.clear();
    for (uint32_t i = 0; i < k; i++) {
        subs.emplace_back(reinterpret_cast<const char*>(b->pat_bytes) + b->pat_off[i], (size_t)(b->pat_off[i + 1] - b->pat_off[i]));
        if (i + 1 < k) raw_gaps.emplace_back(b->gap_min[i], b->gap_max[i]);
    }
    vlgb_batch_free(b);
}

// "97 98 114" -> {97, 98, 114}: `while (stream >> symbol)` of vlg_index.hpp:63-69
inline std::vector<uint64_t> int_symbols(const std::string& s)
{
    std::istringstream in(s);
    std::vector<uint64_t> out;
    uint64_t v;
    while (in >> v) out.push_back(v);
    return out;
}

}  // namespace detail

// ---------------------------------------------------------------------------------------------
// gapped_pattern_query  (vlg_index.hpp:43-106)
// ---------------------------------------------------------------------------------------------
template <typename alphabet_tag>
struct gapped_pattern_query {
    typedef typename std::conditional<alphabet_tag::WIDTH == 8, std::string, std::vector<uint64_t>>::type string_type;

    std::vector<string_type> subpatterns;
    std::vector<std::pair<uint64_t, uint64_t>> gaps;      // start-to-start bounds (min + |P_i|, max + |P_i|), vlg_index.hpp:95
    std::vector<std::pair<uint64_t, uint64_t>> raw_gaps;  // {a,b} as written (what the C ABI takes)

    gapped_pattern_query() {}
    gapped_pattern_query(const char* raw_regexp) { parse(raw_regexp); }
    gapped_pattern_query(const std::string& raw_regexp) { parse(raw_regexp); }

private:
    static void convert(const std::string& s, std::string& out) { out = s; }
    static void convert(const std::string& s, std::vector<uint64_t>& out) { out = detail::int_symbols(s); }
    void parse(const std::string& raw)
    {
        std::vector<std::string> subs;
        detail::split_query(raw, VLGB_SYNTAX_LIBRARY, subs, raw_gaps);
        subpatterns.resize(subs.size());
        for (size_t i = 0; i < subs.size(); i++) convert(subs[i], subpatterns[i]);
        for (size_t i = 0; i < raw_gaps.size(); i++)
            gaps.emplace_back(raw_gaps[i].first + subpatterns[i].size(), raw_gaps[i].second + subpatterns[i].size());
    }
};

// ---------------------------------------------------------------------------------------------
// vlg_index  (vlg_index.hpp:116-199): value semantics; copies share the immutable device shard
// ---------------------------------------------------------------------------------------------
template <typename alphabet_tag = byte_alphabet_tag>
class vlg_index {
public:
    typedef alphabet_tag alphabet_category;
    typedef uint64_t size_type;
    typedef gapped_pattern_query<alphabet_tag> query_type;

    vlg_index() = default;
    explicit vlg_index(vlgb_index* h) : m_h(std::make_shared<detail::index_handle>(h)) {}

    bool empty() const { return !m_h || vlgb_index_size(m_h->h) == 0; }
    size_type size() const { return m_h ? vlgb_index_size(m_h->h) : 0; }  // symbols of the text
    vlgb_index* handle() const { return m_h ? m_h->h : nullptr; }
    uint32_t symbol_width() const { return m_h ? vlgb_index_symbol_width(m_h->h) : 1; }

    // vlg_index::text (vlg_index.hpp:141): the indexed text, copied back from the device on demand
    std::vector<uint8_t> text_bytes() const
    {
        const uint64_t nb = size() * symbol_width();
        std::vector<uint8_t> t(nb);
        if (nb) detail::check(vlgb_index_extract(m_h->h, 0, nb, t.data()));
        return t;
    }

    void swap(vlg_index& o) { m_h.swap(o.m_h); }

    // serialize / load (vlg_index.hpp:183-198): one blob in the library's own format (magic, extents, text)
    size_type serialize(std::ostream& out, void* /*structure_tree_node*/ = nullptr, std::string /*name*/ = "") const
    {
        if (!m_h) return 0;
        const std::string tmp = tmp_name();
        detail::check(vlgb_index_save(m_h->h, tmp.c_str()));
        std::ifstream in(tmp.c_str(), std::ios::binary);
        out << in.rdbuf();
        in.seekg(0, std::ios::end);
        const size_type written = (size_type)in.tellg();
        std::remove(tmp.c_str());
        return written;
    }
    void load(std::istream& in)
    {
        const std::string tmp = tmp_name();
        {
            std::ofstream o(tmp.c_str(), std::ios::binary);
            o << in.rdbuf();
        }
        vlgb_index* h = nullptr;
        const int rc = vlgb_index_load(tmp.c_str(), &h);
        std::remove(tmp.c_str());
        detail::check(rc);
        m_h = std::make_shared<detail::index_handle>(h);
    }

private:
    static std::string tmp_name()
    {
        const char* d = std::getenv("TMPDIR");
        return std::string(d ? d : "/tmp") + "/vlgb200_" + std::to_string((unsigned long long)::getpid()) + "_" +
               std::to_string((unsigned long long)(uintptr_t)&d) + ".idx";
    }
    std::shared_ptr<detail::index_handle> m_h;
};

// construct(idx, file, num_bytes) (vlg_index.hpp:375-392).  num_bytes = 1: raw byte file.  num_bytes = 0 on an
// int-alphabet index: a serialized sdsl int_vector<> (u64 bit length, u8 width, packed words -- int_vector.hpp:584-600).
inline std::vector<uint64_t> read_sdsl_int_vector(const std::string& file);  // below

inline void construct(vlg_index<byte_alphabet_tag>& idx, const std::string& file, uint8_t num_bytes = 1)
{
    if (num_bytes != 1) throw std::logic_error("vlg_index<byte_alphabet_tag>: num_bytes must be 1");
    vlgb_index* h = nullptr;
    detail::check(vlgb_index_build_file(file.c_str(), &h));
    idx = vlg_index<byte_alphabet_tag>(h);
}
inline void construct(vlg_index<int_alphabet_tag>& idx, const std::string& file, uint8_t num_bytes = 0)
{
    std::vector<uint64_t> sym;
    if (num_bytes == 0) {
        sym = read_sdsl_int_vector(file);
    } else {  // num_bytes-wide little-endian integers, like load_vector_from_file (io.hpp:177-230)
        std::ifstream in(file.c_str(), std::ios::binary);
        if (!in) throw std::runtime_error("cannot open " + file);
        std::vector<char> raw((std::istreambuf_iterator<char>(in)), std::istreambuf_iterator<char>());
        sym.resize(raw.size() / num_bytes);
        for (size_t i = 0; i < sym.size(); i++) {
            uint64_t v = 0;
            for (uint8_t b = 0; b < num_bytes; b++) v |= (uint64_t)(uint8_t)raw[i * num_bytes + b] << (8 * b);
            sym[i] = v;
        }
    }
    vlgb_index* h = nullptr;
    detail::check(vlgb_index_build_symbols(sym.data(), sym.size(), 0, sym.size(), sym.size(), &h));
    idx = vlg_index<int_alphabet_tag>(h);
}
// construct_im(idx, data, num_bytes) (construct.hpp:66-73)
inline void construct_im(vlg_index<byte_alphabet_tag>& idx, const std::string& data, uint8_t /*num_bytes*/ = 1)
{
    vlgb_index* h = nullptr;
    detail::check(vlgb_index_build(reinterpret_cast<const uint8_t*>(data.data()), data.size(), 0, data.size(), data.size(), &h));
    idx = vlg_index<byte_alphabet_tag>(h);
}
inline void construct_im(vlg_index<int_alphabet_tag>& idx, const std::vector<uint64_t>& data, uint8_t /*num_bytes*/ = 0)
{
    vlgb_index* h = nullptr;
    detail::check(vlgb_index_build_symbols(data.data(), data.size(), 0, data.size(), data.size(), &h));
    idx = vlg_index<int_alphabet_tag>(h);
}

// ---------------------------------------------------------------------------------------------
// vlg_iterator / locate / count  (vlg_index.hpp:209-373, 395-411)
// ---------------------------------------------------------------------------------------------
// The whole match set of a query is produced by one device search; the iterator walks the k-tuples.
class vlg_iterator : public std::iterator<std::forward_iterator_tag, uint64_t> {
public:
    typedef uint64_t size_type;
    vlg_iterator() : m_tuples(nullptr), m_k(0), m_i(0), m_n(0) {}  // end iterator (vlg_index.hpp:296-300)
    vlg_iterator(std::shared_ptr<detail::result_handle> r, size_type k)
        : m_r(std::move(r)), m_tuples(vlgb_result_tuples(m_r->h, 0)), m_k(k), m_i(0), m_n(vlgb_result_count(m_r->h, 0))
    {
    }
    size_type size() const { return m_k; }                                  // number of subpatterns, vlg_index.hpp:331-334
    size_type operator[](size_type i) const { return m_tuples[m_i * m_k + i]; }  // start of subpattern i, :337-340
    size_type operator*() const { return m_tuples[m_i * m_k]; }            // start of the match, :343-346
    vlg_iterator& operator++() { ++m_i; return *this; }
    vlg_iterator operator++(int) { vlg_iterator t(*this); ++m_i; return t; }
    bool is_end() const { return m_i >= m_n; }
    // like the reference (vlg_index.hpp:359-364) equality is only meaningful between end iterators
    friend bool operator==(const vlg_iterator& a, const vlg_iterator& b) { return a.is_end() && b.is_end(); }
    friend bool operator!=(const vlg_iterator& a, const vlg_iterator& b) { return !(a == b); }

private:
    std::shared_ptr<detail::result_handle> m_r;
    const uint64_t* m_tuples;
    size_type m_k, m_i, m_n;
};

template <class t_iter>
class container {  // include/sdsl/iterators.hpp:169-189
public:
    container(t_iter b, t_iter e) : m_b(std::move(b)), m_e(std::move(e)) {}
    t_iter begin() const { return m_b; }
    t_iter end() const { return m_e; }

private:
    t_iter m_b, m_e;
};

namespace detail {

// encode the subpatterns of a query for `idx`; false = a symbol cannot occur in the text (no match)
inline bool encode_query(const vlg_index<byte_alphabet_tag>&, const gapped_pattern_query<byte_alphabet_tag>& q, packed_batch& pb)
{
    pb.add(q.subpatterns, q.raw_gaps);
    return true;
}
inline bool encode_query(const vlg_index<int_alphabet_tag>& idx, const gapped_pattern_query<int_alphabet_tag>& q, packed_batch& pb)
{
    const uint32_t w = idx.symbol_width();
    std::vector<std::string> subs;
    for (const auto& p : q.subpatterns) {
        std::string e(p.size() * w, '\0');
        const int rc = vlgb_encode_symbols(w, p.data(), p.size(), reinterpret_cast<uint8_t*>(&e[0]));
        if (rc == VLGB_ERR_TOO_LARGE || rc == VLGB_ERR_ZERO_SYMBOL) return false;
        check(rc);
        subs.push_back(std::move(e));
    }
    pb.add(subs, q.raw_gaps);
    return true;
}

template <class t_index>
std::shared_ptr<result_handle> run_query(const t_index& idx, const typename t_index::query_type& q, uint32_t want)
{
    if (!idx.handle()) return nullptr;
    packed_batch pb;
    if (!encode_query(idx, q, pb)) return nullptr;
    vlgb_batch b = pb.view();
    vlgb_result* r = nullptr;
    check(vlgb_search_batch(idx.handle(), &b, nullptr, want, &r));
    auto h = std::make_shared<result_handle>(r);
    const int32_t st = vlgb_result_status(r, 0);
    if (st != VLGB_OK)
        throw std::runtime_error("vlgb200: query rejected (" + std::to_string(st) +
                                 (st == VLGB_ERR_EMPTY_SUB ? ": empty subpattern" : "") + ")");
    return h;
}

}  // namespace detail

template <class t_index>
container<vlg_iterator> locate(const t_index& idx, const typename t_index::query_type& query)
{
    auto r = detail::run_query(idx, query, VLGB_WANT_TUPLES);
    if (!r) return container<vlg_iterator>(vlg_iterator(), vlg_iterator());
    return container<vlg_iterator>(vlg_iterator(r, query.subpatterns.size()), vlg_iterator());
}

template <class t_index>
uint64_t count(const t_index& idx, const typename t_index::query_type& query)
{
    auto r = detail::run_query(idx, query, VLGB_WANT_COUNT);
    return r ? vlgb_result_count(r->h, 0) : 0;
}

// write_structure<HTML_FORMAT>(idx, out) (examples/vlg_matching.cpp:36-37): a small space report.  The
// reference prints the structure tree of text + wavelet tree; here the index is the text shard in HBM.
template <class format, class t_index>
void write_structure(const t_index& idx, std::ostream& out)
{
    const uint64_t bytes = idx.size() * idx.symbol_width();
    if (std::is_same<format, JSON_FORMAT>::value) {
        out << "{\"class_name\":\"vlg_b200::vlg_index\",\"name\":\"vlg_index\",\"size\":" << bytes
            << ",\"children\":[{\"class_name\":\"device_text\",\"name\":\"text\",\"size\":" << bytes << "}]}\n";
    } else {
        out << "<html><head><title>vlg_b200::vlg_index</title></head><body>\n<h1>vlg_b200::vlg_index</h1>\n<table border=\"1\">"
            << "<tr><th>member</th><th>bytes (HBM)</th></tr>\n<tr><td>text (" << idx.size() << " symbols x " << idx.symbol_width()
            << " B)</td><td>" << bytes << "</td></tr>\n<tr><td>total</td><td>" << bytes << "</td></tr></table>\n</body></html>\n";
    }
}

// A serialized sdsl::int_vector<0> (int_vector.hpp:584-600): u64 length in bits, u8 width, then the bits packed
// into little-endian u64 words.  This is also the format of a collection's text.TEXT (create_collection.cpp:79-93).
inline std::vector<uint64_t> read_sdsl_int_vector(const std::string& file)
{
    std::ifstream in(file.c_str(), std::ios::binary);
    if (!in) throw std::runtime_error("cannot open " + file);
    uint64_t bits = 0;
    uint8_t width = 0;
    in.read(reinterpret_cast<char*>(&bits), 8);
    in.read(reinterpret_cast<char*>(&width), 1);
    if (!in || width == 0 || width > 64 || bits % width) throw std::runtime_error(file + ": not a serialized int_vector<>");
    const uint64_t n = bits / width, words = (bits + 63) / 64;
    if (words > (uint64_t(1) << 36)) throw std::bad_alloc();  // same symptom as the reference on a raw byte file
    std::vector<uint64_t> w(words + 1, 0);
    in.read(reinterpret_cast<char*>(w.data()), (std::streamsize)(words * 8));
    if ((uint64_t)in.gcount() != words * 8) throw std::runtime_error(file + ": truncated int_vector<>");
    std::vector<uint64_t> v(n);
    const uint64_t mask = width == 64 ? ~uint64_t(0) : ((uint64_t(1) << width) - 1);
    for (uint64_t i = 0; i < n; i++) {
        const uint64_t bit = i * width, word = bit >> 6, off = bit & 63;
        uint64_t x = w[word] >> off;
        if (off + width > 64) x |= w[word + 1] << (64 - off);
        v[i] = x & mask;
    }
    return v;
}
inline void write_sdsl_int_vector(const std::string& file, const std::vector<uint64_t>& v, uint8_t width)
{
    std::ofstream out(file.c_str(), std::ios::binary);
    if (!out) throw std::runtime_error("cannot create " + file);
    const uint64_t bits = (uint64_t)v.size() * width, words = (bits + 63) / 64;
    std::vector<uint64_t> w(words + 1, 0);
    for (uint64_t i = 0; i < v.size(); i++) {
        const uint64_t bit = i * width, word = bit >> 6, off = bit & 63;
        w[word] |= v[i] << off;
        if (off + width > 64) w[word + 1] |= v[i] >> (64 - off);
    }
    out.write(reinterpret_cast<const char*>(&bits), 8);
    out.write(reinterpret_cast<const char*>(&width), 1);
    out.write(reinterpret_cast<const char*>(w.data()), (std::streamsize)(words * 8));
}

}  // namespace vlg_b200
