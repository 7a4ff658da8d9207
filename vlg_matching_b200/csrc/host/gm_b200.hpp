// gm_b200.hpp — the benchmark surface of /root/reference/benchmark/gapped-matching on top of the C ABI.
//
//   gapped_pattern, gapped_search_result, parse_pattern_file   include/utils.hpp:17-106
//   collection (dir/text.TEXT, dir/{tmp,index,patterns,results})   include/collection.hpp:17-44, src/create_collection.cpp:62-68
//   index_b200: the duck-typed 7-method index interface the drivers instantiate through -DINDEX_TYPE
//       (src/gm_index.cpp:42-56, src/gm_search.cpp:71-104); reference implementation include/index_wcsearch3.hpp:349-423
// Differences, on purpose:
//   * index_b200::search applies every gap's own {a,b} with the lengths of its own subpattern (the library
//     semantics of vlg_index.hpp:95).  index_wcsearch3 applies gaps[0] + |P_1| to every gap
//     (index_wcsearch3.hpp:412-413); both agree for 2 subpatterns and for uniform gaps + lengths.
//   * a subpattern that does not occur gives an empty result (index_wcsearch3 has undefined behaviour there).
//   * search_batch() answers a whole pattern file in one device pass; search() is the one-pattern call.
#pragma once
#include <sys/stat.h>

#include <algorithm>
#include <chrono>
#include <map>

#include "vlg_b200.hpp"

namespace gm_b200 {

typedef std::vector<uint64_t> string_type;

struct gapped_pattern {  // utils.hpp:19-71 (benchmark syntax: no lazy '?', raw gaps)
    std::string raw_regexp;
    std::vector<string_type> subpatterns;
    std::vector<std::string> sub_strs;  // the literal subpattern strings
    std::vector<std::pair<uint64_t, uint64_t>> gaps;
    gapped_pattern(const std::string& p, bool string_patterns) : raw_regexp(p)
    {
        try {
            vlg_b200::detail::split_query(p, VLGB_SYNTAX_BENCHMARK, sub_strs, gaps);
        } catch (const std::runtime_error&) {
            throw std::runtime_error("invalid gap description");  // utils.hpp:57-59
        }
        for (const auto& s : sub_strs)
            subpatterns.push_back(string_patterns ? string_type(s.begin(), s.end()) : vlg_b200::detail::int_symbols(s));
    }
};

struct gapped_search_result {  // utils.hpp:73-80
    std::vector<uint64_t> positions;
    gapped_search_result() = default;
    explicit gapped_search_result(size_t n) { positions.resize(n); }
};

inline bool file_exists(const std::string& f)
{
    struct stat sb;
    return stat(f.c_str(), &sb) == 0 && S_ISREG(sb.st_mode);
}
inline bool directory_exists(const std::string& d)
{
    struct stat sb;
    return stat(d.c_str(), &sb) == 0 && S_ISDIR(sb.st_mode);
}
inline void create_directory(const std::string& d)
{
    if (!directory_exists(d) && mkdir(d.c_str(), 0777) != 0) throw std::runtime_error("could not create directory " + d);
}

// one pattern per line, unparsable lines are skipped (utils.hpp:85-106)
inline std::vector<gapped_pattern> parse_pattern_file(const std::string& file_name, bool string_patterns)
{
    std::vector<gapped_pattern> patterns;
    std::ifstream in(file_name.c_str());
    if (!in) {
        std::cerr << "Cannot open pattern file '" << file_name << "'" << std::endl;
        return patterns;
    }
    std::string line;
    while (std::getline(in, line)) {
        try {
            patterns.emplace_back(line, string_patterns);
        } catch (...) {
            std::cerr << "Could not parse pattern '" << line << "'. Skipped" << std::endl;
        }
    }
    return patterns;
}

struct collection {  // collection.hpp:17-44
    std::string path;
    std::map<std::string, std::string> file_map;
    collection() = default;
    explicit collection(const std::string& p) : path(p + "/")
    {
        if (!directory_exists(path)) throw std::runtime_error("collection path not found.");
        const std::string text = path + "/text.TEXT";
        if (!file_exists(text)) throw std::runtime_error("collection path does not contain text.");
        file_map["TEXT"] = text;
    }
};

class index_b200 {
public:
    typedef uint64_t size_type;
    std::string name() const
    {
#ifdef IDXNAME
        return std::string("VLGB200-") + IDXNAME;
#else
        return "VLGB200";
#endif
    }
    // stands in for sdsl::util::class_to_hash(idx) in the index file name (gm_index.cpp:48)
    std::string class_hash() const { return "b200v2"; }

    index_b200() {}
    explicit index_b200(collection& col)  // index_wcsearch3(collection&), index_wcsearch3.hpp:365-369
    {
        const std::vector<uint64_t> sym = vlg_b200::read_sdsl_int_vector(col.file_map["TEXT"]);
        const bool bytes = std::all_of(sym.begin(), sym.end(), [](uint64_t v) { return v < 256; });
        vlgb_index* h = nullptr;
        if (bytes) {
            std::vector<uint8_t> t(sym.begin(), sym.end());
            vlg_b200::detail::check(vlgb_index_build(t.data(), t.size(), 0, t.size(), t.size(), &h));
        } else {
            vlg_b200::detail::check(vlgb_index_build_symbols(sym.data(), sym.size(), 0, sym.size(), sym.size(), &h));
        }
        m_h = std::make_shared<vlg_b200::detail::index_handle>(h);
        m_symbols = !bytes;
    }

    size_type serialize(std::ostream& out, void* = nullptr, std::string = "") const
    {
        return serialize_handle(out);
    }
    void load(std::istream& in)
    {
        vlg_b200::vlg_index<> tmp;
        tmp.load(in);
        m_keep = tmp;  // shares the device shard
        m_h.reset();
        m_symbols = vlgb_index_symbol_width(handle()) > 1;
    }
    void swap(index_b200& o)
    {
        if (this != &o) {
            m_h.swap(o.m_h);
            m_keep.swap(o.m_keep);
            std::swap(m_symbols, o.m_symbols);
        }
    }

    // index_wcsearch3::info (index_wcsearch3.hpp:388-397): the sum of the subpatterns' SA-range sizes
    std::string info(const gapped_pattern& pat) const
    {
        std::vector<gapped_pattern> one(1, pat);
        vlg_b200::detail::packed_batch pb;
        std::vector<char> possible;
        pack(one, pb, possible);
        if (!possible[0]) return "0";
        vlgb_batch b = pb.view();
        vlgb_result* r = nullptr;
        vlg_b200::detail::check(vlgb_search_batch(handle(), &b, nullptr, VLGB_WANT_COUNT, &r));
        const uint64_t occ = vlgb_result_occurrences(r, 0);
        vlgb_result_free(r);
        return std::to_string(occ);
    }
    void prepare(const gapped_pattern&) {}

    gapped_search_result search(const gapped_pattern& pat) const
    {
        std::vector<gapped_pattern> one(1, pat);
        return search_batch(one)[0];
    }

    // all patterns in one device pass; result i belongs to pats[i]
    std::vector<gapped_search_result> search_batch(const std::vector<gapped_pattern>& pats) const
    {
        std::vector<gapped_search_result> out(pats.size());
        if (pats.empty()) return out;
        vlg_b200::detail::packed_batch pb;
        std::vector<char> possible;
        pack(pats, pb, possible);
        vlgb_batch b = pb.view();
        vlgb_result* r = nullptr;
        vlg_b200::detail::check(vlgb_search_batch(handle(), &b, nullptr, VLGB_WANT_STARTS, &r));
        vlg_b200::detail::result_handle guard(r);
        for (size_t i = 0, q = 0; i < pats.size(); i++) {
            if (!possible[i]) continue;  // holds a symbol that cannot occur: no match
            const uint64_t c = vlgb_result_count(r, (uint32_t)q);
            const uint64_t* s = vlgb_result_starts(r, (uint32_t)q);
            if (c && s) out[i].positions.assign(s, s + c);
            q++;
        }
        return out;
    }

    vlgb_index* handle() const { return m_h ? m_h->h : m_keep.handle(); }

private:
    size_type serialize_handle(std::ostream& out) const
    {
        const std::string tmp = "/tmp/vlgb200_gm_" + std::to_string((unsigned long long)::getpid()) + ".idx";
        vlg_b200::detail::check(vlgb_index_save(handle(), tmp.c_str()));
        std::ifstream in(tmp.c_str(), std::ios::binary);
        out << in.rdbuf();
        in.clear();
        in.seekg(0, std::ios::end);
        const size_type n = (size_type)in.tellg();
        std::remove(tmp.c_str());
        return n;
    }
    void pack(const std::vector<gapped_pattern>& pats, vlg_b200::detail::packed_batch& pb, std::vector<char>& possible) const
    {
        const uint32_t w = vlgb_index_symbol_width(handle());
        possible.assign(pats.size(), 1);
        for (size_t i = 0; i < pats.size(); i++) {
            const gapped_pattern& p = pats[i];
            std::vector<std::string> subs;
            for (const auto& sp : p.subpatterns) {
                std::string e(sp.size() * w, '\0');
                if (!m_symbols && w == 1) {
                    for (size_t j = 0; j < sp.size(); j++) {
                        if (sp[j] == 0 || sp[j] > 255) possible[i] = 0;
                        e[j] = (char)sp[j];
                    }
                } else {
                    const int rc = vlgb_encode_symbols(w, sp.data(), sp.size(), reinterpret_cast<uint8_t*>(&e[0]));
                    if (rc != VLGB_OK) possible[i] = 0;
                }
                subs.push_back(std::move(e));
            }
            if (possible[i]) pb.add(subs, p.gaps);
        }
    }
    std::shared_ptr<vlg_b200::detail::index_handle> m_h;
    vlg_b200::vlg_index<> m_keep;  // set by load()
    bool m_symbols = false;
};

// timings.hpp:33-62
struct timing_results {
    std::vector<std::chrono::nanoseconds> timings;
    void add_timing(std::chrono::nanoseconds d) { timings.push_back(d); }
    struct summary_t {
        std::chrono::nanoseconds total{0}, min{0}, qrt_1st{0}, mean{0}, median{0}, qrt_3rd{0}, max{0};
    };
    summary_t summary()
    {
        summary_t t;
        if (timings.empty()) return t;
        std::sort(timings.begin(), timings.end());
        for (auto d : timings) t.total += d;
        t.min = timings.front();
        t.max = timings.back();
        t.mean = t.total / timings.size();
        t.median = timings[timings.size() / 2];
        t.qrt_1st = timings[timings.size() / 4];
        t.qrt_3rd = timings[timings.size() * 3 / 4];
        return t;
    }
};

}  // namespace gm_b200
