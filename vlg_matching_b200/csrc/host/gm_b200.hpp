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
//     set_gap_semantics(GAPS_WCSEARCH3) (gm_search --wcsearch3-gaps) reproduces index_wcsearch3's rule instead:
//     pinned by tests/golden/wcsearch3_ref.json, generated from the reference's own index_wcsearch3.
//   * a subpattern that does not occur gives an empty result (index_wcsearch3 has undefined behaviour there).
//   * search_batch() answers a whole pattern file in one device pass; search() is the one-pattern call.
#pragma once
#include <sys/stat.h>

#include <pthread.h>

#include <algorithm>
#include <chrono>
#include <map>
#include <thread>

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

// ---- which start-to-start bounds a pattern's levels get ----------------------------------------------------------
// GAPS_PER_GAP (default): level i uses its own gap and its own length, lo_i = a_i + |P_i| (vlg_index.hpp:95).
// GAPS_WCSEARCH3: every level uses ONE bound, [|P_1| + a_1, |P_1| + b_1] (index_wcsearch3.hpp:412-413 hands
// wild_card_match_iterator3 a single (min_gap, max_gap); relax() applies it to every pair of neighbours, :217-236).
// Expressed through the library's per-gap form: a_i' = |P_1| + a_1 - |P_i|, b_i' = |P_1| + b_1 - |P_i|.  A level
// whose subpattern is longer than |P_1| + a_1 would need a negative gap (the reference then lets the occurrences
// overlap): not representable, reported as an error for that pattern.
enum gap_semantics_t { GAPS_PER_GAP = 0, GAPS_WCSEARCH3 = 1 };
inline gap_semantics_t& gap_semantics_ref()
{
    static gap_semantics_t g = GAPS_PER_GAP;
    return g;
}
inline void set_gap_semantics(gap_semantics_t g) { gap_semantics_ref() = g; }
inline std::vector<std::pair<uint64_t, uint64_t>> effective_gaps(const gapped_pattern& p)
{
    if (gap_semantics_ref() == GAPS_PER_GAP || p.gaps.empty()) return p.gaps;
    std::vector<std::pair<uint64_t, uint64_t>> g(p.gaps.size());
    const uint64_t lo = p.subpatterns[0].size() + p.gaps[0].first, hi = p.subpatterns[0].size() + p.gaps[0].second;
    for (size_t i = 0; i < g.size(); i++) {
        const uint64_t len = p.subpatterns[i].size();
        if (len > lo)
            throw std::runtime_error("wcsearch3 gap semantics: subpattern " + std::to_string(i + 1) + " of '" + p.raw_regexp +
                                     "' is longer than |P_1| + min gap (overlapping occurrences are not representable)");
        g[i] = std::make_pair(lo - len, hi - len);
    }
    return g;
}

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
            if (possible[i]) pb.add(subs, effective_gaps(p));
        }
    }
    std::shared_ptr<vlg_b200::detail::index_handle> m_h;
    vlg_b200::vlg_index<> m_keep;  // set by load()
    bool m_symbols = false;
};

// ---- several GPUs in ONE process: one host thread per GPU drives one rank of vlgb_search_sharded (the
// other way to drive it is one process per GPU, bench.py).  The text is cut into n_gpus contiguous shards that overlap
// by the largest span of the pattern file; every rank answers all patterns on its shard; the NCCL exchange of the
// non-overlap cursors (pull_forward state, vlg_index.hpp:253-264) happens inside the library.
struct multi_gpu_run {
    std::vector<gapped_search_result> results;  // per pattern, whole-text positions, ascending
    double seconds = 0;                         // slowest rank, search only (shards resident)
    vlgb_shard_stats stats;
};

inline multi_gpu_run search_batch_multi_gpu(const std::vector<uint8_t>& text, const std::vector<gapped_pattern>& pats, int n_gpus)
{
    multi_gpu_run run;
    memset(&run.stats, 0, sizeof run.stats);
    run.results.resize(pats.size());
    if (pats.empty() || n_gpus < 1) return run;
    if (vlgb_device_count() < n_gpus) throw std::runtime_error("search_batch_multi_gpu: fewer CUDA devices than --gpus");
    vlg_b200::detail::packed_batch pb;
    std::vector<char> possible(pats.size(), 1);
    uint64_t span = 0;
    for (size_t i = 0; i < pats.size(); i++) {
        std::vector<std::string> subs;
        uint64_t sp = 0;
        const std::vector<std::pair<uint64_t, uint64_t>> gaps = effective_gaps(pats[i]);
        for (size_t j = 0; j < pats[i].subpatterns.size(); j++) {
            const string_type& s = pats[i].subpatterns[j];
            std::string e(s.size(), '\0');
            for (size_t c = 0; c < s.size(); c++) {
                if (s[c] == 0 || s[c] > 255) possible[i] = 0;
                e[c] = (char)s[c];
            }
            sp += s.size() + (j < gaps.size() ? gaps[j].second : 0);
            subs.push_back(std::move(e));
        }
        if (possible[i]) {
            pb.add(subs, gaps);
            span = std::max(span, sp);
        }
    }
    vlgb_batch batch = pb.view();
    const uint64_t n = text.size();
    uint8_t id[VLGB_COMM_ID_BYTES] = {0};
    if (n_gpus > 1) vlg_b200::detail::check(vlgb_comm_unique_id(id));
    pthread_barrier_t bar;
    pthread_barrier_init(&bar, nullptr, (unsigned)n_gpus);
    std::vector<std::vector<std::vector<uint64_t>>> part(n_gpus, std::vector<std::vector<uint64_t>>(batch.n_queries));
    std::vector<double> secs(n_gpus, 0.0);
    std::vector<std::string> errors(n_gpus);
    std::vector<vlgb_shard_stats> st(n_gpus);
    auto worker = [&](int r) {
        vlgb_index* shard = nullptr;
        vlgb_result* res = nullptr;
        bool armed = true;  // every thread reaches every barrier, whatever fails
        int rc = vlgb_init(r);
        if (rc == VLGB_OK) rc = vlgb_comm_init(n_gpus > 1 ? id : nullptr, n_gpus, r);
        const uint64_t lo = n * (uint64_t)r / n_gpus, hi = n * (uint64_t)(r + 1) / n_gpus, end = std::min(n, hi + span);
        if (rc == VLGB_OK) rc = vlgb_index_build(text.data() + lo, end - lo, lo, hi - lo, n, &shard);
        if (rc != VLGB_OK) errors[r] = vlgb_last_error();
        pthread_barrier_wait(&bar);
        bool all_ok = true;
        for (int g = 0; g < n_gpus; g++) all_ok = all_ok && errors[g].empty();
        if (all_ok) {
            const vlgb_index* one[1] = {shard};
            auto t0 = std::chrono::high_resolution_clock::now();
            rc = vlgb_search_sharded(one, 1, &batch, VLGB_WANT_STARTS, &res, nullptr, &st[r]);
            secs[r] = std::chrono::duration<double>(std::chrono::high_resolution_clock::now() - t0).count();
            if (rc != VLGB_OK) errors[r] = vlgb_last_error();
            else
                for (uint32_t q = 0; q < batch.n_queries; q++) {
                    const uint64_t c = vlgb_result_count(res, q);
                    const uint64_t* s = vlgb_result_starts(res, q);
                    if (c && s) part[r][q].assign(s, s + c);
                }
        }
        (void)armed;
        vlgb_result_free(res);
        vlgb_index_free(shard);
        vlgb_comm_destroy();
    };
    std::vector<std::thread> pool;
    for (int r = 0; r < n_gpus; r++) pool.emplace_back(worker, r);
    for (auto& t : pool) t.join();
    pthread_barrier_destroy(&bar);
    for (int r = 0; r < n_gpus; r++)
        if (!errors[r].empty()) throw std::runtime_error("rank " + std::to_string(r) + ": " + errors[r]);
    for (size_t i = 0, q = 0; i < pats.size(); i++) {
        if (!possible[i]) continue;
        for (int r = 0; r < n_gpus; r++) run.results[i].positions.insert(run.results[i].positions.end(), part[r][q].begin(), part[r][q].end());
        q++;
    }
    run.seconds = *std::max_element(secs.begin(), secs.end());
    run.stats = st[0];
    return run;
}

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
