// TEST INFRASTRUCTURE ONLY — not product code, never linked into libvlgb200.so.
//
// Thin command-line driver around the UNMODIFIED reference matcher.  It is compiled by
// oracle/Makefile against the headers and lib/*.cpp where they lie under /root/reference
// (nothing is copied), and the binary lands in oracle/_ref/ (git-ignored).  All matching is done by
// the reference's own code:
//   sdsl::vlg_index<>            /root/reference/include/sdsl/vlg_index.hpp:116-199
//   sdsl::construct(vlg_index&)  /root/reference/include/sdsl/vlg_index.hpp:375-392
//   sdsl::locate / sdsl::count   /root/reference/include/sdsl/vlg_index.hpp:395-411
// This file only reads inputs, calls those functions exactly the way examples/vlg_matching.cpp:7-23
// does (count, then locate and it[i] for every subpattern) and prints the tuples.
//
// Usage:
//   ref_vlg locate <text-file> <query-file>            print every match tuple of every query
//   ref_vlg build  <text-file> <index-file>            construct + serialize (vlg_index.hpp:183-191)
//   ref_vlg bench  <text-file|index-file> <query-file> <threads> [--index] [--steps K] [--warmup W]
//                                                       time locate() over all queries, W + K passes in ONE process
//                                                       (the index is loaded once); JSON on stdout
// Query file: one query per line in the LIBRARY syntax  P1.{a,b}?P2...   (vlg_index.hpp:54-105).
// Output of `locate`, one block per query:
//   Q <line-no> <count>            or   E <line-no> <exception text>
//   <p_1> <p_2> ... <p_k>          (count lines)
#include <sdsl/vlg_index.hpp>

#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <string>
#include <thread>
#include <unistd.h>
#include <vector>

using namespace sdsl;

static std::string abspath(const std::string& p)
{
    char* r = realpath(p.c_str(), nullptr);
    if (!r) { perror(p.c_str()); exit(2); }
    std::string s(r);
    free(r);
    return s;
}

static std::vector<std::string> read_lines(const std::string& file)
{
    std::ifstream in(file);
    if (!in) { std::cerr << "cannot open " << file << std::endl; exit(2); }
    std::vector<std::string> v;
    std::string line;
    while (std::getline(in, line)) v.push_back(line);
    return v;
}

// construct() drops its temporary SA/BWT files into the current directory
// (vlg_index.hpp:378), so run it from a private scratch directory.
static void enter_scratch_dir()
{
    char tmpl[] = "/tmp/ref_vlg_XXXXXX";
    char* d = mkdtemp(tmpl);
    if (!d || chdir(d) != 0) { perror("scratch dir"); exit(2); }
}

static void build_from_text(vlg_index<>& idx, const std::string& text_file)
{
    std::string f = abspath(text_file);
    enter_scratch_dir();
    construct(idx, f, 1);
}

int main(int argc, char** argv)
{
    if (argc < 4) {
        std::cerr << "usage: ref_vlg locate|build|bench ..." << std::endl;
        return 2;
    }
    std::string mode = argv[1];

    if (mode == "locate") {
        auto queries = read_lines(abspath(argv[3]));
        vlg_index<> idx;
        build_from_text(idx, argv[2]);
        for (size_t q = 0; q < queries.size(); ++q) {
            try {
                auto cnt = count(idx, queries[q]);
                std::cout << "Q " << q << " " << cnt << "\n";
                auto res = locate(idx, queries[q]);
                for (auto it = res.begin(); it != res.end(); ++it) {
                    for (size_t i = 0; i < it.size(); ++i)
                        std::cout << (i ? " " : "") << it[i];
                    std::cout << "\n";
                }
            } catch (const std::exception& e) {
                std::cout << "E " << q << " " << e.what() << "\n";
            }
        }
        return 0;
    }

    if (mode == "build") {
        std::string out = argv[3];
        if (out[0] != '/') {
            char cwd[4096];
            if (!getcwd(cwd, sizeof cwd)) return 2;
            out = std::string(cwd) + "/" + out;
        }
        vlg_index<> idx;
        auto t0 = std::chrono::steady_clock::now();
        build_from_text(idx, argv[2]);
        double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        std::ofstream ofs(out, std::ios::binary);
        auto bytes = idx.serialize(ofs);
        std::cout << "{\"build_s\": " << dt << ", \"index_bytes\": " << bytes << "}" << std::endl;
        return 0;
    }

    if (mode == "bench") {
        if (argc < 5) return 2;
        auto queries = read_lines(abspath(argv[3]));
        int threads = atoi(argv[4]);
        bool from_index = false;
        int steps = 1, warmup = 0;
        for (int a = 5; a < argc; ++a) {
            std::string opt = argv[a];
            if (opt == "--index") from_index = true;
            else if (opt == "--steps" && a + 1 < argc) steps = atoi(argv[++a]);
            else if (opt == "--warmup" && a + 1 < argc) warmup = atoi(argv[++a]);
        }
        vlg_index<> idx;
        auto t0 = std::chrono::steady_clock::now();
        if (from_index) {
            std::ifstream ifs(argv[2], std::ios::binary);
            if (!ifs) { std::cerr << "cannot open index" << std::endl; return 2; }
            idx.load(ifs);
        } else {
            build_from_text(idx, argv[2]);
        }
        double build_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();

        // The matcher is single-threaded and const: split the query list over
        // worker threads that share the one index.
        std::vector<double> step_s;
        uint64_t r = 0, c = 0;
        double search_s = 0;
        for (int pass = 0; pass < warmup + steps; ++pass) {
            std::atomic<size_t> next(0);
            std::vector<uint64_t> nres(threads, 0), csum(threads, 0);
            auto t1 = std::chrono::steady_clock::now();
            std::vector<std::thread> pool;
            for (int t = 0; t < threads; ++t) {
                pool.emplace_back([&, t]() {
                    size_t q;
                    while ((q = next.fetch_add(1)) < queries.size()) {
                        try {
                            auto res = locate(idx, queries[q]);
                            for (auto it = res.begin(); it != res.end(); ++it) {
                                nres[t]++;
                                csum[t] += *it;  // same checksum as gm_search.cpp:110-114
                            }
                        } catch (const std::exception&) {
                        }
                    }
                });
            }
            for (auto& th : pool) th.join();
            double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t1).count();
            r = c = 0;
            for (int t = 0; t < threads; ++t) { r += nres[t]; c += csum[t]; }
            if (pass >= warmup) { step_s.push_back(dt); search_s += dt; }
        }
        search_s /= (double)step_s.size();
        std::cout << "{\"build_s\": " << build_s << ", \"search_s\": " << search_s << ", \"steps\": " << step_s.size()
                  << ", \"queries\": " << queries.size() << ", \"threads\": " << threads
                  << ", \"text_bytes\": " << idx.text.size()
                  << ", \"num_results\": " << r << ", \"checksum\": " << c << "}" << std::endl;
        return 0;
    }

    std::cerr << "unknown mode " << mode << std::endl;
    return 2;
}
