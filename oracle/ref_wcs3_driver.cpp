// TEST INFRASTRUCTURE ONLY — not product code, never linked into libvlgb200.so.
//
// Command-line driver around the reference's BENCHMARK matcher, index_wcsearch3
// (/root/reference/benchmark/gapped-matching/include/index_wcsearch3.hpp:349-423: its search() hands
// wild_card_match_iterator3 ONE (min_gap, max_gap) = |P_1| + gaps[0] for every level, :412-413).  Used to generate
// tests/golden/wcsearch3_ref.json, which pins the "wcsearch3 gap semantics" switch of the B200 benchmark shim.
// All matching is done by the reference's own code.  Two things differ from the reference's own build, both forced
// by what is (not) under /root/reference:
//   * the header is compiled from a copy made by oracle/Makefile into _ref/gen/ with the one-token fix SURVEY.md
//     Appendix C records (`sdsl::range_type(sp, ep)` -> `sdsl::range_type{{sp, ep}}`, :302: std::array has no such
//     constructor, the header does not compile as it is);
//   * index_wcsearch3's constructor builds a csa_wt through libdivsufsort, which is not vendored.  The index is
//     therefore built by sdsl::construct(vlg_index&) (vlg_index.hpp:375-392, qsufsort) -- the same text + wavelet tree
//     over the suffix array, same wt type -- serialized as a matching_index<> (index_wcsearch3.hpp:75-84) and handed
//     to index_wcsearch3::load().
//
// Usage:  ref_wcs3 <text-file> <query-file>      one query per line in the BENCHMARK syntax  P1.{a,b}P2...  (k >= 2)
// Output, one line per query:  Q <line-no> <count> <pos_1> <pos_2> ...   or   X <line-no> reference died (...)
#include <sdsl/vlg_index.hpp>

#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <sstream>
#include <string>
#include <sys/wait.h>
#include <unistd.h>
#include <vector>

#define IDXNAME "ORACLE"
#include "index_wcsearch3.hpp"  // the patched copy in _ref/gen (see above)

int main(int argc, char** argv)
{
    if (argc < 3) {
        std::cerr << "usage: ref_wcs3 <text-file> <query-file>" << std::endl;
        return 2;
    }
    char* rp = realpath(argv[1], nullptr);
    if (!rp) { perror(argv[1]); return 2; }
    const std::string text_file(rp);
    std::vector<std::string> queries;
    {
        std::ifstream in(argv[2]);
        std::string line;
        while (std::getline(in, line)) queries.push_back(line);
    }
    // construct() drops temporary files into the current directory (vlg_index.hpp:378)
    char tmpl[] = "/tmp/ref_wcs3_XXXXXX";
    if (!mkdtemp(tmpl) || chdir(tmpl) != 0) { perror("scratch dir"); return 2; }
    sdsl::vlg_index<> vi;
    sdsl::construct(vi, text_file, 1);
    index_wcsearch3 idx;
    {
        sdsl::int_vector<0> text(vi.text.size(), 0, 8);
        for (size_t i = 0; i < vi.text.size(); i++) text[i] = vi.text[i];
        matching_index<> mi(text, vi.wt);
        std::stringstream ss;
        mi.serialize(ss);
        idx.load(ss);
    }
    std::ofstream devnull("/dev/null");
    std::streambuf* old = std::cerr.rdbuf(devnull.rdbuf());  // search() prints every query to stderr (:409)
    // One child process per query.  wild_card_match_iterator3::next() calls relax() without checking that the first
    // walker still has nodes (index_wcsearch3.hpp:217-236 after the pull-forward of :272-273): once the last match
    // has used up the first list, it reads dfs_stack.back() of an empty vector -- undefined behaviour that usually
    // ends the iteration but depends on the heap.  Forking keeps every query's heap identical and turns a crash into
    // an "X" line instead of losing the whole run.
    std::cout.flush();
    for (size_t q = 0; q < queries.size(); q++) {
        const pid_t pid = fork();
        if (pid == 0) {
            gapped_pattern pat(queries[q], true);
            if (pat.subpatterns.size() < 2) {  // gaps[0] does not exist (:412): undefined in the reference
                std::cout << "E " << q << " k<2" << std::endl;
                _exit(0);
            }
            const gapped_search_result res = idx.search(pat);
            std::ostringstream os;
            os << "Q " << q << " " << res.positions.size();
            for (uint64_t p : res.positions) os << " " << p;
            os << "\n";
            const std::string line = os.str();
            (void)!write(1, line.data(), line.size());
            _exit(0);
        }
        int status = 0;
        waitpid(pid, &status, 0);
        if (!WIFEXITED(status) || WEXITSTATUS(status) != 0) {
            std::ostringstream os;
            os << "X " << q << " reference died (signal " << (WIFSIGNALED(status) ? WTERMSIG(status) : 0) << ")\n";
            const std::string line = os.str();
            (void)!write(1, line.data(), line.size());
        }
    }
    std::cerr.rdbuf(old);
    (void)!chdir("/");
    rmdir(tmpl);
    return 0;
}
