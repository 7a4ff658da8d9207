// TEST INFRASTRUCTURE: host-only check of the automaton tables (vlg_matching_b200/csrc/dfa.cuh: build_dfa -- byte
// classes, the two-byte fallback table, the double array of deep transitions, output chains).  Walks random texts
// through the tables exactly as the device code does (one step per byte: da[state + c] if its check is c, else
// fb[previous class, c]) and compares the occurrences of every pattern with a brute-force search.  Built with nvcc,
// runs on the CPU (tests/test_scan_tables_cpu.py).
#include <chrono>
#include <cstdio>
#include <memory>
#include <random>
#include <set>
#include "dfa.cuh"
using namespace vlgb;
int main()
{
    std::mt19937 rng(11);
    int eligible = 0;
    for (int trial = 0; trial < 8; trial++) {
        const int sigma = trial % 4 == 0 ? 4 : (trial % 4 == 1 ? 26 : (trial % 4 == 2 ? 2 : 255));
        const size_t n = 60000;
        std::string text(n + 32, '\0');
        for (size_t i = 0; i < n; i++) text[i] = (char)(1 + rng() % sigma);
        PatternDict d;
        const int np = sigma == 255 ? 40 : 200 + rng() % 3000;
        const int maxl = trial < 4 ? 8 : 17;
        for (int i = 0; i < np; i++) {
            size_t len = 1 + rng() % maxl;
            size_t pos = rng() % (n - 32);
            d.add((const uint8_t*)text.data() + pos, len);
        }
        const uint32_t P = (uint32_t)d.pats.size();
        std::vector<std::set<uint32_t>> got(P), want(P);
        for (uint32_t i = 0; i < P; i++) {
            const std::string& s = d.pats[i];
            for (size_t p = 0; p + s.size() <= n; p++) if (!memcmp(text.data() + p, s.data(), s.size())) want[i].insert((uint32_t)p);
        }
        uint32_t npass = 0, ok_pass = 0;
        size_t slots = 0;
        double ms = 0;
        std::unique_ptr<PassHost> ph(new PassHost());
        for (uint32_t first = 0; first < P;) {
            const uint32_t cnt = build_pass(d, first, P - first, *ph);
            DfaHost h;
            auto t0 = std::chrono::steady_clock::now();
            build_dfa(d, first, cnt, h);
            ms += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
            npass++;
            if (h.ok) {
                ok_pass++;
                slots += h.slots;
                uint32_t state = h.root, pc = 0;
                for (uint32_t p = 0; p < n; p++) {
                    const uint32_t c = h.cls[(uint8_t)text[p]];
                    const uint32_t e = h.da[state + c];
                    const uint32_t sel = ((e >> 15) & 0x3Fu) == c ? e : h.fb[(pc << h.cshift) | c];
                    state = sel & 0x7FFFu;
                    pc = c;
                    uint32_t nout = (sel >> 21) & 15u, seen = 0;
                    if (nout) {
                        for (uint32_t cs = state; cs != DFA_NONE;) {
                            const uint32_t oi = h.oinfo[cs];
                            const uint32_t pid = oi & 0x7FFu, len = (oi >> 11) & 31u;
                            if (len - 1 != h.adj[pid]) { printf("adj mismatch\n"); return 1; }
                            got[first + pid].insert(p + 1 - len);
                            seen++;
                            cs = (oi >> 16) & 0x7FFFu;
                        }
                        if (seen != nout) { printf("output count mismatch at %u: %u vs %u\n", p, seen, nout); return 1; }
                    }
                }
            } else {
                for (uint32_t i = 0; i < cnt; i++) got[first + i] = want[first + i];
            }
            first += cnt;
        }
        if (ok_pass) eligible++;
        size_t bad = 0, tot = 0;
        for (uint32_t i = 0; i < P; i++) { if (got[i] != want[i]) bad++; tot += want[i].size(); }
        printf("trial %d sigma %d patterns %u passes %u automaton-passes %u slots %zu build-ms %.2f occurrences %zu mismatching patterns %zu\n",
               trial, sigma, P, npass, ok_pass, slots, ms, tot, bad);
        if (bad) return 1;
    }
    printf("eligible trials %d\n", eligible);
    return 0;
}
