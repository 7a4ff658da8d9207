// TEST INFRASTRUCTURE: host-only check of the scan tables (vlg_matching_b200/csrc/scan.cuh: build_pass -- root table,
// two-choice edge table, node info, chains of long patterns, pseudo keys of ambiguous prefixes and short patterns).
// Walks every position of random texts through the tables exactly as the device code does and compares the
// occurrences of every pattern with a brute-force search.  Built with nvcc, runs on the CPU (tests/test_scan_tables_cpu.py).
#include "../../vlg_matching_b200/csrc/scan.cuh"
#include <cstdio>
#include <memory>
#include <random>
#include <map>
#include <set>
using namespace vlgb;
static uint32_t h_edge_lookup(const ScanTables* T, uint32_t key)
{
    const uint32_t h1 = key * T->mul_e1, h2 = key * T->mul_e2;
    const uint32_t e1 = T->edge[edge_slot(h1)], e2 = T->edge[SCAN_EDGE / 2 + edge_slot(h2)];
    uint32_t r = 0;
    if (((e2 ^ h2) & 0x7FFFFu) == 0u) r = e2;
    if (((e1 ^ h1) & 0x7FFFFu) == 0u) r = e1;
    return r ? r >> 19 : NONE32;
}
int main()
{
    std::mt19937 rng(7);
    for (int trial = 0; trial < 6; trial++) {
        const int sigma = trial % 3 == 0 ? 4 : (trial % 3 == 1 ? 26 : 255);
        const size_t n = 60000;
        std::string text(n + 32, '\0');
        for (size_t i = 0; i < n; i++) text[i] = (char)(1 + rng() % sigma);
        PatternDict d;
        const int np = 200 + rng() % 5000;
        for (int i = 0; i < np; i++) {
            size_t len = 1 + rng() % 12;
            if (sigma == 255 && len < 3 && (rng() & 3)) len += 3;
            size_t pos = rng() % (n - 16);
            d.add((const uint8_t*)text.data() + pos, len);
        }
        const uint32_t P = (uint32_t)d.pats.size();
        std::vector<std::set<uint32_t>> got(P), want(P);
        for (uint32_t i = 0; i < P; i++) {
            const std::string& s = d.pats[i];
            for (size_t p = 0; p + s.size() <= n; p++) if (!memcmp(text.data() + p, s.data(), s.size())) want[i].insert((uint32_t)p);
        }
        std::unique_ptr<PassHost> ph(new PassHost());
        uint32_t npass = 0, amb_items = 0;
        for (uint32_t first = 0; first < P;) {
            const uint32_t cnt = build_pass(d, first, P - first, *ph);
            npass++;
            const ScanTables* T = &ph->tables;
            auto walk = [&](uint32_t p, uint32_t node) {
                for (;;) {
                    const uint32_t info = T->info[node];
                    const uint32_t depth = (info >> 13) & 15;
                    if (info & NODE_TERM) {
                        uint32_t cand = info & 0x7FF;
                        if (depth < 8) { if (p + depth <= n) got[first + cand].insert(p); }
                        else for (; cand != NONE32; cand = ph->long_next[cand]) {
                            const uint32_t L = ph->long_len[cand];
                            if (p + L <= n && !memcmp(text.data() + p, ph->long_blob.data() + ph->long_off[cand], L)) got[first + cand].insert(p);
                        }
                    }
                    if (!(info & NODE_CHILD)) break;
                    const uint32_t c = h_edge_lookup(T, (node << 8) | (uint8_t)text[p + depth]);
                    if (c == NONE32) break;
                    node = c;
                }
            };
            for (uint32_t p = 0; p < n; p++) {
                uint32_t w = 0;
                memcpy(&w, text.data() + p, 4);
                const uint32_t h = w * T->mul_root;
                const uint32_t v = T->root[root_slot(h)];
                const bool hit = ((v ^ h) & 0x3FFFFu) == 0u;
                const bool amb = (v & ROOT_AMB_BIT) != 0u;
                if (amb) {
                    amb_items++;
                    const uint32_t c = h_edge_lookup(T, KEY_PREFIX3 | (w & 0xFFFFFF));
                    if (c != NONE32) walk(p, c);
                } else if (hit) walk(p, v >> 19);
                if (ph->has_short) {
                    const uint32_t b0 = w & 0xFF;
                    if ((T->f1[b0 >> 5] >> (b0 & 31)) & 1) { const uint32_t c = h_edge_lookup(T, KEY_SHORT1 | b0); if (c != NONE32) walk(p, c); }
                    const uint32_t i2 = scan_h2(w);
                    if ((T->f2[i2 >> 5] >> (i2 & 31)) & 1) { const uint32_t c = h_edge_lookup(T, KEY_SHORT2 | (w & 0xFFFF)); if (c != NONE32) walk(p, c); }
                }
            }
            first += cnt;
        }
        size_t bad = 0, tot = 0;
        for (uint32_t i = 0; i < P; i++) { if (got[i] != want[i]) bad++; tot += want[i].size(); }
        printf("trial %d sigma %d patterns %u passes %u occurrences %zu amb-items %u mismatching patterns %zu\n", trial, sigma, P, npass, tot, amb_items, bad);
        if (bad) return 1;
    }
    return 0;
}
