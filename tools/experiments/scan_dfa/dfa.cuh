// dfa.cuh — the scan's automaton kernel: the patterns of a pass as ONE deterministic automaton (Aho-Corasick closure),
// every LANE walking its own 64-byte piece of the text with one table step per byte.  Same contract as
// scan_kernel<SCAN_STREAM> (scan.cuh): one hit stream per row in text order, occurrence counts per (row group, pattern),
// halo counts -- scan_scatter_kernel then writes the lists.  Replaces, for a whole batch, forward_search
// (suffix_array_algorithm.hpp:48-115) + the in-order enumeration of the SA range (wt_helper.hpp:726-785).
//
// Why: the trie/queue kernel spends 2.4 warp instructions per text byte on BASELINE config 2 (one queue item per trie
// node reached: 0.44 per byte).  An automaton spends a FIXED ~20 thread instructions per byte whatever the text
// matches (0.6 warp instructions per byte) plus ~0.3 per byte to turn the recorded states into stream records.
//
// Tables of a pass (one bulk copy into shared memory):
//   cls[256]      byte -> class; 0 = "occurs in no pattern", 1..k the pattern bytes (k <= 62)
//   fb[side*side] the state reached when no longer context matters: fb[a << cshift | b] = the deepest trie node that is a
//                 suffix of the two bytes (a, b) -- every transition whose target has depth <= 2 is served from here
//   da[slots]     double array of the transitions whose target has depth >= 3: a state IS its base offset, the entry
//                 of (state, c) sits at da[state + c] and carries c as its check (bases are unique, so (slot, c) names
//                 the owner).  Transitions inherited through failure links are materialised (17 k entries for the
//                 4466 states of config 2), so a step is ONE dependent table load: no failure walk.
//   entry layout (da and fb alike): bits 0..14 next state | 15..20 check class (fb: 0) | 21..24 number of patterns that
//                 end in the next state (itself or a proper suffix), at most DFA_MAX_OUT
//   oinfo[slots]  (global memory, read when records are emitted) per state with outputs: pattern id | length << 11 |
//                 next state of the output chain << 16 (0x7FFF = end)
// Kernel: one CTA per row GROUP (16 rows = 16 warps); a warp takes its row in tiles of 32 x 64 bytes.
//   walk   every lane walks its 64 bytes after a warm-up over the (longest pattern - 1) bytes before them (a lane owns the
//          occurrences that END in its piece).  Each occurrence becomes a 32-bit note (state | position << 15 | index in
//          the state's output chain << 21) in the warp's staging block, row = the occurrence's index within the lane.
//   emit   the notes of all lanes in LANE-MAJOR order (= text order: what the scatter needs) are turned into stream
//          records 32 at a time: owner lane by a 5-step search over the lanes' prefix sums (shuffles), pattern id from
//          oinfo[], coalesced stores to the row's stream, counts in the group's shared-memory counters.
//          The staging block has R <= 64 rows (what the tables leave of the shared memory); a lane with more than R
//          occurrences in its 64 bytes (dense texts) is served in further rounds: the walk is repeated and keeps
//          occurrences [R r, R r + R) of every lane.
// Stream entries of this kernel hold the END position of the occurrence relative to the row start; the scatter subtracts
// length - 1 (ScanArgs::end_adj).  A pass that is not eligible (more than 62 distinct pattern bytes, patterns longer than
// 17 bytes, W-byte symbols, tables beyond the shared-memory budget) takes the trie/queue kernel.
#pragma once
#include "../../../vlg_matching_b200/csrc/scan.cuh"

namespace vlgb {

constexpr int DFA_WARPS = SCAN_GROUP_ROWS;   // one CTA = one row group
constexpr int DFA_C = 64;                    // text bytes per lane per tile
constexpr int DFA_TILE = 32 * DFA_C;
constexpr int DFA_MAXLEN = 17;               // warm-up = the 16 bytes before a lane's piece
constexpr uint32_t DFA_NONE = 0x7FFFu;
constexpr uint32_t DFA_MAX_SLOTS = 32768 - 64;
constexpr uint32_t DFA_MAX_OUT = 14;
constexpr size_t DFA_SMEM_MAX = 227 * 1024;
constexpr uint32_t DFA_MIN_ROWS = 24;        // rows of a warp's staging block (32 notes each): as many as fit, <= 64

struct DfaHost {
    bool ok = false;
    uint32_t cshift = 0, slots = 0, maxlen = 0, root = 0;
    std::vector<uint8_t> cls;      // 256
    std::vector<uint32_t> fb;      // (1 << 2*cshift)
    std::vector<uint32_t> da;      // slots (multiple of 4)
    std::vector<uint32_t> oinfo;   // slots
    std::vector<uint16_t> adj;     // per pattern of the pass: length - 1
    size_t table_bytes() const { return 256 + fb.size() * 4 + da.size() * 4; }
};

inline size_t dfa_smem_bytes(size_t table_bytes, uint32_t s_stride, uint32_t rows)
{
    return table_bytes + 16 + (size_t)s_stride * 4 + (size_t)DFA_WARPS * rows * 128;
}
// rows of the staging blocks that the shared memory left by the tables allows (0 = the pass does not fit)
inline uint32_t dfa_rows_for(size_t table_bytes, uint32_t s_stride)
{
    const size_t fixed = dfa_smem_bytes(table_bytes, s_stride, 0);
    if (fixed + (size_t)DFA_WARPS * DFA_MIN_ROWS * 128 > DFA_SMEM_MAX) return 0;
    return (uint32_t)std::min<size_t>(DFA_C, (DFA_SMEM_MAX - fixed) / ((size_t)DFA_WARPS * 128));
}

// ---- host side: the automaton of a pass ------------------------------------------------------------------------
// Builds the tables for dictionary patterns [first, first + count).  h.ok = false when the pass is not eligible.
inline void build_dfa(const PatternDict& d, uint32_t first, uint32_t count, DfaHost& h)
{
    h.ok = false;
    h.cls.assign(256, 0);
    uint32_t k = 0, maxlen = 0;
    for (uint32_t i = 0; i < count; i++) {
        const std::string& s = d.pats[first + i];
        if (s.empty() || s.size() > (size_t)DFA_MAXLEN) return;
        maxlen = std::max<uint32_t>(maxlen, (uint32_t)s.size());
        for (unsigned char b : s) {
            if (b == 0) return;  // the bytes before the first shard are read as zeros: they must belong to no pattern
            if (!h.cls[b]) {
                if (k == 62) return;
                h.cls[b] = (uint8_t)++k;
            }
        }
    }
    const uint32_t nc = k + 1;
    uint32_t cshift = 1;
    while ((1u << cshift) < nc) cshift++;
    // ---- trie: children as (class -> node) lists kept sorted by class
    struct Node {
        std::vector<std::pair<uint8_t, uint32_t>> ch;
        uint32_t fail = 0, depth = 0, pid = NONE32, olink = NONE32, nout = 0;
        std::vector<std::pair<uint8_t, uint32_t>> exc;  // (class, target node) of the transitions with target depth >= 3
    };
    std::vector<Node> t(1);
    auto child = [&](uint32_t s, uint8_t c) -> uint32_t {
        for (auto& pr : t[s].ch)
            if (pr.first == c) return pr.second;
        return NONE32;
    };
    for (uint32_t i = 0; i < count; i++) {
        const std::string& s = d.pats[first + i];
        uint32_t cur = 0;
        for (unsigned char b : s) {
            const uint8_t c = h.cls[b];
            uint32_t nx = child(cur, c);
            if (nx == NONE32) {
                nx = (uint32_t)t.size();
                t[cur].ch.emplace_back(c, nx);
                t.emplace_back();
                t[nx].depth = t[cur].depth + 1;
            }
            cur = nx;
        }
        t[cur].pid = i;  // distinct patterns: one per node
    }
    const uint32_t N = (uint32_t)t.size();
    if (N > DFA_MAX_SLOTS) return;
    for (auto& nd : t) std::sort(nd.ch.begin(), nd.ch.end());
    // ---- breadth first: failure links, output links, the materialised deep transitions
    std::vector<uint32_t> order;
    order.reserve(N);
    order.push_back(0);
    for (size_t qi = 0; qi < order.size(); qi++) {
        const uint32_t s = order[qi];
        Node& ns = t[s];
        if (s) {
            const Node& nf = t[ns.fail];
            ns.olink = nf.pid != NONE32 ? ns.fail : nf.olink;
            ns.nout = (ns.pid != NONE32 ? 1u : 0u) + (ns.olink != NONE32 ? t[ns.olink].nout : 0u);
            if (ns.nout > DFA_MAX_OUT) return;
            // exc(s) = exc(fail(s)) overridden by s's own children (their depth is >= 3 iff depth(s) >= 2)
            if (ns.depth >= 2) {
                size_t a = 0, b = 0;
                const auto& fe = nf.exc;
                while (a < fe.size() || b < ns.ch.size()) {
                    if (b == ns.ch.size() || (a < fe.size() && fe[a].first < ns.ch[b].first)) ns.exc.push_back(fe[a++]);
                    else {
                        if (a < fe.size() && fe[a].first == ns.ch[b].first) a++;
                        ns.exc.push_back(ns.ch[b++]);
                    }
                }
            }
        }
        for (auto& pr : ns.ch) {
            const uint32_t c = pr.second;
            // fail(child) = delta(fail(s), class): walk the failure chain of s
            uint32_t f = s ? ns.fail : NONE32;
            uint32_t to = 0;
            while (f != NONE32) {
                const uint32_t g = child(f, pr.first);
                if (g != NONE32) { to = g; break; }
                f = f ? t[f].fail : NONE32;
            }
            t[c].fail = to;
            order.push_back(c);
        }
    }
    // ---- double array: first fit, largest rows first; every state gets a UNIQUE base (states without deep transitions
    // too).  Occupancy as bit words: a row (<= 63 classes) is one 64-bit mask, "does it fit at base b" one shifted
    // compare; candidate bases are taken from the FREE slots only (bit scan).
    std::vector<uint32_t> by_size(N);
    for (uint32_t i = 0; i < N; i++) by_size[i] = i;
    std::stable_sort(by_size.begin(), by_size.end(), [&](uint32_t a, uint32_t b) { return t[a].exc.size() > t[b].exc.size(); });
    const uint32_t words = (DFA_MAX_SLOTS + 64) / 64 + 2;
    std::vector<uint64_t> slot_used(words, 0), base_used(words, 0);
    auto window = [&](const std::vector<uint64_t>& v, uint32_t b) {
        const uint32_t w = b >> 6, sh = b & 63;
        return sh ? (v[w] >> sh) | (v[w + 1] << (64 - sh)) : v[w];
    };
    auto next_free = [&](const std::vector<uint64_t>& v, uint32_t from) {  // first clear bit at or after `from`
        uint32_t w = from >> 6;
        uint64_t m = ~v[w] & (~0ull << (from & 63));
        while (!m) m = ~v[++w];
        return (w << 6) + (uint32_t)__builtin_ctzll(m);
    };
    std::vector<uint32_t> base(N, 0);
    uint32_t hi_slot = 0, first_free = 0, leaf_cursor = 0, prev_b = 0;
    for (uint32_t s : by_size) {
        const auto& e = t[s].exc;
        uint32_t b;
        if (e.empty()) {
            b = leaf_cursor = next_free(base_used, leaf_cursor);
            if (b >= DFA_MAX_SLOTS) return;
        } else {
            uint64_t row = 0;
            for (auto& pr : e) row |= 1ull << pr.first;
            const uint32_t c0 = e[0].first;
            first_free = next_free(slot_used, first_free);
            // rows of several entries do not retry the crowded front every time: they start a little before the place
            // the previous row went to (rows come largest first, what rejected the last one mostly rejects this one);
            // single entries fill the holes from the front
            uint32_t f = std::max(first_free, c0);  // the slot the row's first class would take
            if (e.size() > 1 && prev_b > 256) f = std::max(f, prev_b - 256 + c0);
            for (;; f++) {
                f = next_free(slot_used, f);
                b = f - c0;
                if (b + 63 >= DFA_MAX_SLOTS) return;
                if ((base_used[b >> 6] >> (b & 63)) & 1u) continue;
                if (!(window(slot_used, b) & row)) break;
            }
            const uint32_t w = b >> 6, sh = b & 63;
            slot_used[w] |= row << sh;
            if (sh) slot_used[w + 1] |= row >> (64 - sh);
            if (e.size() > 1) prev_b = b;
        }
        base_used[b >> 6] |= 1ull << (b & 63);
        base[s] = b;
        hi_slot = std::max(hi_slot, b + 64);
    }
    h.slots = (hi_slot + 3) & ~3u;
    h.cshift = cshift;
    h.maxlen = maxlen;
    h.root = base[0];
    auto entry = [&](uint32_t target, uint32_t chk) { return base[target] | (chk << 15) | (t[target].nout << 21); };
    h.da.assign(h.slots, 0x3Fu << 15);  // empty: a check no class has
    for (uint32_t s = 0; s < N; s++)
        for (auto& pr : t[s].exc) h.da[base[s] + pr.first] = entry(pr.second, pr.first);
    const uint32_t side = 1u << cshift;
    h.fb.assign((size_t)side * side, entry(0, 0));
    for (uint32_t a = 0; a < side; a++)
        for (uint32_t b = 1; b < nc; b++) {
            uint32_t node = 0;
            const uint32_t nb = child(0, (uint8_t)b);
            if (nb != NONE32) node = nb;
            if (a >= 1 && a < nc) {
                const uint32_t na = child(0, (uint8_t)a);
                if (na != NONE32) {
                    const uint32_t nab = child(na, (uint8_t)b);
                    if (nab != NONE32) node = nab;
                }
            }
            h.fb[(size_t)a * side + b] = entry(node, 0);
        }
    h.oinfo.assign(h.slots, 0);
    for (uint32_t s = 0; s < N; s++) {
        if (!t[s].nout) continue;
        const uint32_t src = t[s].pid != NONE32 ? s : t[s].olink;  // the first pattern that ends here
        const uint32_t nxt = t[src].olink != NONE32 ? base[t[src].olink] : DFA_NONE;
        h.oinfo[base[s]] = t[src].pid | (t[src].depth << 11) | (nxt << 16);
    }
    h.adj.resize(count);
    for (uint32_t i = 0; i < count; i++) h.adj[i] = (uint16_t)(d.pats[first + i].size() - 1);
    h.ok = true;
}

// ---- device side -----------------------------------------------------------------------------------------------
struct DfaArgs {
    const uint8_t* tables;   // cls[256] | fb[side * side] | da[slots]  (device, 16-byte aligned, table_bytes % 16 == 0)
    const uint32_t* oinfo;   // [slots]
    uint32_t table_bytes, da_off;  // da starts at tables + da_off (fb at tables + 256)
    uint32_t cshift, root, warm;   // warm = longest pattern - 1 (<= 16): bytes walked before a lane's piece
    uint32_t rows;                 // rows of a warp's staging block = occurrences per lane noted in one round
};

// one automaton step on text byte `b`: sel = the entry taken (next state | check | outputs << 21)
#define VLGB_DFA_STEP(b)                                                        \
    {                                                                           \
        const uint32_t c_ = cls[(b)];                                           \
        const uint32_t e_ = da[state + c_];                                     \
        sel = (((e_ >> 15) ^ c_) & 0x3Fu) == 0u ? e_ : fb[(pc << cshift) | c_]; \
        state = sel & 0x7FFFu;                                                  \
        pc = c_;                                                                \
    }

// 16 bytes of a lane's piece (quarter q).  Occurrence number i of the lane (i counts from the start of the piece) is
// noted in row i - base of the staging block if that is one of its 64 rows; the column is rotated by the row so that
// the emit phase, which reads consecutive rows of ONE lane, does not hit one bank.
template <bool FULL>
__device__ __forceinline__ void dfa_walk16(const uint4 v, const int q, const uint32_t nst, const uint8_t* cls, const uint32_t* fb,
                                           const uint32_t* da, const uint32_t cshift, uint32_t& state, uint32_t& pc, uint32_t& c,
                                           const uint32_t base, const uint32_t rows, uint32_t* st, const uint32_t lane)
{
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 16; i++) {
        const int j = q * 16 + i;
        const uint32_t b = (w[i >> 2] >> ((i & 3) * 8)) & 0xFFu;
        uint32_t sel;
        VLGB_DFA_STEP(b)
        uint32_t no = (sel >> 21) & 15u;
        if (!FULL && (uint32_t)j >= nst) no = 0u;
        if (no) {
            const uint32_t note = state | ((uint32_t)j << 15);
            uint32_t row = c - base;
            if (row < rows) st[row * 32 + ((lane + row) & 31u)] = note;
            for (uint32_t t = 1; t < no; t++) {  // further patterns that end here: suffixes of the first one
                row++;
                if (row < rows) st[row * 32 + ((lane + row) & 31u)] = note | (t << 21);
            }
            c += no;
        }
    }
}

__device__ __forceinline__ uint32_t warp_inclusive_sum(uint32_t v, uint32_t lane)
{
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t o = __shfl_up_sync(0xffffffffu, v, d);
        if (lane >= (uint32_t)d) v += o;
    }
    return v;
}

__global__ void __launch_bounds__(DFA_WARPS * 32, 1) scan_dfa_kernel(ScanArgs a, DfaArgs f)
{
    extern __shared__ __align__(128) uint8_t smem_raw[];
    const uint8_t* cls = smem_raw;
    const uint32_t* fb = reinterpret_cast<const uint32_t*>(smem_raw + 256);
    const uint32_t* da = reinterpret_cast<const uint32_t*>(smem_raw + f.da_off);
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem_raw + f.table_bytes);
    uint32_t* gcnt = reinterpret_cast<uint32_t*>(smem_raw + f.table_bytes + 16);
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t* st = reinterpret_cast<uint32_t*>(smem_raw + f.table_bytes + 16 + (size_t)a.s_stride * 4 + (size_t)warp * f.rows * 128);

    // stage the tables: ONE bulk asynchronous copy, completion on an mbarrier
    if (threadIdx.x == 0) {
        mbar_init(bar, 1);
        mbar_fence_init();
    }
    for (uint32_t i = threadIdx.x; i < a.s_stride; i += DFA_WARPS * 32) gcnt[i] = 0u;
    __syncthreads();
    if (threadIdx.x == 0) {
        mbar_expect_tx(bar, f.table_bytes);
        bulk_g2s(smem_raw, f.tables, f.table_bytes, bar);
    }
    const uint32_t row = blockIdx.x * DFA_WARPS + warp;
    const bool live = row < a.rows;
    const uint32_t lo = live ? row * a.chunk : 0u;
    const uint32_t hi = live ? (uint32_t)min((uint64_t)lo + a.chunk, (uint64_t)a.valid_len) : 0u;
    const uint32_t cshift = f.cshift;
    uint32_t* stream_row = live ? a.stream + (size_t)row * a.stream_cap : nullptr;
    uint32_t nh = 0;

    // the lane's text: the 16 bytes before its piece (warm-up) and the piece itself, in registers
    auto load16 = [&](uint64_t at) {
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (at < a.valid_len) v = __ldg(reinterpret_cast<const uint4*>(a.text + at));
        return v;
    };
    uint4 tb4[5];
    auto load_tile = [&](uint32_t tb) {
        const uint64_t cb = (uint64_t)tb + lane * DFA_C;
        if (tb < hi && cb < hi) {
            tb4[0] = cb >= 16 ? load16(cb - 16) : make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
            for (int q = 0; q < 4; q++) tb4[1 + q] = load16(cb + 16 * q);
        }
    };
    load_tile(lo);
    mbar_wait(bar, 0);

    for (uint32_t tb = lo; tb < hi; tb += DFA_TILE) {
        const uint32_t cb = tb + lane * DFA_C;  // no wrap: tb + 2048 stays below the shard limit (2^32 - 4096)
        const uint32_t nst = cb < hi ? min((uint32_t)DFA_C, hi - cb) : 0u;
        const uint32_t rel0 = tb - lo;
        uint32_t c = 0, inclC = 0;
        for (uint32_t base = 0;; base += f.rows) {
            // ---- walk: occurrences [base, base + 64) of every lane are noted
            uint32_t state = f.root, pc = 0u;
            c = 0;
            if (nst) {
                const uint32_t wv[4] = {tb4[0].x, tb4[0].y, tb4[0].z, tb4[0].w};
#pragma unroll
                for (int k = 0; k < 16; k++) {
                    if ((uint32_t)k + f.warm >= 16u) {  // uniform
                        const uint32_t b = (wv[k >> 2] >> ((k & 3) * 8)) & 0xFFu;
                        uint32_t sel;
                        VLGB_DFA_STEP(b)
                    }
                }
                if (tb + DFA_TILE <= hi) {  // uniform: every lane has a whole piece
#pragma unroll
                    for (int q = 0; q < 4; q++) dfa_walk16<true>(tb4[1 + q], q, nst, cls, fb, da, cshift, state, pc, c, base, f.rows, st, lane);
                } else {
#pragma unroll
                    for (int q = 0; q < 4; q++) dfa_walk16<false>(tb4[1 + q], q, nst, cls, fb, da, cshift, state, pc, c, base, f.rows, st, lane);
                }
            }
            const bool more = __any_sync(0xffffffffu, c > base + f.rows);
            if (!more) load_tile(tb + DFA_TILE);  // the next tile's text is in flight while this one's records are emitted
            if (base == 0u) inclC = warp_inclusive_sum(c, lane);
            const uint32_t cr = c > base ? min(c - base, f.rows) : 0u;  // notes of this round
            const uint32_t inclR = base == 0u && !more ? inclC : warp_inclusive_sum(cr, lane);
            const uint32_t totalR = __shfl_sync(0xffffffffu, inclR, 31);
            __syncwarp();
            // ---- emit: the round's notes in lane-major (= text) order, 32 per iteration
            for (uint32_t r0 = 0; r0 < totalR; r0 += 32) {
                const uint32_t r = r0 + lane;
                uint32_t L = 0;
#pragma unroll
                for (int s = 16; s >= 1; s >>= 1) {
                    const uint32_t probe = __shfl_sync(0xffffffffu, inclR, (L + s - 1) & 31u);
                    if (probe <= r) L += s;
                }
                L = min(L, 31u);
                const uint32_t eR = __shfl_sync(0xffffffffu, inclR - cr, L), eC = __shfl_sync(0xffffffffu, inclC - c, L);
                if (r < totalR) {
                    const uint32_t k = r - eR;
                    const uint32_t w = st[k * 32 + ((L + k) & 31u)];
                    uint32_t oi = __ldg(f.oinfo + (w & 0x7FFFu));
                    for (uint32_t t = (w >> 21) & 15u; t; t--) oi = __ldg(f.oinfo + ((oi >> 16) & 0x7FFFu));
                    const uint32_t pid = oi & 0x7FFu, len = (oi >> 11) & 31u;
                    const uint32_t endrel = rel0 + L * DFA_C + ((w >> 15) & 63u);
                    const uint32_t at = nh + eC + base + k;
                    if (at < a.stream_cap) stream_row[at] = endrel | (pid << 16);
                    atomicAdd(gcnt + pid, 1u);
                    if (lo + endrel + 1u - len >= a.owned_len) atomicAdd(a.halo + pid, 1u);  // a handful per shard
                }
            }
            __syncwarp();
            if (!more) break;
        }
        nh += __shfl_sync(0xffffffffu, inclC, 31);
    }
    if (live && lane == 0) {
        a.row_hits[row] = nh;
        if (nh > a.stream_cap) atomicAdd(a.overflow, 1u);
    }
    __syncthreads();
    if (blockIdx.x * DFA_WARPS < a.rows)
        for (uint32_t i = threadIdx.x; i < a.s_stride; i += DFA_WARPS * 32) a.gmatrix[(size_t)blockIdx.x * a.s_stride + i] = gcnt[i];
}
#undef VLGB_DFA_STEP

}  // namespace vlgb
