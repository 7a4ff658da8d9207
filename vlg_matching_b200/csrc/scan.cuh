// scan.cuh — multi-pattern text scan: text shard -> one sorted occurrence list per distinct
// subpattern (north-star (2), "coalesced text-scan path").  Replaces, for a whole batch at once, what
// the reference obtains per subpattern from forward_search (suffix_array_algorithm.hpp:48-115) plus the
// in-order WT enumeration of the SA range (wt_helper.hpp:726-785): the ascending list of text
// positions where the subpattern occurs.
//
// Layout / algorithm
//  * the patterns of a pass (<= SCAN_S, <= SCAN_NODES trie nodes) form a trie over their first min(len, 8) bytes.
//    Nodes exist from depth 3 on (depth = bytes matched); a node knows whether a pattern ends there (its id) and whether
//    it has children.  Patterns longer than 8 bytes hang off their depth-8 node as a chain and are verified byte by byte.
//  * shared memory per CTA (one bulk copy, ScanTables):
//      root[16384]  u32, indexed by a 14-bit hash of the first THREE bytes; the entry holds the other 10 bits of the
//                   (bijective) hash and the id of the depth-3 node, so one load decides EXACTLY whether a position can
//                   start an occurrence.  Two pattern prefixes that hash to the same entry make it "ambiguous": such
//                   positions resolve their prefix through the edge table instead (a few percent of the prefixes).
//      edge[16384]  u32, two-choice (cuckoo) hash of (parent node, next byte) -> child node; the key is 25 bits, the
//                   hash a bijection on 25 bits, so (slot, 12-bit check) identifies the key: a lookup is two independent
//                   loads and two compares, never a walk and never a verify.  Also holds the pseudo keys of ambiguous
//                   prefixes and of 1- and 2-byte patterns.
//      info[8192]   u32 per node: depth, "a pattern ends here", "has children", pattern id.
//    plus one work queue per warp and the count lines of the CTA's two row groups.
//  * every warp owns one contiguous chunk of the shard ("row") and walks it 256 positions per step: lane l holds bytes
//    [8l, 8l+12) of the step in registers (one coalesced 8-byte load + a shuffle) and looks its 8 positions up in
//    root[] -- ~9 instructions and one shared-memory load per position.  Positions that hit become work ITEMS
//    (position | node), appended to the warp's queue in position order by a prefix sum, no divergent loop.
//  * the queue is drained in rounds of 32 items, every lane busy with ONE item: read the node, emit an occurrence if
//    a pattern ends there, look up the child for the next text byte and append the child as a new item.  Every item
//    costs one emit and one lookup whatever the pattern lengths are; there is no per-class loop with idle lanes and
//    no re-hash of a whole window per (position, length) pair.  Items of one depth stay in position order (FIFO, lane
//    order), a pattern ends at exactly one depth, so every list receives its positions ascending.
//  * an occurrence is counted in the group's shared-memory counters and appended to the row's hit stream
//    (SCAN_STREAM); the stream is later scattered into the lists by scan_scatter_kernel.  The legacy modes
//    (SCAN_COUNT / SCAN_FILL: per-row count matrix by global reductions, second scan writing at the row cursors) serve
//    passes without stream memory and rows that find more occurrences than their stream holds.
//  * measured alternatives that lost (profiles/): one queue entry per (position, length class) pair probed
//    in a hash table of whole keys (3.4 warp instructions per text byte, 0.81 ms for BASELINE config 2), per-position
//    entries with a class loop in the drain (idle lanes), probing in place without a queue (lane imbalance),
//    global reductions into a per-row count matrix (1.2 cycles of the load/store pipe per LANE).
//  * algorithmic bytes per pass: shard bytes read + 4 B per occurrence written.
#pragma once
#include <algorithm>
#include <string>

#include "common.cuh"
#include "tma.cuh"

namespace vlgb {

constexpr int SCAN_S = 2048;       // patterns per pass (11-bit ids)
constexpr int SCAN_NODES = 8192;   // trie nodes per pass (13-bit ids)
constexpr int SCAN_ROOT = 16384;   // root entries
constexpr int SCAN_EDGE = 16384;   // edge slots, two halves of 8192
constexpr int SCAN_EDGE_MAX = 6600;  // entries the host puts into the edge table (40 % load: cuckoo settles)
constexpr int SCAN_WARPS = 32;     // warps per CTA (one CTA per SM)
constexpr int SCAN_PPL = 8;        // positions per lane per step
constexpr int SCAN_STEP = 32 * SCAN_PPL;
constexpr int SCAN_F2_WORDS = 64;  // 2048 bits
constexpr int SCAN_F1_WORDS = 8;   // 256 bits
constexpr int SCAN_QCAP = 256;     // items (u32) of a warp's circular work queue
constexpr int SCAN_RB = 64;        // rows per row-block in the legacy count-matrix scan
constexpr uint32_t SCAN_MAX_CHUNK = 65280;  // positions per row: an item keeps its position in 16 bits
static_assert(SCAN_S <= 2048, "pattern ids are 11 bits");

// All three tables are laid out so that a lookup is "multiply, shift, load, xor-and, compare":
// root entry: bits 19..31 node id | bit 18 ambiguous (several prefixes share the entry: resolve through the edge
//             table) | bits 8..17 the 10 check bits of the prefix hash | bits 0..7 zero (0xFF = empty, 0xAA = ambiguous:
//             the hash has zeros there, so neither ever compares equal)
// edge entry: bits 19..31 child node id | bits 7..18 the 12 check bits of the key hash | bits 0..6 zero (0x7F = empty)
// node info : bits 0..10 pattern id, bit 11 a pattern ends here (depth 8: head of the chain of patterns that share
//             these 8 bytes), bit 12 has children, bits 13..16 depth
// work item : bits 0..15 position - row start | bits 16..17 kind (0 node / prefix, 1 one-byte, 2 two-byte pattern)
//             | bit 18 ambiguous prefix | bits 19..31 node id   -- (root entry & 0xFFFC0000) | position is an item,
//             and so is (edge entry & 0xFFF80000) | position
constexpr uint32_t ROOT_EMPTY = 0xFFu, ROOT_AMB = 0x400AAu, ROOT_AMB_BIT = 1u << 18, EDGE_EMPTY = 0x7Fu;
constexpr uint32_t NODE_TERM = 1u << 11, NODE_CHILD = 1u << 12;
// edge keys (25 bits): (node << 8 | byte) for real edges; pseudo keys for what the root table cannot resolve
constexpr uint32_t KEY_SHORT1 = 1u << 21;  // | b0            -> node of the 1-byte pattern
constexpr uint32_t KEY_SHORT2 = 1u << 22;  // | b0 | b1 << 8  -> node of the 2-byte pattern
constexpr uint32_t KEY_PREFIX3 = 1u << 24; // | 24-bit prefix -> depth-3 node of an ambiguous prefix
constexpr uint32_t ITEM_SHORT1 = 1u << 16, ITEM_SHORT2 = 2u << 16, ITEM_SPECIAL = 7u << 16;

struct ScanTables {  // one pass; same layout on host, in global memory and in shared memory
    uint32_t root[SCAN_ROOT];
    uint32_t edge[SCAN_EDGE];
    uint32_t info[SCAN_NODES];
    uint32_t f2[SCAN_F2_WORDS];  // 2-byte patterns: bitmap over an 11-bit hash of the two bytes (candidates only)
    uint32_t f1[SCAN_F1_WORDS];  // 1-byte patterns: bitmap over the byte
    // multipliers of the three bijective hashes, pre-shifted (the host re-seeds when the cuckoo placement does not settle)
    uint32_t mul_root;  // odd << 8: (prefix24 * mul_root) has the 24-bit bijection in bits 8..31
    uint32_t mul_e1;    // odd << 7: (key25 * mul_e)  has the 25-bit bijection in bits 7..31
    uint32_t mul_e2;
    uint32_t pad;
};

// root: h = w * mul_root -- slot = top 14 bits, check = the next 10, low 8 bits zero (w's top byte falls out)
__host__ __device__ inline uint32_t root_slot(uint32_t h) { return h >> 18; }
__host__ __device__ inline uint32_t root_word(uint32_t h, uint32_t node) { return (node << 19) | (h & 0x3FF00u); }
// edge: h = key * mul_e -- slot = top 13 bits (inside one half of the table), check = the next 12, low 7 bits zero
__host__ __device__ inline uint32_t edge_slot(uint32_t h) { return h >> 19; }
__host__ __device__ inline uint32_t edge_word(uint32_t h, uint32_t child) { return (child << 19) | (h & 0x7FF80u); }
__host__ __device__ inline uint32_t scan_h2(uint32_t w) { return ((w & 0xFFFFu) * 0x9E3779B1u) >> 21; }

struct ScanArgs {
    const uint8_t* text;     // shard start (device, 16-byte aligned), >= 16 readable bytes after valid_len
    uint32_t valid_len;      // bytes of the shard that may be matched (owned + halo)
    uint32_t owned_len;      // occurrences at positions >= owned_len lie in the halo: counted per pattern in `halo`
    uint32_t chunk;          // positions per row (multiple of SCAN_STEP, <= SCAN_MAX_CHUNK)
    uint32_t rows;
    uint32_t s_act;          // patterns in this pass
    uint32_t s_stride;       // row stride of the count matrix (multiple of 32)
    uint32_t has_short;      // the pass holds 1- or 2-byte patterns
    uint32_t has_long;       // the pass holds patterns of 8 bytes and more
    uint32_t align;          // bit j set = position j of a lane's eight is symbol-aligned (all ones for byte texts;
                             // W-byte symbols start at multiples of W)
    const ScanTables* tables;
    const uint32_t* long_len;  // [s_act] full length (patterns of >= 8 bytes only)
    const uint32_t* long_off;  // [s_act] offset into long_blob
    const uint32_t* long_next; // [s_act] next pattern with the same first 8 bytes (NONE32 = end of chain)
    const uint8_t* long_blob;
    uint32_t* matrix;        // LEGACY passes only: [rows][s_stride] counts (zeroed by the host first) / write cursors
                             // inside each list (the re-scan advances them)
    uint32_t* gmatrix;       // [groups][s_stride], one row per GROUP of SCAN_GROUP_ROWS consecutive rows: occurrence
                             // counts (stream pass out), then the group's write cursor inside each list
    const uint64_t* list_base;  // pass 2: [s_act] first word of every list inside `out`
    uint32_t* out;           // pass 2: positions (u32, shard-local)
    // hit stream: the stream pass records every occurrence of a row as (position - row start) | pid << 16, in the
    // order it was found (ascending per pattern), so that pass 2 only has to scatter it (scan_scatter_kernel).
    uint32_t* stream;        // [rows][stream_cap] or nullptr
    uint32_t stream_cap;
    uint32_t* row_hits;      // [rows] occurrences found in the row (stream pass out)
    uint32_t* overflow;      // number of rows with row_hits > stream_cap (stream pass out)
    uint32_t* halo;          // [s_stride] occurrences per pattern in the halo (counting passes out; zeroed by the host).
                             // The halo is the last ~1 KB of a shard: list length - halo count = the OWNED prefix of the
                             // list, known with the list sizes, no second look at the lists
    uint32_t only_overflow;  // unused (kept for the ABI of the args block)
};

static_assert(sizeof(ScanTables) % 16 == 0 && sizeof(ScanTables) < (1u << 20), "one bulk copy (mbarrier tx count < 2^20)");
constexpr int SCAN_GROUP_ROWS = 16;  // rows (= warps) that share one line of the count matrix: half a scan CTA
constexpr int SCAN_CTA_GROUPS = SCAN_WARPS / SCAN_GROUP_ROWS;
constexpr size_t SCAN_SMEM = sizeof(ScanTables) + (size_t)SCAN_WARPS * SCAN_QCAP * sizeof(uint32_t) + 16 +
                             (size_t)SCAN_CTA_GROUPS * SCAN_S * sizeof(uint32_t);
static_assert(SCAN_SMEM <= 227 * 1024, "scan CTA: tables + queues + group counters must fit the 227 KB of an SM");
// what a scan pass does with an occurrence
constexpr int SCAN_STREAM = 0;  // default: count it in the group's shared-memory counters, append it to the row's hit stream
constexpr int SCAN_COUNT = 1;   // legacy pass 1: count it in the per-row matrix (global reduction), no stream
constexpr int SCAN_FILL = 2;    // legacy pass 2: write it at the row's cursor

// four text bytes at an arbitrary position (two aligned words, one funnel shift)
__device__ __forceinline__ uint32_t text_u32(const uint8_t* text, uint32_t p)
{
    const uint32_t* t32 = reinterpret_cast<const uint32_t*>(text + (p & ~3u));
    return __funnelshift_r(__ldg(t32), __ldg(t32 + 1), (p & 3u) * 8u);
}

// two-choice lookup of a 25-bit key: the matching edge entry (child node in bits 19..31) or 0
__device__ __forceinline__ uint32_t edge_lookup(const ScanTables* T, uint32_t key, uint32_t m1, uint32_t m2)
{
    const uint32_t h1 = key * m1, h2 = key * m2;
    const uint32_t e1 = T->edge[edge_slot(h1)], e2 = T->edge[SCAN_EDGE / 2 + edge_slot(h2)];
    uint32_t r = 0;
    if (((e2 ^ h2) & 0x7FFFFu) == 0u) r = e2;
    if (((e1 ^ h1) & 0x7FFFFu) == 0u) r = e1;
    return r;
}

// what happens to the occurrences a round found: pid != NONE32 on the lanes that hold one, p = its position
template <int MODE>
__device__ __forceinline__ void scan_emit(const ScanArgs& a, uint32_t hits, uint32_t pid, uint32_t p, uint32_t rel, uint32_t* mrow,
                                          uint32_t lane, uint32_t lt_mask, uint32_t* stream_row, uint32_t& nh)
{
    if (MODE == SCAN_STREAM) {
        // bump the (group, pattern) count -- a shared-memory atomic, where the global reduction of the legacy pass
        // costs ~1.2 cycles of the load/store pipe per LANE (tools/microbench) -- and append the hit to the row's
        // stream at its place in lane order
        const uint32_t at = nh + __popc(hits & lt_mask);
        if (pid != NONE32) {
            atomicAdd(mrow + pid, 1u);
            if (at < a.stream_cap) stream_row[at] = rel | (pid << 16);
            if (p >= a.owned_len) atomicAdd(a.halo + pid, 1u);  // a handful per shard
        }
        nh += __popc(hits);
    } else if (MODE == SCAN_COUNT) {
        if (pid != NONE32) {
            atomicAdd(mrow + pid, 1u);  // fire-and-forget L2 reduction
            if (p >= a.owned_len) atomicAdd(a.halo + pid, 1u);
        }
    } else {
        // legacy pass 2: the matrix entry is the row's write cursor inside the list; rank among equal patterns of
        // the round = lane order = position order
        const uint32_t grp = __match_any_sync(0xffffffffu, pid);
        const int leader = __ffs(grp) - 1;
        uint32_t old = 0;
        if (pid != NONE32 && (int)lane == leader) old = atomicAdd(mrow + pid, (uint32_t)__popc(grp));
        old = __shfl_sync(0xffffffffu, old, leader);
        if (pid != NONE32) a.out[__ldg(a.list_base + pid) + old + __popc(grp & lt_mask)] = p;
    }
}

// One drain round: the first n (<= 32) queue items starting at qhead, one per lane, all lanes converged.
//   A. node items read their node; if a pattern ends there the occurrence is emitted (warp-wide, in lane order: items
//      of one depth sit in the queue in position order, and a pattern ends at exactly one depth).  Depth-8 nodes carry
//      the chain of patterns that share their 8 bytes: every member is verified against the text, one per emit round.
//   B. one lookup per item: the child for the next text byte (node items with children), or the pseudo key of an
//      ambiguous prefix / a short pattern.  Lanes that find a node append it to the queue as a new item.
// `slow` (warp-uniform, fixed for the row): the row ends within 8 bytes of the text's end, so occurrences are checked
// against it, or the pass holds patterns of 8 bytes and more (chains to verify).  Returns the number of items appended.
template <int MODE, bool FULL>
__device__ __forceinline__ uint32_t scan_drain_round(const ScanArgs& a, const ScanTables* T, uint32_t* q, uint32_t qhead, uint32_t n,
                                                     uint32_t qtail, uint32_t row_lo, uint32_t* mrow, uint32_t lane, uint32_t lt_mask,
                                                     uint32_t* stream_row, uint32_t& nh, uint32_t m1, uint32_t m2, bool slow)
{
    const bool active = FULL || lane < n;
    const uint32_t e = active ? q[(qhead + lane) & (SCAN_QCAP - 1)] : ITEM_SPECIAL;
    const uint32_t rel = e & 0xFFFFu, p = row_lo + rel;
    const bool special = (e & ITEM_SPECIAL) != 0u;
    const uint32_t info = special ? 0u : T->info[e >> 19];
    const uint32_t depth = (info >> 13) & 15u;
    // ---- A. emit
    const bool term = (info & NODE_TERM) != 0u;
    const uint32_t terms = __ballot_sync(0xffffffffu, term);
    if (terms) {
        if (!slow) {
            scan_emit<MODE>(a, terms, term ? (info & 0x7FFu) : NONE32, p, rel, mrow, lane, lt_mask, stream_row, nh);
        } else {
            uint32_t cand = term ? (info & 0x7FFu) : NONE32;
            for (;;) {
                uint32_t pid = NONE32;
                if (cand != NONE32) {
                    if (depth < 8u) {
                        if ((uint64_t)p + depth <= a.valid_len) pid = cand;
                        cand = NONE32;
                    } else {
                        // patterns of >= 8 bytes that share these 8 bytes: one candidate per emit round, hit or miss --
                        // lanes on the same node walk the same chain, so a pattern is emitted by all its lanes
                        // together and its list stays sorted
                        const uint32_t L = a.long_len[cand];
                        bool ok = (uint64_t)p + L <= a.valid_len;
                        if (ok) {
                            const uint8_t* pat = a.long_blob + a.long_off[cand];
                            const uint8_t* txt = a.text + p;
                            for (uint32_t j = 8; j < L; j++)
                                if (__ldg(txt + j) != __ldg(pat + j)) { ok = false; break; }
                        }
                        if (ok) pid = cand;
                        cand = a.long_next[cand];
                    }
                }
                const uint32_t hits = __ballot_sync(0xffffffffu, pid != NONE32);
                if (hits) scan_emit<MODE>(a, hits, pid, p, rel, mrow, lane, lt_mask, stream_row, nh);
                if (!__any_sync(0xffffffffu, cand != NONE32)) break;  // only depth-8 lanes can still have candidates
            }
        }
    }
    // ---- B. one lookup per item: node items look for the child under the next text byte; the rare kinds (ambiguous
    // prefix, short pattern) look their first bytes up under a pseudo key -- one code path, the key is selected
    uint32_t child = 0;
    if ((info & NODE_CHILD) || (special && active)) {
        const uint32_t w = text_u32(a.text, p + depth);  // depth = 0 for the rare kinds
        uint32_t key = ((e >> 11) & 0x1FFF00u) | (w & 0xFFu);
        if (special) key = (e & ROOT_AMB_BIT) ? (KEY_PREFIX3 | (w & 0xFFFFFFu)) : (e & ITEM_SHORT1) ? (KEY_SHORT1 | (w & 0xFFu)) : (KEY_SHORT2 | (w & 0xFFFFu));
        child = edge_lookup(T, key, m1, m2);
    }
    const uint32_t found = __ballot_sync(0xffffffffu, child != 0u);
    if (child) q[(qtail + __popc(found & lt_mask)) & (SCAN_QCAP - 1)] = (child & 0xFFF80000u) | rel;
    __syncwarp();
    return __popc(found);
}

template <int MODE, bool SHORT>
__global__ void __launch_bounds__(SCAN_WARPS * 32, 1) scan_kernel(ScanArgs a)
{
    extern __shared__ __align__(128) uint8_t smem_raw[];
    ScanTables* T = reinterpret_cast<ScanTables*>(smem_raw);
    uint32_t* queues = reinterpret_cast<uint32_t*>(smem_raw + sizeof(ScanTables));
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem_raw + sizeof(ScanTables) + (size_t)SCAN_WARPS * SCAN_QCAP * sizeof(uint32_t));
    uint32_t* gcnt = reinterpret_cast<uint32_t*>(smem_raw + sizeof(ScanTables) + (size_t)SCAN_WARPS * SCAN_QCAP * sizeof(uint32_t) + 16);

    // stage the pass tables (160 KB): ONE bulk asynchronous copy issued by one thread, completion on an mbarrier --
    // the copy engine moves the bytes while the warps set up
    if (threadIdx.x == 0) {
        mbar_init(bar, 1);
        mbar_fence_init();
    }
    if (MODE == SCAN_STREAM)
        for (uint32_t i = threadIdx.x; i < (uint32_t)SCAN_CTA_GROUPS * SCAN_S; i += SCAN_WARPS * 32) gcnt[i] = 0u;
    __syncthreads();
    if (threadIdx.x == 0) {
        mbar_expect_tx(bar, (uint32_t)sizeof(ScanTables));
        bulk_g2s(T, a.tables, (uint32_t)sizeof(ScanTables), bar);
    }
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t* q = queues + warp * SCAN_QCAP;  // circular queue of work items
    mbar_wait(bar, 0);

    const uint32_t row = blockIdx.x * SCAN_WARPS + warp;
    const bool live = row < a.rows;
    if (MODE != SCAN_STREAM && !live) return;  // the stream pass meets again at the barrier before the counter flush
    // where an occurrence is counted: the group's shared-memory counters (stream pass) / the row's line of the matrix
    uint32_t* mrow = MODE == SCAN_STREAM ? gcnt + (warp / SCAN_GROUP_ROWS) * SCAN_S : a.matrix + (size_t)row * a.s_stride;

    const uint32_t lo = live ? row * a.chunk : 0u;
    const uint32_t hi = live ? (uint32_t)min((uint64_t)lo + a.chunk, (uint64_t)a.valid_len) : 0u;  // 64-bit: lo + chunk may pass 2^32 in the last row
    // occurrences need a second look when one that starts in this row can reach past the text, or when the pass holds
    // patterns of 8 bytes and more (their chains are verified byte by byte)
    const bool slow = (uint64_t)hi + 8u > a.valid_len || a.has_long != 0u;
    const uint32_t lt_mask = (1u << lane) - 1u;
    const uint32_t mul_root = T->mul_root, m1 = T->mul_e1, m2 = T->mul_e2;
    uint32_t qhead = 0, qcount = 0;  // warp-uniform
    uint32_t nh = 0;                 // occurrences recorded in the row's hit stream (stream pass)
    uint32_t* stream_row = (MODE == SCAN_STREAM && live) ? a.stream + (size_t)row * a.stream_cap : nullptr;

    auto drain_full_rounds = [&]() {
        while (qcount >= 32) {
            const uint32_t added = scan_drain_round<MODE, true>(a, T, q, qhead, 32u, qhead + qcount, lo, mrow, lane, lt_mask, stream_row, nh, m1, m2, slow);
            qhead = (qhead + 32) & (SCAN_QCAP - 1);
            qcount += added - 32;
        }
    };
    // appends one item per set bit of `m` (bit j = position j of this lane) in (lane, j) = position order; item j is
    // (rel0 + j) | hi[j].  At most SCAN_QCAP - 32 items go in at once: a denser step is taken in two halves (of the lanes).
    auto enqueue = [&](uint32_t m, uint32_t rel0, const uint32_t (&hi16)[SCAN_PPL]) {
        const uint32_t all = __popc(m);
        uint32_t incl = all;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= (uint32_t)d) incl += o;
        }
        const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
        if (total == 0) return;
        if (total <= SCAN_QCAP - 32) {
            uint32_t idx = qhead + qcount + incl - all;
#pragma unroll
            for (int j = 0; j < SCAN_PPL; j++)
                if ((m >> j) & 1u) q[idx++ & (SCAN_QCAP - 1)] = (rel0 + j) | hi16[j];
            qcount += total;
            __syncwarp();
            drain_full_rounds();
        } else {
            // lanes 0..15 first, then lanes 16..31: queue order stays position order
            for (uint32_t half = 0; half < 2; half++) {
                const uint32_t mh = (lane >> 4) == half ? m : 0u;
                const uint32_t mine = __popc(mh);
                uint32_t inc2 = mine;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const uint32_t o = __shfl_up_sync(0xffffffffu, inc2, d);
                    if (lane >= (uint32_t)d) inc2 += o;
                }
                uint32_t idx = qhead + qcount + inc2 - mine;
#pragma unroll
                for (int j = 0; j < SCAN_PPL; j++)
                    if ((mh >> j) & 1u) q[idx++ & (SCAN_QCAP - 1)] = (rel0 + j) | hi16[j];
                qcount += __shfl_sync(0xffffffffu, inc2, 31);
                __syncwarp();
                drain_full_rounds();
            }
        }
    };

    // the text is read one step ahead: a step's 8 bytes per lane are in flight while the previous step is matched
    auto load8 = [&](uint32_t at) {
        uint2 v = make_uint2(0u, 0u);
        if (at < a.valid_len) v = __ldg(reinterpret_cast<const uint2*>(a.text + at));
        return v;
    };
    uint2 own = load8(lo + 8 * lane);
    for (uint32_t base = lo; base < hi; base += SCAN_STEP) {
        const uint32_t off = base + 8 * lane;
        const uint2 ahead = load8(off + SCAN_STEP);  // the next step (the row's last step: look-ahead bytes only)
        // ---- 12 bytes per lane: own 8 + the next lane's first 4 (lane 31: the next step's first 4)
        uint32_t nx = __shfl_down_sync(0xffffffffu, own.x, 1);
        const uint32_t nx31 = __shfl_sync(0xffffffffu, ahead.x, 0);
        if (lane == 31) nx = nx31;
        uint32_t wl[SCAN_PPL];  // low 32 bits of the window at position j
        wl[0] = own.x;
        wl[1] = __funnelshift_r(own.x, own.y, 8);
        wl[2] = __funnelshift_r(own.x, own.y, 16);
        wl[3] = __funnelshift_r(own.x, own.y, 24);
        wl[4] = own.y;
        wl[5] = __funnelshift_r(own.y, nx, 8);
        wl[6] = __funnelshift_r(own.y, nx, 16);
        wl[7] = __funnelshift_r(own.y, nx, 24);
        // positions of this lane that may start an occurrence: symbol-aligned and, in the last step of the row, < hi
        // (the others belong to the next row or to nobody)
        uint32_t okpos = a.align;
        if (base + SCAN_STEP > hi) {
            const uint32_t left = off < hi ? min(hi - off, 8u) : 0u;
            okpos &= (1u << left) - 1u;
        }
        // ---- root lookup: exact on the first three bytes (the window's fourth byte falls out of the product)
        uint32_t m = 0, hi16[SCAN_PPL];
#pragma unroll
        for (int j = 0; j < SCAN_PPL; j++) {
            const uint32_t h = wl[j] * mul_root;
            const uint32_t v = T->root[root_slot(h)];
            if (((v ^ h) & 0x3FFFFu) == 0u) m |= 1u << j;   // a valid entry with my check bits
            if (v & ROOT_AMB_BIT) m |= 1u << j;             // several prefixes share the entry: resolved through the edge table
            hi16[j] = v & 0xFFFC0000u;
        }
        const uint32_t rel0 = off - lo;
        enqueue(m & okpos, rel0, hi16);
        if (SHORT) {
            uint32_t mm1 = 0, mm2 = 0;
#pragma unroll
            for (int j = 0; j < SCAN_PPL; j++) {
                const uint32_t b0 = wl[j] & 0xFFu;
                mm1 |= ((T->f1[b0 >> 5] >> (b0 & 31)) & 1u) << j;
                const uint32_t i2 = scan_h2(wl[j]);
                mm2 |= ((T->f2[i2 >> 5] >> (i2 & 31)) & 1u) << j;
            }
#pragma unroll
            for (int j = 0; j < SCAN_PPL; j++) hi16[j] = ITEM_SHORT1;
            enqueue(mm1 & okpos, rel0, hi16);
#pragma unroll
            for (int j = 0; j < SCAN_PPL; j++) hi16[j] = ITEM_SHORT2;
            enqueue(mm2 & okpos, rel0, hi16);
        }
        own = ahead;
    }
    // the tail: rounds until the queue runs dry (a round can append what it finds)
    while (qcount) {
        const uint32_t n = min(qcount, 32u);
        const uint32_t added = scan_drain_round<MODE, false>(a, T, q, qhead, n, qhead + qcount, lo, mrow, lane, lt_mask, stream_row, nh, m1, m2, slow);
        qhead = (qhead + n) & (SCAN_QCAP - 1);
        qcount += added - n;
    }
    if (MODE == SCAN_STREAM) {
        if (live && lane == 0) {
            a.row_hits[row] = nh;
            if (nh > a.stream_cap) atomicAdd(a.overflow, 1u);
        }
        // the CTA's groups publish their counts: one line of the group matrix each, plain coalesced stores
        __syncthreads();
        for (uint32_t g = 0; g < (uint32_t)SCAN_CTA_GROUPS; g++) {
            const uint32_t grp = blockIdx.x * SCAN_CTA_GROUPS + g;
            if (grp * SCAN_GROUP_ROWS >= a.rows) break;
            for (uint32_t i = threadIdx.x; i < a.s_stride; i += SCAN_WARPS * 32) a.gmatrix[(size_t)grp * a.s_stride + i] = gcnt[g * SCAN_S + i];
        }
    }
}

// Pass 2 of the stream path: no text, no tables -- every recorded occurrence goes to
//     list_base[pid] + group cursor[pid] + (occurrences of pid earlier in the group).
// One CTA per GROUP of SCAN_GROUP_ROWS consecutive rows, rows taken in order, so the cursors of all patterns stay in
// shared memory (base[]) for the whole group and every list receives the group's positions as one ascending run:
// all writes into a 32-byte sector happen within microseconds of each other and merge in L2.  (One warp per row with
// 4736 rows in flight left every sector partially written for ~1 ms: 4x the algorithmic DRAM traffic, profiles/r01b.)
//
// A row's stream is taken in blocks of SCAT_WARPS x 32 x SCAT_ROUNDS hits, one slice of whole rounds per warp:
//   1. rank + count: round by round every warp ranks the hits of its slice among equal patterns and counts them in
//      its PRIVATE 16-bit counters: all lanes read their counter, then all lanes add 1 (shared-memory atomic without
//      return value, ~1 cycle of the load/store pipe per warp instruction, tools/microbench), then all lanes read again.
//      rank = the value read first; a counter that moved by more than 1 is shared by several lanes of the round -- the
//      order in which the hardware serialises them is not defined by the ISA and positions must land in lane (= text)
//      order, so only those lanes run MATCH.ANY and add the number of earlier lanes with the same pattern (MATCH costs
//      ~2 cycles per DISTINCT value among the participating lanes, so it is kept to the few repeated patterns of a round).
//   2. the counters become exclusive prefixes over the slices; the block totals advance base[pid] at the next block.
//   3. scatter: out[base[pid] + prefix[slice][pid] + rank] = position -- independent gathers and stores.
constexpr int SCAT_WARPS = 16;
constexpr int SCAT_ROUNDS = 16;
constexpr int SCAT_BLOCK = SCAT_WARPS * 32 * SCAT_ROUNDS;  // hits per block: every counter stays below 2^15 (marker bit)
inline size_t scat_smem_bytes(uint32_t s_act)
{
    const size_t words = (s_act + 1) >> 1;
    return 2 * words * sizeof(uint64_t) + (size_t)(SCAT_WARPS + 1) * words * sizeof(uint32_t);
}
constexpr size_t SCAT_SMEM = (size_t)(SCAT_WARPS + 1) * SCAN_S * sizeof(uint16_t) + (size_t)SCAN_S * sizeof(uint64_t);  // the largest pass
static_assert(SCAT_BLOCK < (1 << 16), "16-bit counter halves");

__global__ void __launch_bounds__(SCAT_WARPS * 32) scan_scatter_kernel(ScanArgs a)
{
    extern __shared__ __align__(128) uint8_t smem_raw[];
    // shared memory is sized by the patterns of the pass (scat_smem_bytes): 2 * words cursors, (SCAT_WARPS + 1) * words
    // counter words (two patterns each): one line per slice and one for the block totals
    const uint32_t words = (a.s_act + 1) >> 1;
    uint64_t* base = reinterpret_cast<uint64_t*>(smem_raw);  // [2 * words] next free word of pid's list for this group
    uint32_t* cnt32 = reinterpret_cast<uint32_t*>(smem_raw + (size_t)2 * words * sizeof(uint64_t));  // [SCAT_WARPS][words]
    uint32_t* tot32 = cnt32 + SCAT_WARPS * words;  // [words] hits of the block per pattern
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t* cur = cnt32 + warp * words;
    const volatile uint32_t* vcur = cur;
    const uint32_t lt_mask = (1u << lane) - 1u;
    const uint32_t grp = blockIdx.x;
    const uint32_t* gm = a.gmatrix + (size_t)grp * a.s_stride;
    for (uint32_t i = threadIdx.x; i < a.s_act; i += SCAT_WARPS * 32) base[i] = __ldg(a.list_base + i) + __ldg(gm + i);
    for (uint32_t i = threadIdx.x; i < (uint32_t)(SCAT_WARPS + 1) * words; i += SCAT_WARPS * 32) cnt32[i] = 0u;
    const uint32_t row_end = min(a.rows, (grp + 1) * (uint32_t)SCAN_GROUP_ROWS);
    for (uint32_t row = grp * SCAN_GROUP_ROWS; row < row_end; row++) {
        const uint32_t nh = a.row_hits[row];
        const uint32_t* st = a.stream + (size_t)row * a.stream_cap;
        const uint32_t lo = row * a.chunk;
        for (uint32_t b0 = 0; b0 < nh; b0 += SCAT_BLOCK) {
            const uint32_t nb = min((uint32_t)SCAT_BLOCK, nh - b0);
            const uint32_t per = ((nb + SCAT_WARPS - 1) / SCAT_WARPS + 31) & ~31u;  // hits per slice, whole rounds
            const uint32_t i0 = b0 + min(nb, warp * per), i1 = b0 + min(nb, (warp + 1) * per);
            uint32_t e[SCAT_ROUNDS], rk[SCAT_ROUNDS];
#pragma unroll
            for (int r = 0; r < SCAT_ROUNDS; r++) e[r] = i0 + r * 32 + lane < i1 ? __ldg(st + i0 + r * 32 + lane) : NONE32;
            __syncthreads();  // the previous block is scattered (first block: base[] and the counters are set up)
            // the group cursors advance by the previous block's totals; then all slices start from zero (one thread per
            // counter word, the same thread for all slices)
            for (uint32_t i = threadIdx.x; i < words; i += SCAT_WARPS * 32) {
                const uint32_t t = tot32[i];
                base[2 * i] += t & 0xFFFFu;
                base[2 * i + 1] += t >> 16;
#pragma unroll
                for (int w = 0; w < SCAT_WARPS; w++) cnt32[w * words + i] = 0u;
            }
            __syncthreads();
            // ---- 1. rank of every hit inside its slice, counting as it goes: the slice's counter read BEFORE the round's
            // increments + the lanes of the round that hold the same pattern and come earlier
#pragma unroll
            for (int r = 0; r < SCAT_ROUNDS; r++) {
                rk[r] = 0;
                if (i0 + r * 32 >= i1) continue;  // uniform
                const bool valid = e[r] != NONE32;
                const uint32_t pid = e[r] >> 16, w = valid ? e[r] >> 17 : 0u, sh = (e[r] >> 12) & 16u;
                uint32_t before = valid ? (vcur[w] >> sh) & 0xFFFFu : 0u;  // every lane reads before any lane adds
                __syncwarp();
                if (valid) atomicAdd(&cur[w], 1u << sh);
                __syncwarp();
                // a counter that moved by more than 1: other lanes of the round hold the same pattern
                const bool dup = valid && ((vcur[w] >> sh) & 0xFFFFu) - before > 1u;
                const uint32_t m = __ballot_sync(0xffffffffu, dup);
                if (dup) before += __popc(__match_any_sync(m, pid) & lt_mask);
                rk[r] = before;
            }
            __syncthreads();
            // ---- 2. exclusive prefix over the slices; the block totals wait for the next block's cursor advance
            for (uint32_t i = threadIdx.x; i < words; i += SCAT_WARPS * 32) {
                uint32_t run = 0;
#pragma unroll
                for (int w = 0; w < SCAT_WARPS; w++) {
                    const uint32_t v = cnt32[w * words + i];
                    cnt32[w * words + i] = run;
                    run += v;
                }
                tot32[i] = run;
            }
            __syncthreads();
            // ---- 3. scatter: independent gathers and stores, nothing serialises the rounds any more
#pragma unroll
            for (int r = 0; r < SCAT_ROUNDS; r++) {
                if (e[r] != NONE32) {
                    const uint32_t pid = e[r] >> 16;
                    const uint32_t at = ((cur[pid >> 1] >> ((pid & 1u) * 16u)) & 0xFFFFu) + rk[r];
                    a.out[base[pid] + at] = lo + (e[r] & 0xFFFFu);
                }
            }
        }
    }
}

// gm[g][pid] (occurrences of pid in row group g) -> exclusive prefix over the groups (in place) = the group's write
// cursor inside list pid; total[pid] = list length.  32 consecutive patterns x 32 segments of groups per CTA: loads
// are coalesced over the patterns, the segments are scanned in shared memory.
__global__ void __launch_bounds__(1024) scan_groupbase_kernel(uint32_t* gm, uint32_t ngroups, uint32_t stride, uint32_t* total)
{
    __shared__ uint32_t seg_sum[32][33];
    const uint32_t pid = blockIdx.x * 32 + threadIdx.x, seg = threadIdx.y;
    const uint32_t per = (ngroups + 31) / 32;
    const uint32_t g0 = min(ngroups, seg * per), g1 = min(ngroups, g0 + per);
    uint32_t s = 0;
    if (pid < stride)
        for (uint32_t g = g0; g < g1; g++) s += gm[(size_t)g * stride + pid];
    seg_sum[seg][threadIdx.x] = s;
    __syncthreads();
    uint32_t run = 0;
    for (uint32_t k = 0; k < seg; k++) run += seg_sum[k][threadIdx.x];
    if (pid < stride) {
        for (uint32_t g = g0; g < g1; g++) {
            const uint32_t v = gm[(size_t)g * stride + pid];
            gm[(size_t)g * stride + pid] = run;
            run += v;
        }
        if (seg == 31) total[pid] = run;
    }
}

// ---- count-matrix column scans -----------------------------------------------------------------
// part[rb][pid] = sum of matrix[r][pid] over the rows of row-block rb
__global__ void scan_colsum_kernel(const uint32_t* matrix, uint32_t rows, uint32_t stride, uint32_t* part)
{
    uint32_t pid = blockIdx.x * blockDim.x + threadIdx.x;
    if (pid >= stride) return;
    uint32_t r0 = blockIdx.y * SCAN_RB, r1 = min(rows, r0 + SCAN_RB);
    uint32_t s = 0;
#pragma unroll 16
    for (uint32_t r = r0; r < r1; r++) s += matrix[(size_t)r * stride + pid];
    part[(size_t)blockIdx.y * stride + pid] = s;
}

// part -> exclusive prefix over row-blocks (in place), total[pid] = list length
__global__ void scan_colbase_kernel(uint32_t* part, uint32_t nrb, uint32_t stride, uint32_t* total)
{
    uint32_t pid = blockIdx.x * blockDim.x + threadIdx.x;
    if (pid >= stride) return;
    uint32_t run = 0;
    for (uint32_t rb = 0; rb < nrb; rb++) {
        uint32_t v = part[(size_t)rb * stride + pid];
        part[(size_t)rb * stride + pid] = run;
        run += v;
    }
    total[pid] = run;
}

// matrix[r][pid] <- part[rb][pid] + (prefix of the counts inside the row-block) = write cursor of row r
// inside list pid
__global__ void scan_rowoff_kernel(uint32_t* matrix, uint32_t rows, uint32_t stride, const uint32_t* part)
{
    uint32_t pid = blockIdx.x * blockDim.x + threadIdx.x;
    if (pid >= stride) return;
    uint32_t r0 = blockIdx.y * SCAN_RB, r1 = min(rows, r0 + SCAN_RB);
    uint32_t run = part[(size_t)blockIdx.y * stride + pid];
    for (uint32_t r = r0; r < r1; r += 16) {  // 16 independent loads in flight, then the running sum
        uint32_t c[16];
#pragma unroll
        for (int i = 0; i < 16; i++) c[i] = r + i < r1 ? matrix[(size_t)(r + i) * stride + pid] : 0u;
#pragma unroll
        for (int i = 0; i < 16; i++) {
            if (r + i < r1) matrix[(size_t)(r + i) * stride + pid] = run;
            run += c[i];
        }
    }
}

// owned[pid] = number of positions < owned_len in list pid (lists are sorted)
__global__ void scan_owned_kernel(const uint32_t* pos, const uint64_t* base, const uint32_t* total, uint32_t s_act,
                                  uint32_t owned_len, uint32_t* owned)
{
    uint32_t pid = blockIdx.x * blockDim.x + threadIdx.x;
    if (pid >= s_act) return;
    const uint32_t* L = pos + base[pid];
    uint32_t lo = 0, hi = total[pid];
    while (lo < hi) {
        uint32_t mid = lo + ((hi - lo) >> 1);
        if (L[mid] < owned_len) lo = mid + 1; else hi = mid;
    }
    owned[pid] = lo;
}

// ---- host side: pattern dictionary and pass tables ----------------------------------------------
struct PatternDict {
    uint64_t stamp = 0;                          // identity of a finished dictionary (0 = none): same stamp, same patterns
    std::vector<std::string> pats;               // distinct subpatterns of the batch
    // open addressing over (hash of the bytes) -> index in pats; no allocation per lookup
    std::vector<uint32_t> table;
    uint32_t mask = 0;
    static uint64_t hash_bytes(const uint8_t* p, uint64_t len)
    {
        uint64_t h = 0xcbf29ce484222325ull ^ len;
        uint64_t i = 0;
        for (; i + 8 <= len; i += 8) {
            uint64_t w;
            memcpy(&w, p + i, 8);
            h = (h ^ w) * 0x9E3779B97F4A7C15ull;
            h ^= h >> 29;
        }
        uint64_t w = 0;
        if (i < len) memcpy(&w, p + i, (size_t)(len - i));
        h = (h ^ w) * 0x9E3779B97F4A7C15ull;
        return h ^ (h >> 32);
    }
    void grow()
    {
        const uint32_t cap = mask ? (mask + 1) * 2 : 4096;
        table.assign(cap, NONE32);
        mask = cap - 1;
        for (uint32_t k = 0; k < pats.size(); k++) {
            uint32_t at = (uint32_t)hash_bytes(reinterpret_cast<const uint8_t*>(pats[k].data()), pats[k].size()) & mask;
            while (table[at] != NONE32) at = (at + 1) & mask;
            table[at] = k;
        }
    }
    uint32_t add(const uint8_t* p, uint64_t len)
    {
        if (pats.size() * 2 >= table.size()) grow();
        uint32_t at = (uint32_t)hash_bytes(p, len) & mask;
        while (table[at] != NONE32) {
            const std::string& s = pats[table[at]];
            if (s.size() == len && memcmp(s.data(), p, (size_t)len) == 0) return table[at];
            at = (at + 1) & mask;
        }
        const uint32_t k = (uint32_t)pats.size();
        pats.emplace_back(reinterpret_cast<const char*>(p), (size_t)len);
        table[at] = k;
        return k;
    }
};

// ---- host side: the trie of a pass and its tables ------------------------------------------------
struct PassHost {
    ScanTables tables;
    std::vector<uint32_t> long_len, long_off, long_next;
    std::string long_blob;
    uint32_t first = 0, count = 0;  // patterns [first, first+count) of the dictionary
    bool has_short = false, has_long = false;
};

// two-choice placement of (key -> child) in the edge table, moving earlier entries to their other home when both
// homes are taken.  keys[] / childs[] hold every entry placed so far, where[] the slot of each.  False = did not settle.
struct EdgeBuilder {
    std::vector<uint32_t> keys, childs;
    std::vector<int32_t> owner;  // per slot: index into keys, -1 = empty
    uint32_t m1 = 0, m2 = 0;
    void reset(uint32_t mul1, uint32_t mul2)
    {
        owner.assign(SCAN_EDGE, -1);
        m1 = mul1;
        m2 = mul2;
    }
    uint32_t home(uint32_t k, int which) const
    {
        return which == 0 ? edge_slot(keys[k] * m1) : (uint32_t)(SCAN_EDGE / 2) + edge_slot(keys[k] * m2);
    }
    bool place(uint32_t k)
    {
        uint32_t cur = k;
        int which = 0;
        for (int kicks = 0; kicks < 256; kicks++) {
            const uint32_t s0 = home(cur, which);
            if (owner[s0] < 0) { owner[s0] = (int32_t)cur; return true; }
            const uint32_t s1 = home(cur, which ^ 1);
            if (owner[s1] < 0) { owner[s1] = (int32_t)cur; return true; }
            // evict the resident of s0; it moves to ITS other home
            const uint32_t ev = (uint32_t)owner[s0];
            owner[s0] = (int32_t)cur;
            cur = ev;
            which = s0 < (uint32_t)(SCAN_EDGE / 2) ? 1 : 0;
        }
        return false;
    }
    bool settle()
    {
        for (uint32_t k = 0; k < keys.size(); k++)
            if (!place(k)) return false;
        return true;
    }
    void write(ScanTables& t) const
    {
        for (uint32_t s = 0; s < (uint32_t)SCAN_EDGE; s++) {
            if (owner[s] < 0) {
                t.edge[s] = EDGE_EMPTY;
                continue;
            }
            const uint32_t k = (uint32_t)owner[s];
            t.edge[s] = edge_word(keys[k] * (s < (uint32_t)(SCAN_EDGE / 2) ? m1 : m2), childs[k]);
        }
    }
};

// Builds the tables of the pass that starts at dictionary pattern `first`: as many of the next patterns (<= SCAN_S) as
// fit SCAN_NODES trie nodes and SCAN_EDGE_MAX edge-table entries.  Returns how many patterns it took (>= 1).
inline uint32_t build_pass(const PatternDict& d, uint32_t first, uint32_t max_count, PassHost& ph)
{
    ph.first = first;
    ph.long_len.clear();
    ph.long_off.clear();
    ph.long_next.clear();
    ph.long_blob.clear();
    ph.has_short = ph.has_long = false;
    std::vector<uint32_t> info;  // per node
    info.reserve(SCAN_NODES);
    // (kind, parent / prefix, byte) -> node id: open addressing in a table that is reused from pass to pass
    struct NodeMap {
        std::vector<uint64_t> key;  // key + 1, 0 = empty
        std::vector<uint32_t> val;
        std::vector<uint32_t> used;
        NodeMap() : key(1u << 15, 0), val(1u << 15, 0) {}
        void clear()
        {
            for (uint32_t i : used) key[i] = 0;
            used.clear();
        }
        // slot of k: the entry if present, else the empty slot where it belongs
        uint32_t slot(uint64_t k) const
        {
            uint32_t at = (uint32_t)((k * 0x9E3779B97F4A7C15ull) >> 49);
            while (key[at] && key[at] != k + 1) at = (at + 1) & ((1u << 15) - 1u);
            return at;
        }
        void put(uint32_t at, uint64_t k, uint32_t v)
        {
            key[at] = k + 1;
            val[at] = v;
            used.push_back(at);
        }
    };
    static thread_local NodeMap node_of;
    node_of.clear();
    std::vector<std::pair<uint32_t, uint32_t>> prefix3;  // (24-bit prefix, node) of every depth-3 node
    EdgeBuilder eb;
    auto new_node = [&](uint32_t depth) {
        info.push_back(depth << 13);
        return (uint32_t)info.size() - 1;
    };
    new_node(0);  // node 0 is never a child: an all-zero edge entry means "not found"
    uint32_t count = 0;
    for (; count < max_count && count < (uint32_t)SCAN_S; count++) {
        const std::string& s = d.pats[first + count];
        const uint32_t len = (uint32_t)s.size();
        // worst case this pattern adds 6 nodes and 6 edge entries (+ its prefix, should the root entry turn ambiguous)
        if (count && (info.size() + 6 > (size_t)SCAN_NODES || eb.keys.size() + prefix3.size() / 8 + 8 > (size_t)SCAN_EDGE_MAX)) break;
        auto b = [&](uint32_t j) { return (uint32_t)(uint8_t)s[j]; };
        uint32_t node;
        if (len <= 2) {
            ph.has_short = true;
            const uint32_t key = len == 1 ? (KEY_SHORT1 | b(0)) : (KEY_SHORT2 | b(0) | (b(1) << 8));
            node = new_node(len);  // distinct patterns: always a new node
            eb.keys.push_back(key);
            eb.childs.push_back(node);
        } else {
            const uint32_t p24 = b(0) | (b(1) << 8) | (b(2) << 16);
            const uint64_t k3 = (uint64_t)1 << 40 | p24;
            const uint32_t at3 = node_of.slot(k3);
            if (!node_of.key[at3]) {
                node = new_node(3);
                node_of.put(at3, k3, node);
                prefix3.emplace_back(p24, node);
            } else {
                node = node_of.val[at3];
            }
            for (uint32_t dpt = 4; dpt <= std::min<uint32_t>(len, 8); dpt++) {
                const uint32_t key = (node << 8) | b(dpt - 1);
                const uint32_t atc = node_of.slot(key);
                if (!node_of.key[atc]) {
                    const uint32_t child = new_node(dpt);
                    node_of.put(atc, key, child);
                    eb.keys.push_back(key);
                    eb.childs.push_back(child);
                    info[node] |= NODE_CHILD;
                    node = child;
                } else {
                    node = node_of.val[atc];
                }
            }
        }
        ph.long_len.push_back(0);
        ph.long_off.push_back(0);
        ph.long_next.push_back(NONE32);
        if (len >= 8) {
            ph.has_long = true;
            ph.long_len[count] = len;
            ph.long_off[count] = (uint32_t)ph.long_blob.size();
            ph.long_blob += s;
            if (info[node] & NODE_TERM) {  // append to the chain of the patterns that share these 8 bytes
                uint32_t k = info[node] & 0x7FFu;
                while (ph.long_next[k] != NONE32) k = ph.long_next[k];
                ph.long_next[k] = count;
            } else {
                info[node] |= NODE_TERM | count;
            }
        } else {
            info[node] |= NODE_TERM | count;  // a shorter pattern owns its node alone
        }
    }
    ph.count = count;
    // ---- root table: the multiplier (of a few) that leaves the fewest prefixes ambiguous
    static const uint32_t muls[8] = {0x9E3779B1u, 0x85EBCA77u, 0xC2B2AE3Du, 0x27D4EB2Fu, 0x165667B1u, 0xD3A2646Du, 0xFD7046C5u, 0xB55A4F09u};
    uint32_t best_mul = muls[0], best_amb = ~0u;
    std::vector<uint16_t> fill(SCAN_ROOT);
    for (uint32_t mul : muls) {
        std::fill(fill.begin(), fill.end(), 0);
        uint32_t amb = 0;
        for (auto& pr : prefix3) {
            uint16_t& f = fill[root_slot(pr.first * (mul << 8))];
            if (f == 1) amb += 2; else if (f > 1) amb += 1;
            if (f < 3) f++;
        }
        if (amb < best_amb) { best_amb = amb; best_mul = mul; }
        if (!amb) break;
    }
    memset(&ph.tables, 0, sizeof ph.tables);
    ph.tables.mul_root = best_mul << 8;
    std::fill(fill.begin(), fill.end(), 0);
    for (auto& pr : prefix3) {
        uint16_t& f = fill[root_slot(pr.first * ph.tables.mul_root)];
        if (f < 3) f++;
    }
    for (uint32_t i = 0; i < (uint32_t)SCAN_ROOT; i++) ph.tables.root[i] = ROOT_EMPTY;
    for (auto& pr : prefix3) {
        const uint32_t h = pr.first * ph.tables.mul_root;
        if (fill[root_slot(h)] == 1) {
            ph.tables.root[root_slot(h)] = root_word(h, pr.second);
        } else {
            ph.tables.root[root_slot(h)] = ROOT_AMB;
            eb.keys.push_back(KEY_PREFIX3 | pr.first);
            eb.childs.push_back(pr.second);
        }
    }
    for (uint32_t i = 0; i < (uint32_t)info.size(); i++) ph.tables.info[i] = info[i];
    for (uint32_t i = 0; i < count; i++) {
        const std::string& s = d.pats[first + i];
        if (s.size() == 1) {
            const uint32_t b0 = (uint8_t)s[0];
            ph.tables.f1[b0 >> 5] |= 1u << (b0 & 31);
        } else if (s.size() == 2) {
            const uint32_t i2 = scan_h2((uint32_t)(uint8_t)s[0] | ((uint32_t)(uint8_t)s[1] << 8));
            ph.tables.f2[i2 >> 5] |= 1u << (i2 & 31);
        }
    }
    // ---- edge table
    for (uint32_t seed = 0;; seed++) {
        // odd multipliers, pre-shifted by 7: key * m carries the 25-bit bijection in bits 7..31
        const uint32_t m1 = ((0x9E3779B1u + seed * 0x7F4A7C15u * 2u) | 1u) << 7, m2 = ((0x85EBCA77u + seed * 0x3C6EF372u) | 1u) << 7;
        eb.reset(m1, m2);
        if (eb.settle()) {
            ph.tables.mul_e1 = m1;
            ph.tables.mul_e2 = m2;
            eb.write(ph.tables);
            break;
        }
        if (seed > 16) {  // 40 % load, two choices: practically never; take fewer patterns rather than fail
            if (count <= 1) throw CudaError{cudaErrorUnknown, __FILE__, __LINE__};
            return build_pass(d, first, count * 3 / 4, ph);
        }
    }
    return count;
}

}  // namespace vlgb
