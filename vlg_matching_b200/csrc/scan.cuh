// scan.cuh — multi-pattern text scan: text shard -> one sorted occurrence list per distinct
// subpattern (north-star (2), "coalesced text-scan path").  Replaces, for a whole batch at once, what
// the reference obtains per subpattern from forward_search (suffix_array_algorithm.hpp:48-115) plus the
// in-order WT enumeration of the SA range (wt_helper.hpp:726-785): the ascending list of text
// positions where the subpattern occurs.
//
// Layout / algorithm
//  * up to SCAN_S patterns per pass; patterns are keyed by (class, first min(len,8) bytes) where
//    class = min(len, 8).  Class <= 7 keys are exact; class 8 entries (len >= 8) are verified byte by byte.
//  * shared memory per CTA: two-choice (cuckoo) hash table of the pass (SoA, two halves of SCAN_H / 2 slots, built on
//    the host so that every key sits in one of its two home slots: a probe is two independent slot reads, no walk), the prefilter
//    f3 (hashed first 3 bytes -> byte with bit c-1 set for every class c = 3..8 that has a pattern with that
//    prefix) refined by f4 (hashed first 4 bytes, classes 4..8); f1/f2 bitmaps for 1- and 2-byte patterns;
//    one work queue per warp.
//  * every warp owns one contiguous chunk of the shard ("row") and walks it 256 positions per step:
//    lane l holds bytes [8l, 8l+12) of the step in registers (one coalesced 8-byte load + a shuffle), looks
//    up the class-mask byte of its 8 positions and packs them into 64 bits.  A warp prefix sum of the
//    popcounts gives every lane its place in (POSITION, CLASS) ORDER; one short loop over the set bits
//    appends 4-byte items (position in row | (class - 1) << 16), one per probe, to the warp's circular queue.
//  * the queue is drained only in full rounds of 32 items (leftovers wait for the next step), all lanes
//    busy with one probe each: re-read the item's 8-byte window from the text (L1), mask it to the lane's
//    class length, one hash probe, then a warp-wide emit.
//  * pass 1 counts occurrences per (row, pattern) with fire-and-forget reductions into the (L2-resident)
//    count matrix and records every hit in the row's stream; the column-wise exclusive scan of the matrix
//    gives every row its write cursor per pattern; pass 2 scatters the stream (scan_scatter_kernel) -- or, for
//    rows whose stream overflowed, repeats the walk -- writing positions at cursor + rank (rank among equal
//    patterns of the round by __match_any_sync), which makes every list sorted and the result deterministic.
//  * measured alternatives that lost: per-position queue entries with a class loop in the drain (idle lanes in
//    every class visit), shared-memory row counters (16 warps/SM instead of 32), one warp per row in the scatter
//    (4x DRAM traffic), and probing every lane's candidates in place without a queue (lane imbalance: 1.44 ms
//    against 0.87 ms for the queue on BASELINE config 2).
//  * algorithmic bytes per pass: 2 x shard bytes read + 4 B per occurrence written (+ count matrix).
#pragma once
#include <algorithm>
#include <string>

#include "common.cuh"
#include "tma.cuh"

namespace vlgb {

constexpr int SCAN_S = 2048;       // patterns per pass
constexpr int SCAN_H = 8192;       // hash slots (4 B each; load factor <= 25%)
constexpr int SCAN_WARPS = 32;     // warps per CTA (one CTA per SM: the pass tables take half its shared memory)
constexpr int SCAN_PPL = 8;        // positions per lane per step
constexpr int SCAN_STEP = 32 * SCAN_PPL;
constexpr int SCAN_F3 = 65536;     // 3-byte prefilter entries (u8: bit c-1 set = class c present, c = 3..8)
constexpr int SCAN_F4 = 32768;     // 4-byte prefilter entries (u8: bit c-1 set = class c present, c = 4..8)
constexpr int SCAN_F2_WORDS = 64;  // 2048 bits
constexpr int SCAN_F1_WORDS = 8;   // 256 bits
constexpr int SCAN_QCAP = 512;     // entries (u32) of a warp's circular work queue
constexpr int SCAN_RB = 64;        // rows per row-block in the count-matrix scan
constexpr uint32_t SCAN_MAX_CHUNK = 65280;  // u16 per-row counters

// Hash slot word: 0 = empty, else  1 << 31 | tag (17 bits) << 14 | (class - 1) << 11 | local pid (11 bits),
// class = min(len, 8).  Two-choice (cuckoo) placement: a key lives in slot[h & (H/2-1)] or in
// slot[H/2 + (h2 & (H/2-1))]; a probe reads both words (independent loads) and compares the 8-byte key (by pid) only
// on a tag hit.  Patterns of class 8 that share their first 8 bytes are chained through long_next (one table entry).
constexpr uint32_t SCAN_PID_MASK = 0x7FFu;
static_assert(SCAN_S <= 2048, "pid field of the slot word is 11 bits");

struct ScanTables {  // one pass; same layout on host, in global memory and in shared memory
    uint32_t slot[SCAN_H];
    uint32_t klo[SCAN_S];   // first min(len, 8) bytes of pattern pid, zero padded
    uint32_t khi[SCAN_S];
    uint8_t f3[SCAN_F3];
    uint8_t f4[SCAN_F4];
    uint32_t f2[SCAN_F2_WORDS];
    uint32_t f1[SCAN_F1_WORDS];
    uint32_t seed;          // mixed into the hash; the host re-seeds when a cuckoo insertion does not settle
    uint32_t pad[3];
};

// 32-bit mix of (key, class): low 12 bits = first home slot, bits 13..29 = tag
__host__ __device__ inline uint32_t scan_mix(uint32_t klo, uint32_t khi, uint32_t c, uint32_t seed)
{
    uint32_t h = (klo * 0x9E3779B1u) ^ ((khi + c + seed) * 0x85EBCA77u);
    h ^= h >> 15;
    h *= 0x2C1B3C6Du;
    h ^= h >> 13;
    return h;
}
// second home slot (upper half of the table), from bits the tag does not determine
__host__ __device__ inline uint32_t scan_slot2(uint32_t h)
{
    uint32_t g = h * 0xC2B2AE35u;
    g ^= g >> 16;
    return (uint32_t)(SCAN_H / 2) + (g & (uint32_t)(SCAN_H / 2 - 1));
}
__host__ __device__ inline uint32_t scan_slot_word(uint32_t h, uint32_t c)  // pid field left 0
{
    return 0x80000000u | (((h >> 13) & 0x1FFFFu) << 14) | ((c - 1u) << 11);
}
__host__ __device__ inline uint32_t scan_h3(uint32_t w) { return ((w & 0xFFFFFFu) * 0x9E3779B1u) >> 16; }
__host__ __device__ inline uint32_t scan_h4(uint32_t w) { return (w * 0x85EBCA77u) >> 17; }
__host__ __device__ inline uint32_t scan_h2(uint32_t w) { return ((w & 0xFFFFu) * 0x9E3779B1u) >> 21; }

struct ScanArgs {
    const uint8_t* text;     // shard start (device, 16-byte aligned), >= 16 readable bytes after valid_len
    uint32_t valid_len;      // bytes of the shard that may be matched (owned + halo)
    uint32_t chunk;          // positions per row (multiple of SCAN_STEP, <= SCAN_MAX_CHUNK)
    uint32_t rows;
    uint32_t s_act;          // patterns in this pass
    uint32_t s_stride;       // row stride of the count matrix (multiple of 32)
    uint32_t has_short;      // the pass holds 1- or 2-byte patterns
    uint32_t align0, align1; // byte masks of the positions j = 0..3 / 4..7 of a lane that are symbol-aligned
                             // (all ones for byte texts; W-byte symbols start at multiples of W)
    const ScanTables* tables;
    const uint32_t* long_len;  // [s_act] full length (class 8 only: len >= 8)
    const uint32_t* long_off;  // [s_act] offset into long_blob
    const uint32_t* long_next; // [s_act] next pattern of class 8 with the same first 8 bytes (NONE32 = end of chain)
    const uint8_t* long_blob;
    uint32_t* matrix;        // LEGACY passes only: [rows][s_stride] counts (zeroed by the host first) / write cursors
                             // inside each list (the re-scan advances them)
    uint32_t* gmatrix;       // [groups][s_stride], one row per GROUP of SCAN_GROUP_ROWS consecutive rows: occurrence
                             // counts (stream pass out), then the group's write cursor inside each list
    const uint64_t* list_base;  // pass 2: [s_act] first word of every list inside `out`
    uint32_t* out;           // pass 2: positions (u32, shard-local)
    // hit stream: pass 1 records every occurrence of a row as (position - row start) | pid << 16, in the order
    // it was found (ascending per pattern), so that pass 2 only has to scatter it (scan_scatter_kernel).  A row
    // with more than stream_cap occurrences is re-scanned by the FILL kernel instead.
    uint32_t* stream;        // [rows][stream_cap] or nullptr
    uint32_t stream_cap;
    uint32_t* row_hits;      // [rows] occurrences found in the row (pass 1 out)
    uint32_t* overflow;      // number of rows with row_hits > stream_cap (pass 1 out)
    uint32_t only_overflow;  // FILL kernel: skip the rows the scatter kernel handles
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

// One drain round: the first n (<= 32) queue entries starting at qhead, one per lane, all lanes converged.
// Entry = (position - row start) | (class - 1) << 16: ONE probe, so a lane never idles while the warp visits a
// class it does not carry.  The 8-byte window of the position is re-read from the text (L1-resident: the
// warp loaded that line moments ago), masked to the lane's class length and looked up in the hash table (one
// 4-byte slot word per probe; the key itself is compared only on a tag hit).  Hits are ranked among equal
// patterns in lane order = (position, class) order.  A pattern belongs to exactly one class and the queue is
// FIFO per class, so every list receives its positions in ascending order.
template <int MODE>
__device__ __forceinline__ void scan_drain_round(const ScanArgs& a, const ScanTables* T, const uint32_t* q,
                                                 uint32_t qhead, uint32_t n, uint32_t row_lo, uint32_t* mrow,
                                                 uint32_t lane, uint32_t lt_mask, uint32_t* stream_row, uint32_t& nh)
{
    uint32_t p = 0, c = 1, klo = 0, khi = 0;
    bool more = lane < n;
    if (more) {
        const uint32_t e = q[(qhead + lane) & (SCAN_QCAP - 1)];
        p = row_lo + (e & 0xFFFFu);
        c = ((e >> 16) & 7u) + 1u;
        const uint32_t* t32 = reinterpret_cast<const uint32_t*>(a.text + (p & ~3u));
        const uint32_t w0 = __ldg(t32), w1 = __ldg(t32 + 1), w2 = __ldg(t32 + 2);
        const uint32_t sh = (p & 3u) * 8u;
        const uint32_t wlo = __funnelshift_r(w0, w1, sh), whi = __funnelshift_r(w1, w2, sh);
        // bytes [0, c) of the window: 64-bit mask 2^(8c) - 1 split in two words (c = 8: all ones)
        const uint32_t mlo = c >= 4 ? 0xFFFFFFFFu : ((1u << (8 * c)) - 1u);
        const uint32_t mhi = c >= 8 ? 0xFFFFFFFFu : (c > 4 ? ((1u << (8 * (c - 4))) - 1u) : 0u);
        klo = wlo & mlo;
        khi = whi & mhi;
    }
    // ---- one probe: both home slots (independent loads), the key itself only on a tag hit
    uint32_t cand = NONE32;
    if (more) {
        const uint32_t h = scan_mix(klo, khi, c, T->seed);
        const uint32_t want = scan_slot_word(h, c);
        const uint32_t s1 = T->slot[h & (uint32_t)(SCAN_H / 2 - 1)], s2 = T->slot[scan_slot2(h)];
        if ((s1 & ~SCAN_PID_MASK) == want) {
            const uint32_t k = s1 & SCAN_PID_MASK;
            if (T->klo[k] == klo && T->khi[k] == khi) cand = k;
        }
        if (cand == NONE32 && (s2 & ~SCAN_PID_MASK) == want) {
            const uint32_t k = s2 & SCAN_PID_MASK;
            if (T->klo[k] == klo && T->khi[k] == khi) cand = k;
        }
        more = cand != NONE32;
    }
    for (;;) {
        uint32_t pid = NONE32;
        if (more) {
            bool ok;
            if (c < 8) {
                ok = (uint64_t)p + c <= a.valid_len;
                more = false;  // classes 1..7: the key is the pattern
            } else {
                // class 8: the patterns that share these 8 bytes are chained; one candidate per emit, hit or miss --
                // lanes holding the same key walk the same chain, so a pattern is emitted by all its lanes together
                // and its list stays sorted
                const uint32_t L = a.long_len[cand];
                ok = (uint64_t)p + L <= a.valid_len;
                if (ok) {
                    const uint8_t* pat = a.long_blob + a.long_off[cand];
                    const uint8_t* txt = a.text + p;
                    for (uint32_t j = 8; j < L; j++)
                        if (__ldg(txt + j) != __ldg(pat + j)) { ok = false; break; }
                }
            }
            if (ok) pid = cand;
            if (c >= 8) {
                cand = a.long_next[cand];
                more = cand != NONE32;
            }
        }
        // ---- warp-wide emit (all 32 lanes take part)
        const uint32_t hits = __ballot_sync(0xffffffffu, pid != NONE32);
        if (hits) {
            if (MODE == SCAN_STREAM) {
                // bump the (group, pattern) count -- a shared-memory atomic: ~1 cycle of the SM's load/store pipe per
                // warp instruction, where the global reduction of the legacy pass costs ~1.2 cycles per LANE
                // (tools/microbench, profiles/) -- and append the hit to the row's stream at its place in
                // lane = (position, class) order
                if (pid != NONE32) atomicAdd(mrow + pid, 1u);
                const uint32_t at = nh + __popc(hits & lt_mask);
                if (pid != NONE32 && at < a.stream_cap) stream_row[at] = (p - row_lo) | (pid << 16);
                nh += __popc(hits);
            } else if (MODE == SCAN_COUNT) {
                if (pid != NONE32) atomicAdd(mrow + pid, 1u);  // fire-and-forget L2 reduction
            } else {
                // pass 2 (rows without a usable stream): the matrix entry is the row's write cursor inside the
                // list; rank among equal patterns of the round = lane order = position order
                const uint32_t grp = __match_any_sync(0xffffffffu, pid);
                const int leader = __ffs(grp) - 1;
                uint32_t old = 0;
                if (pid != NONE32 && (int)lane == leader) old = atomicAdd(mrow + pid, (uint32_t)__popc(grp));
                old = __shfl_sync(0xffffffffu, old, leader);
                if (pid != NONE32) a.out[__ldg(a.list_base + pid) + old + __popc(grp & lt_mask)] = p;
            }
        }
        if (!__any_sync(0xffffffffu, more)) break;  // only class-8 lanes can still have candidates
    }
}

template <int MODE, bool SHORT>
__global__ void __launch_bounds__(SCAN_WARPS * 32, 1) scan_kernel(ScanArgs a)
{
    extern __shared__ __align__(128) uint8_t smem_raw[];
    ScanTables* T = reinterpret_cast<ScanTables*>(smem_raw);
    uint32_t* queues = reinterpret_cast<uint32_t*>(smem_raw + sizeof(ScanTables));
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem_raw + sizeof(ScanTables) + (size_t)SCAN_WARPS * SCAN_QCAP * sizeof(uint32_t));
    uint32_t* gcnt = reinterpret_cast<uint32_t*>(smem_raw + sizeof(ScanTables) + (size_t)SCAN_WARPS * SCAN_QCAP * sizeof(uint32_t) + 16);

    // stage the pass tables (144 KB): ONE bulk asynchronous copy issued by one thread, completion on an mbarrier --
    // the copy engine moves the bytes while the warps set up (a per-thread load/store loop cost 9 trips of
    // LDG.128 + STS.128 for every one of the 1024 threads)
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
    uint32_t* q = queues + warp * SCAN_QCAP;        // circular queue of (position - row start) | (class - 1) << 16
    mbar_wait(bar, 0);

    const uint32_t row = blockIdx.x * SCAN_WARPS + warp;
    bool live = row < a.rows;
    if (MODE == SCAN_FILL && live && a.only_overflow && a.row_hits[row] <= a.stream_cap) live = false;  // scattered from the stream instead
    if (MODE != SCAN_STREAM && !live) return;  // the stream pass meets again at the barrier before the counter flush
    // where an occurrence is counted: the group's shared-memory counters (stream pass) / the row's line of the matrix
    uint32_t* mrow = MODE == SCAN_STREAM ? gcnt + (warp / SCAN_GROUP_ROWS) * SCAN_S : a.matrix + (size_t)row * a.s_stride;

    const uint32_t lo = live ? row * a.chunk : 0u;
    const uint32_t hi = live ? (uint32_t)min((uint64_t)lo + a.chunk, (uint64_t)a.valid_len) : 0u;  // 64-bit: lo + chunk may pass 2^32 in the last row
    const uint32_t lt_mask = (1u << lane) - 1u;
    uint32_t qhead = 0, qcount = 0;  // warp-uniform
    uint32_t nh = 0;                 // occurrences recorded in the row's hit stream (stream pass)
    uint32_t* stream_row = (MODE == SCAN_STREAM && live) ? a.stream + (size_t)row * a.stream_cap : nullptr;

    for (uint32_t base = lo; base < hi; base += SCAN_STEP) {
        // ---- 16 bytes per lane: own 8 + the next lane's 8
        const uint32_t off = base + 8 * lane;
        uint2 own = make_uint2(0u, 0u);
        if (off < a.valid_len) own = __ldg(reinterpret_cast<const uint2*>(a.text + off));
        uint32_t nx = __shfl_down_sync(0xffffffffu, own.x, 1);
        if (lane == 31) {
            nx = 0u;
            if (base + SCAN_STEP < a.valid_len) nx = __ldg(reinterpret_cast<const uint32_t*>(a.text + base + SCAN_STEP));
        }
        // ---- prefilter: one class-mask byte per position (bit c-1 = a pattern of class c may start here)
        uint32_t wl[SCAN_PPL];  // low 32 bits of the window at position j
        wl[0] = own.x;
        wl[1] = __funnelshift_r(own.x, own.y, 8);
        wl[2] = __funnelshift_r(own.x, own.y, 16);
        wl[3] = __funnelshift_r(own.x, own.y, 24);
        wl[4] = own.y;
        wl[5] = __funnelshift_r(own.y, nx, 8);
        wl[6] = __funnelshift_r(own.y, nx, 16);
        wl[7] = __funnelshift_r(own.y, nx, 24);
        uint32_t s0 = 0, s1 = 0;
#pragma unroll
        for (int j = 0; j < SCAN_PPL; j++) {
            // classes >= 4 must also pass the 4-byte filter: most 3-byte prefixes of frequent words are shared by
            // longer patterns that do not occur here, and every miss costs a full probe sequence
            uint32_t mj = T->f3[scan_h3(wl[j])] & (T->f4[scan_h4(wl[j])] | 0x07u);
            if (SHORT) {
                const uint32_t b0 = wl[j] & 0xFFu;
                mj |= (T->f1[b0 >> 5] >> (b0 & 31)) & 1u;
                const uint32_t i2 = scan_h2(wl[j]);
                mj |= ((T->f2[i2 >> 5] >> (i2 & 31)) & 1u) << 1;
            }
            if (j < 4) s0 |= mj << (8 * j); else s1 |= mj << (8 * (j - 4));
        }
        s0 &= a.align0;
        s1 &= a.align1;
        if (base + SCAN_STEP > hi) {  // last step of the row: positions >= hi belong to the next row (or nobody)
            const uint32_t left = off < hi ? hi - off : 0u;  // my positions j < left are inside
            if (left < 4) { s1 = 0u; s0 &= left ? (0xFFFFFFFFu >> (8 * (4 - left))) : 0u; }
            else if (left < 8) s1 &= left > 4 ? (0xFFFFFFFFu >> (8 * (8 - left))) : 0u;
        }
        // ---- one queue entry per set mask bit = one (position, class) probe, in (position, class) order.  The
        // queue starts a step with fewer than 32 leftovers, so up to SCAN_QCAP - 32 new entries go in at once
        // (the common case); a step with more probes (small alphabets: every position carries several
        // classes) is enqueued class by class, at most 256 entries at a time.  Entries of one class stay in
        // FIFO = position order either way.
        const uint32_t rel0 = off - lo;
        const uint32_t all = __popc(s0) + __popc(s1);
        uint32_t incl = all;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= (uint32_t)d) incl += o;
        }
        const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
        if (total == 0) continue;
        if (total <= SCAN_QCAP - 32) {
            uint32_t idx = qhead + qcount + incl - all;
            for (uint32_t m = s0; m; m &= m - 1) {
                const uint32_t b = __ffs(m) - 1;
                q[idx++ & (SCAN_QCAP - 1)] = (rel0 + (b >> 3)) | ((b & 7u) << 16);
            }
            for (uint32_t m = s1; m; m &= m - 1) {
                const uint32_t b = __ffs(m) - 1;
                q[idx++ & (SCAN_QCAP - 1)] = (rel0 + 4 + (b >> 3)) | ((b & 7u) << 16);
            }
            qcount += total;
            __syncwarp();
            while (qcount >= 32) {  // drain full rounds; fewer than 32 leftovers wait for the next step
                scan_drain_round<MODE>(a, T, q, qhead, 32u, lo, mrow, lane, lt_mask, stream_row, nh);
                qhead = (qhead + 32) & (SCAN_QCAP - 1);
                qcount -= 32;
            }
            __syncwarp();
        } else {
            const uint32_t present = __reduce_or_sync(0xffffffffu, s0 | s1);
            uint32_t classes = (present | (present >> 8) | (present >> 16) | (present >> 24)) & 0xFFu;
            while (classes) {
                const uint32_t cb = __ffs(classes) - 1;  // class - 1, warp-uniform
                classes &= classes - 1;
                const uint32_t t0 = (s0 >> cb) & 0x01010101u, t1 = (s1 >> cb) & 0x01010101u;
                const uint32_t mine = __popc(t0) + __popc(t1);
                uint32_t inc2 = mine;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const uint32_t o = __shfl_up_sync(0xffffffffu, inc2, d);
                    if (lane >= (uint32_t)d) inc2 += o;
                }
                uint32_t idx = qhead + qcount + inc2 - mine;
                for (uint32_t m = t0; m; m &= m - 1) q[idx++ & (SCAN_QCAP - 1)] = (rel0 + ((__ffs(m) - 1) >> 3)) | (cb << 16);
                for (uint32_t m = t1; m; m &= m - 1) q[idx++ & (SCAN_QCAP - 1)] = (rel0 + 4 + ((__ffs(m) - 1) >> 3)) | (cb << 16);
                qcount += __shfl_sync(0xffffffffu, inc2, 31);
                __syncwarp();
                while (qcount >= 32) {
                    scan_drain_round<MODE>(a, T, q, qhead, 32u, lo, mrow, lane, lt_mask, stream_row, nh);
                    qhead = (qhead + 32) & (SCAN_QCAP - 1);
                    qcount -= 32;
                }
                __syncwarp();
            }
        }
    }
    if (qcount) scan_drain_round<MODE>(a, T, q, qhead, qcount, lo, mrow, lane, lt_mask, stream_row, nh);
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
//   1. histogram: every warp counts its slice per pattern in its PRIVATE 16-bit counters (shared-memory atomics
//      without return value; measured ~1 cycle of the load/store pipe per warp instruction, tools/microbench),
//   2. the counters become exclusive prefixes over the slices; base[pid] advances by the block total,
//   3. scatter: rank among equal patterns = atomic fetch-add on the private counter.  The hardware serialises the
//      lanes of one instruction that hit the same counter, but in an order the ISA does not define, and positions
//      must land in lane (= text) order.  So after the fetch-add every lane re-reads its counter: a lane whose counter
//      moved by more than its own increment shares the pattern with other lanes of the round.  Those lanes raise a
//      marker bit in the counter (that also tells the one lane of the group that saw a difference of exactly 1),
//      and only they run MATCH.ANY to rank themselves in lane order -- MATCH costs ~2 cycles per DISTINCT value among
//      the participating lanes, so it is kept to the few repeated (frequent) patterns of a round.
constexpr int SCAT_WARPS = 16;
constexpr int SCAT_ROUNDS = 16;
constexpr int SCAT_BLOCK = SCAT_WARPS * 32 * SCAT_ROUNDS;  // hits per block: every counter stays below 2^15 (marker bit)
constexpr size_t SCAT_SMEM = (size_t)SCAT_WARPS * SCAN_S * sizeof(uint16_t) + (size_t)SCAN_S * sizeof(uint64_t);
static_assert(SCAT_BLOCK < (1 << 15), "counter halves carry a marker in bit 15");

__global__ void __launch_bounds__(SCAT_WARPS * 32) scan_scatter_kernel(ScanArgs a)
{
    extern __shared__ __align__(128) uint8_t smem_raw[];
    uint64_t* base = reinterpret_cast<uint64_t*>(smem_raw);  // [SCAN_S] next free word of pid's list for this group
    uint32_t* cnt32 = reinterpret_cast<uint32_t*>(smem_raw + (size_t)SCAN_S * sizeof(uint64_t));  // [SCAT_WARPS][SCAN_S / 2], two patterns per word
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t* cur = cnt32 + warp * (SCAN_S / 2);
    const volatile uint32_t* vcur = cur;
    const uint32_t lt_mask = (1u << lane) - 1u;
    const uint32_t words = (a.s_act + 1) >> 1;  // counter words in use
    const uint32_t grp = blockIdx.x;
    const uint32_t* gm = a.gmatrix + (size_t)grp * a.s_stride;
    for (uint32_t i = threadIdx.x; i < a.s_act; i += SCAT_WARPS * 32) base[i] = __ldg(a.list_base + i) + __ldg(gm + i);
    for (uint32_t i = threadIdx.x; i < (uint32_t)SCAT_WARPS * (SCAN_S / 2); i += SCAT_WARPS * 32) cnt32[i] = 0u;
    const uint32_t row_end = min(a.rows, (grp + 1) * (uint32_t)SCAN_GROUP_ROWS);
    for (uint32_t row = grp * SCAN_GROUP_ROWS; row < row_end; row++) {
        const uint32_t nh = a.row_hits[row];
        const uint32_t* st = a.stream + (size_t)row * a.stream_cap;
        const uint32_t lo = row * a.chunk;
        for (uint32_t b0 = 0; b0 < nh; b0 += SCAT_BLOCK) {
            const uint32_t nb = min((uint32_t)SCAT_BLOCK, nh - b0);
            const uint32_t per = ((nb + SCAT_WARPS - 1) / SCAT_WARPS + 31) & ~31u;  // hits per slice, whole rounds
            const uint32_t i0 = b0 + min(nb, warp * per), i1 = b0 + min(nb, (warp + 1) * per);
            uint32_t e[SCAT_ROUNDS];
#pragma unroll
            for (int r = 0; r < SCAT_ROUNDS; r++) e[r] = i0 + r * 32 + lane < i1 ? __ldg(st + i0 + r * 32 + lane) : NONE32;
            __syncthreads();  // the previous block is scattered (first block: base[] and the counters are set up)
            // the last slice's counters now hold the previous block's totals: the group cursors advance; then all
            // slices start from zero (one thread per counter word, the same thread for all slices)
            for (uint32_t i = threadIdx.x; i < words; i += SCAT_WARPS * 32) {
                const uint32_t t = cnt32[(SCAT_WARPS - 1) * (SCAN_S / 2) + i];
                base[2 * i] += t & 0x7FFFu;
                base[2 * i + 1] += (t >> 16) & 0x7FFFu;
#pragma unroll
                for (int w = 0; w < SCAT_WARPS; w++) cnt32[w * (SCAN_S / 2) + i] = 0u;
            }
            __syncthreads();
            // ---- 1. histogram of the slice
#pragma unroll
            for (int r = 0; r < SCAT_ROUNDS; r++)
                if (e[r] != NONE32) atomicAdd(&cur[e[r] >> 17], 1u << ((e[r] >> 12) & 16u));
            __syncthreads();
            // ---- 2. exclusive prefix over the slices
            for (uint32_t i = threadIdx.x; i < words; i += SCAT_WARPS * 32) {
                uint32_t run = 0;
#pragma unroll
                for (int w = 0; w < SCAT_WARPS; w++) {
                    const uint32_t v = cnt32[w * (SCAN_S / 2) + i];
                    cnt32[w * (SCAN_S / 2) + i] = run;
                    run += v;
                }
            }
            __syncthreads();
            // ---- 3. scatter: rank = fetch-add on the slice's counter, put into lane order where lanes share a pattern
#pragma unroll
            for (int r = 0; r < SCAT_ROUNDS; r++) {
                if (i0 + r * 32 >= i1) break;  // uniform
                const bool valid = e[r] != NONE32;
                const uint32_t pid = e[r] >> 16, w = (e[r] >> 17) & (SCAN_S / 2 - 1), sh = (e[r] >> 12) & 16u;
                uint32_t rk = 0, now = 0;
                if (valid) rk = (atomicAdd(&cur[w], 1u << sh) >> sh) & 0x7FFFu;
                __syncwarp();
                if (valid) now = (vcur[w] >> sh) & 0x7FFFu;
                bool dup = valid && now - rk > 1u;
                if (__any_sync(0xffffffffu, dup)) {
                    if (dup) atomicOr(&cur[w], 0x8000u << sh);
                    __syncwarp();
                    const bool last = valid && !dup && ((vcur[w] >> sh) & 0x8000u);
                    __syncwarp();
                    if (dup) atomicAnd(&cur[w], ~(0x8000u << sh));
                    dup |= last;
                    const uint32_t m = __ballot_sync(0xffffffffu, dup);
                    if (dup) {
                        const uint32_t g = __match_any_sync(m, pid);
                        rk = now - __popc(g) + __popc(g & lt_mask);
                    }
                    __syncwarp();
                }
                if (valid) a.out[base[pid] + rk] = lo + (e[r] & 0xFFFFu);
            }
        }
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

// Places entry `word` (tag | class | pid) of hash h into one of its two home slots, moving earlier entries to their
// other home when both are taken (cuckoo).  `hs[pid]` = hash of every placed entry.  False if it does not settle.
inline bool cuckoo_insert(ScanTables& t, std::vector<uint32_t>& hs, uint32_t word, uint32_t h)
{
    uint32_t cur_word = word, cur_h = h;
    uint32_t slot = cur_h & (uint32_t)(SCAN_H / 2 - 1);
    for (int kicks = 0; kicks < 512; kicks++) {
        if (!t.slot[slot]) {
            t.slot[slot] = cur_word;
            return true;
        }
        const uint32_t other = slot < (uint32_t)(SCAN_H / 2) ? scan_slot2(cur_h) : (cur_h & (uint32_t)(SCAN_H / 2 - 1));
        if (!t.slot[other]) {
            t.slot[other] = cur_word;
            return true;
        }
        // evict the resident of `slot`, put the current entry there, continue with the evicted one at ITS other home
        const uint32_t ev = t.slot[slot];
        t.slot[slot] = cur_word;
        cur_word = ev;
        cur_h = hs[ev & SCAN_PID_MASK];
        const uint32_t home1 = cur_h & (uint32_t)(SCAN_H / 2 - 1);
        slot = slot == home1 ? scan_slot2(cur_h) : home1;
    }
    return false;
}

struct PassHost {
    ScanTables tables;
    std::vector<uint32_t> long_len, long_off, long_next;
    std::string long_blob;
    uint32_t first = 0, count = 0;  // patterns [first, first+count) of the dictionary
    bool has_short = false;
};

inline void build_pass(const PatternDict& d, uint32_t first, uint32_t count, PassHost& ph)
{
    ph.first = first;
    ph.count = count;
    std::vector<uint32_t> hs(count, 0);
    for (uint32_t seed = 0;; seed++) {
        memset(&ph.tables, 0, sizeof ph.tables);
        ph.tables.seed = seed * 0x9E3779B9u;
        ph.long_len.assign(count, 0);
        ph.long_off.assign(count, 0);
        ph.long_next.assign(count, NONE32);
        ph.long_blob.clear();
        ph.has_short = false;
        bool settled = true;
        for (uint32_t i = 0; i < count && settled; i++) {
            const std::string& s = d.pats[first + i];
            uint32_t len = (uint32_t)s.size();
            uint32_t c = len < 8 ? len : 8;
            uint64_t key = 0;
            for (uint32_t j = 0; j < len && j < 8; j++) key |= (uint64_t)(uint8_t)s[j] << (8 * j);
            uint32_t klo = (uint32_t)key, khi = (uint32_t)(key >> 32);
            const uint32_t h = scan_mix(klo, khi, c, ph.tables.seed);
            hs[i] = h;
            ph.tables.klo[i] = klo;
            ph.tables.khi[i] = khi;
            uint32_t w32 = klo;
            if (c <= 2) ph.has_short = true;
            if (c == 1) {
                ph.tables.f1[(w32 & 0xFF) >> 5] |= 1u << (w32 & 31);
            } else if (c == 2) {
                uint32_t i2 = scan_h2(w32);
                ph.tables.f2[i2 >> 5] |= 1u << (i2 & 31);
            } else {
                ph.tables.f3[scan_h3(w32)] |= (uint8_t)(1u << (c - 1));
                if (c >= 4) ph.tables.f4[scan_h4(w32)] |= (uint8_t)(1u << (c - 1));
            }
            bool chained = false;
            if (c == 8) {
                ph.long_len[i] = len;
                ph.long_off[i] = (uint32_t)ph.long_blob.size();
                ph.long_blob += s;
                // patterns with the same first 8 bytes share ONE table entry: append to the chain of the resident
                const uint32_t want = scan_slot_word(h, c);
                const uint32_t homes[2] = {h & (uint32_t)(SCAN_H / 2 - 1), scan_slot2(h)};
                for (uint32_t hm : homes) {
                    const uint32_t w = ph.tables.slot[hm];
                    if (w && (w & ~SCAN_PID_MASK) == want) {
                        uint32_t k = w & SCAN_PID_MASK;
                        if (ph.tables.klo[k] == klo && ph.tables.khi[k] == khi) {
                            while (ph.long_next[k] != NONE32) k = ph.long_next[k];
                            ph.long_next[k] = i;
                            chained = true;
                            break;
                        }
                    }
                }
            }
            if (!chained) settled = cuckoo_insert(ph.tables, hs, scan_slot_word(h, c) | i, h);
        }
        if (settled) return;
        if (seed > 64) throw CudaError{cudaErrorUnknown, __FILE__, __LINE__};  // 2048 keys in 8192 slots: never
    }
}

}  // namespace vlgb
