// join.cuh — gap-bounded join of sorted occurrence lists (north-star (3)).
//
// Replaces vlg_iterator::relax()/next() (include/sdsl/vlg_index.hpp:227-291) for a whole chunk of
// queries at once, in the list formulation of SURVEY.md A.2: for level i = k-2 .. 0 of every query,
// an occurrence x of P_i is VIABLE iff the smallest viable occurrence y of P_{i+1} with
// y >= x + lo_i exists and y <= x + hi_i   (lo_i = a_i + |P_i|, hi_i = b_i + |P_i|, vlg_index.hpp:95).
//
// One "depth" d = k-1-i is processed per launch set for ALL queries of the chunk that have such a level.
// The slots of a depth are the concatenation of the left lists (one segment per query), cut into tiles of
// 2048 slots (256 threads x 8 consecutive slots; a tile costs a fixed ~10 us of dependent memory latency whatever its
// size, so many small tiles in flight beat few large ones):
//   probe   : j = lower_bound(V_right, x + lo); viable iff V_right[j] <= x + hi.  Viable slots are compacted
//             inside the tile (order kept) into a staging area; viable count per tile
//   prefix  : exclusive scan of the tile counts (prefix.cuh; ~1 value per 2048 slots)
//   vseg    : vseg[seg] = rank of the segment's first slot, so that every query finds its segment inside the
//             compacted arrays and no size ever has to travel to the host between depths
//   compact : vpos / vnxt / vlast [tile prefix + i] = staged x / j / last position of the tuple
// (Streaming larger windows through shared memory chunk by chunk was measured slower than galloping in place;
// a single-pass variant with a decoupled look-back was measured slower too: probe times per tile vary widely and
// the look-back makes every wave wait for its slowest tile.)
//
// Both lists are ascending and x + lo is monotone in x, so a tile of consecutive slots maps to one contiguous
// window of the right list (a merge-path partition): the window is staged in shared memory when it is small,
// else the lanes gallop from their neighbours' answers.
// Algorithmic bytes per depth: 4 B per slot (x) + 4 B per right-list element read; per VIABLE slot 12 B staged,
// 12 B read back (+ 4 B gather of the tuple end at depth > 1) and 12 B written.
#pragma once
#include "common.cuh"
#include "tma.cuh"

namespace vlgb {

constexpr int JOIN_THREADS = 256;
constexpr int MAX_K = 64;  // subpatterns per query (reference grid goes to 32, num_subpatterns.config:1-5)

// Everything a probe CTA needs to know about its tile, in one 64-byte record (join_tileseg_kernel): the CTA
// starts with two independent loads instead of a chain of five dependent ones.
struct __align__(16) JoinTile {
    const uint32_t* left;  // first slot of the tile inside its segment's left list (meaningful if single)
    const uint32_t* R;     // right list of the segment of the tile's first slot
    uint32_t rlen;
    uint32_t j_first;      // lower_bound of the tile's first need in R
    uint32_t seg;          // segment of the tile's first slot
    uint32_t single;       // that segment covers the whole tile
    uint64_t glo, ghi;     // start-to-start bounds of the segment, clamped to 2^33
    uint64_t pad[2];
};
static_assert(sizeof(JoinTile) == 64, "JoinTile is one 64-byte record");

// Per-depth description of the chunk (device arrays, nseg entries unless noted)
struct DepthDesc {
    uint32_t nseg;               // queries taking part at this depth (a prefix of the chunk: sorted by k desc)
    uint32_t total_slots;        // slot_off[nseg]
    const uint32_t* slot_off;    // [nseg+1] prefix of left-list lengths
    const uint32_t* const* left; // left list of each segment (scan output; level 0: owned prefix only)
    // right side: depth 1 -> the full scan list of the query's last subpattern
    const uint32_t* const* right_list;  // depth 1 only, else nullptr
    const uint32_t* right_len;          // depth 1 only
    // right side: depth > 1 -> the compacted viable list of the previous depth
    const uint32_t* prev_vpos;
    const uint32_t* prev_vlast;
    const uint32_t* prev_vseg;   // [prev nseg] first index of each segment inside the previous depth's arrays
    const uint32_t* prev_vlen;   // [prev nseg] its length
    const uint64_t* lo;          // start-to-start lower bound of this level (saturated)
    const uint64_t* hi;
    // outputs
    uint32_t* stage_x;           // [total_slots] viable slots of every tile, compacted inside the tile:
    uint32_t* stage_j;           //   left position / index into the right list / partner position
    uint32_t* stage_last;
    uint32_t* seg_rank;          // [nseg+1] viable slots of the segment's first tile before the segment's first slot
    uint32_t* tile_count;        // [tiles] viable slots per tile
    const uint32_t* tile_prefix; // [tiles+1] exclusive scan of tile_count
    JoinTile* tiles;             // [tiles+1] per-tile records (join_tileseg_kernel); the last one is a sentinel
    unsigned long long* dbg;     // measurement hook (VLGB_JOIN_DEBUG): per tile kind {sum of ns, tiles}, else nullptr
    uint32_t* total_out;         // number of viable slots of the depth
    uint32_t* vpos;              // [total_slots] compacted
    uint32_t* vnxt;
    uint32_t* vlast;
    uint32_t* vseg;              // [nseg+1]
};

// largest s in [lo_s, hi_s] with off[s] <= slot (off ascending, off[hi_s+1] > slot guaranteed)
__device__ __forceinline__ uint32_t find_segment(const uint32_t* off, uint32_t lo_s, uint32_t hi_s, uint32_t slot)
{
    while (lo_s < hi_s) {
        uint32_t mid = (lo_s + hi_s + 1) >> 1;
        if (__ldg(off + mid) <= slot) lo_s = mid; else hi_s = mid - 1;
    }
    return lo_s;
}

// first index in [from, len) whose value is >= need, galloping from `from` (list ascending)
__device__ __forceinline__ uint32_t gallop_lower_bound(const uint32_t* __restrict__ R, uint32_t from, uint32_t len,
                                                       uint64_t need)
{
    if (from >= len) return len;
    if ((uint64_t)__ldg(R + from) >= need) return from;
    // invariant: R[lo] < need; answer in (lo, hi]
    uint32_t lo = from, step = 1, hi;
    for (;;) {
        uint64_t nx = (uint64_t)lo + step;
        if (nx >= len) { hi = len; break; }
        if ((uint64_t)__ldg(R + (uint32_t)nx) >= need) { hi = (uint32_t)nx; break; }
        lo = (uint32_t)nx;
        step <<= 1;
    }
    while (hi - lo > 1) {
        uint32_t mid = lo + ((hi - lo) >> 1);
        if ((uint64_t)__ldg(R + mid) >= need) hi = mid; else lo = mid;
    }
    return hi;
}

__device__ __forceinline__ uint32_t full_lower_bound(const uint32_t* __restrict__ R, uint32_t len, uint64_t need)
{
    uint32_t lo = 0, hi = len;
    while (lo < hi) {
        uint32_t mid = lo + ((hi - lo) >> 1);
        if ((uint64_t)__ldg(R + mid) < need) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// lower_bound by a whole warp: 32 probes per step cut the range 33-fold (5 dependent loads for 2^24 elements
// instead of 24).  All lanes must call it with the same arguments; every lane returns the answer.
__device__ __forceinline__ uint32_t warp_lower_bound(const uint32_t* __restrict__ R, uint32_t len, uint64_t need, uint32_t lane)
{
    uint32_t lo = 0, hi = len;  // the answer lies in [lo, hi]
    while (hi - lo > 32) {
        const uint32_t step = (hi - lo + 32) / 33;
        const uint32_t idx = lo + (lane + 1) * step - 1;
        const bool ge = idx >= hi || (uint64_t)__ldg(R + idx) >= need;
        const uint32_t b = __ballot_sync(0xffffffffu, ge);
        const uint32_t f = b ? (uint32_t)__ffs(b) - 1u : 32u;  // first probe that is >= need
        const uint32_t nlo = lo + f * step;
        if (f < 32) hi = min(hi, lo + (f + 1) * step - 1);
        lo = nlo;
    }
    const uint32_t idx = lo + lane;
    const bool ge = idx >= hi || (uint64_t)__ldg(R + idx) >= need;
    const uint32_t b = __ballot_sync(0xffffffffu, ge);
    return b ? min(hi, lo + (uint32_t)__ffs(b) - 1u) : hi;
}

constexpr int JF_THREADS = 256;
constexpr int JF_ITEMS = 8;                          // CONSECUTIVE slots per thread
constexpr int JF_TILE = JF_THREADS * JF_ITEMS;       // slots per tile
constexpr int JF_WIN = 8192;                         // right-list window staged in shared memory (elements)
constexpr int JF_WARPS = JF_THREADS / 32;

// Per-tile records: segment of the tile's first slot, its lists and bounds, and the lower_bound of that slot's
// need in the segment's right list -- the tile's window into the right list starts there and ends at the next
// tile's value when that tile continues the same segment.  One thread per tile, all tiles in flight at once:
// inside the probe kernel the same search would stall a whole CTA behind a chain of dependent loads.
__global__ void join_tileseg_kernel(DepthDesc d, uint32_t ntiles)
{
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t > ntiles) return;
    JoinTile jt;
    jt.pad[0] = jt.pad[1] = 0;
    if (t == ntiles) {  // sentinel: "the next tile starts another segment"
        jt.left = jt.R = nullptr;
        jt.rlen = jt.j_first = jt.single = 0;
        jt.seg = d.nseg - 1;
        jt.glo = jt.ghi = 0;
        d.tiles[t] = jt;
        return;
    }
    const uint32_t slot = t * (uint32_t)JF_TILE;
    const uint32_t nslots = min((uint32_t)JF_TILE, d.total_slots - slot);
    const uint32_t seg = find_segment(d.slot_off, 0, d.nseg - 1, slot);
    const uint32_t so = __ldg(d.slot_off + seg);
    jt.seg = seg;
    jt.single = __ldg(d.slot_off + seg + 1) >= slot + nslots;
    if (d.right_list) {
        jt.R = d.right_list[seg];
        jt.rlen = d.right_len[seg];
    } else {
        jt.R = d.prev_vpos + __ldg(d.prev_vseg + seg);
        jt.rlen = __ldg(d.prev_vlen + seg);
    }
    jt.left = d.left[seg] + (slot - so);
    // positions are < 2^32: bounds clamped to 2^33 compare the same and cannot overflow a 64-bit add
    const uint64_t cap = 1ull << 33;
    jt.glo = d.lo[seg] < cap ? d.lo[seg] : cap;
    jt.ghi = d.hi[seg] < cap ? d.hi[seg] : cap;
    jt.j_first = full_lower_bound(jt.R, jt.rlen, (uint64_t)__ldg(jt.left) + jt.glo);
    d.tiles[t] = jt;
}

// lower_bound in shared memory, galloping from `from` (answers of consecutive slots are a few elements apart)
__device__ __forceinline__ uint32_t smem_gallop(const uint32_t* w, uint32_t from, uint32_t len, uint32_t need)
{
    if (from >= len || w[from] >= need) return min(from, len);
    uint32_t lo = from, step = 1, hi;  // w[lo] < need; answer in (lo, hi]
    for (;;) {
        const uint32_t nx = lo + step;
        if (nx >= len) { hi = len; break; }
        if (w[nx] >= need) { hi = nx; break; }
        lo = nx;
        step <<= 1;
    }
    while (hi - lo > 1) {
        const uint32_t mid = (lo + hi) >> 1;
        if (w[mid] >= need) hi = mid; else lo = mid;
    }
    return hi;
}

// x + bound as a 32-bit value: list entries are < 2^32 - 4096, so anything larger compares like 0xFFFFFFFF
__device__ __forceinline__ uint32_t clamp_add32(uint32_t x, uint64_t bound)
{
    const uint64_t v = x + bound;  // bound <= 2^33: no overflow
    return v > 0xFFFFFFFFull ? 0xFFFFFFFFu : (uint32_t)v;
}

// A tile = JF_TILE consecutive slots, JF_ITEMS consecutive ones per thread.  Inside a tile of one segment the
// part of the right list that can be hit, R[lb(need of first slot) .. lb(need of last slot)], is staged in
// shared memory when it is at most JF_WIN elements (left and right lists of similar density): a thread
// binary-searches its first slot there and gallops on for the next ones.  Otherwise (dense right list) lane 0
// of every warp gallops from the tile's first answer, the other lanes from lane 0's, every thread's later
// slots from its previous answer.
// The viable slots are compacted INSIDE the tile (order kept) into stage_{x,j,last}[tile start + rank]: only
// viable slots ever reach memory (a sixth of the slots on BASELINE config 2), join_compact_kernel then moves
// tile_count[tile] records per tile to their final place.
__global__ void __launch_bounds__(JF_THREADS, 4) join_probe_kernel(DepthDesc d)
{
    __shared__ __align__(128) uint32_t s_win_raw[JF_WIN + 8];  // + the elements before / after the window that alignment drags in
    __shared__ uint32_t s_wsum[32];
    __shared__ __align__(8) uint64_t s_bar;
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) {
        mbar_init(&s_bar, 1);
        mbar_fence_init();
    }
    __syncthreads();
    const uint32_t tile = blockIdx.x;
    const uint32_t slot0 = tile * JF_TILE;
    const uint32_t nslots = min((uint32_t)JF_TILE, d.total_slots - slot0);
    unsigned long long dbg_t0 = 0;
    if (d.dbg && tid == 0) asm volatile("mov.u64 %0, %globaltimer;" : "=l"(dbg_t0));
    const JoinTile jt = d.tiles[tile], jn = d.tiles[tile + 1];  // two independent 64-byte loads
    const uint32_t seg_lo = jt.seg;  // segment of the tile's first slot
    // upper bound for the segment searches of this tile's slots: the segment of the next tile's first slot
    const uint32_t seg_hi = jn.seg;
    const bool single = jt.single;   // segment seg_lo covers the whole tile

    // ---- single-segment tile: the reachable part of the right list, R1[j_first, j_first + wlen)
    const uint32_t* R1 = jt.R;
    const uint32_t rlen1 = jt.rlen, j_first = jt.j_first;
    uint32_t wlen = 0;
    if (single) {
        // the next tile's first need is >= every need of this tile, if it continues the segment
        const bool cont = slot0 + nslots < d.total_slots && seg_hi == seg_lo;
        const uint32_t j_end = cont ? min(rlen1, jn.j_first + 1) : rlen1;
        wlen = j_end - j_first;
    }
    const bool use_win = single && wlen <= (uint32_t)JF_WIN;
    // the window as ONE bulk asynchronous copy (TMA 1-D): source rounded out to 16-byte boundaries, `lead` elements
    // of slack in front; one thread arms the barrier and issues it, the left positions below are fetched while the
    // copy engine moves the window (was: up to 32 LDGSTS per thread)
    const uint32_t lead = use_win ? (uint32_t)((reinterpret_cast<uintptr_t>(R1 + j_first) & 15u) >> 2) : 0u;
    const uint32_t* s_win = s_win_raw + lead;
    if (use_win && wlen && tid == 0) {
        const uint32_t bytes = ((lead + wlen + 3u) & ~3u) * 4u;
        mbar_expect_tx(&s_bar, bytes);
        bulk_g2s(s_win_raw, R1 + j_first - lead, bytes, &s_bar);
    }

    // ---- probe
    const uint32_t sbase = slot0 + tid * JF_ITEMS;
    uint32_t xs[JF_ITEMS], js[JF_ITEMS], ls[JF_ITEMS], segs[JF_ITEMS];
    uint32_t okbits = 0;
    if (single) {
        const uint32_t* L = jt.left + tid * JF_ITEMS;
        const uint64_t glo = jt.glo, ghi = jt.ghi;
        const uint32_t live = sbase < slot0 + nslots ? min((uint32_t)JF_ITEMS, slot0 + nslots - sbase) : 0u;
#pragma unroll
        for (int i = 0; i < JF_ITEMS; i++) {
            xs[i] = (uint32_t)i < live ? __ldg(L + i) : 0u;
            segs[i] = seg_lo;
        }
        if (use_win) {
            if (wlen) mbar_wait(&s_bar, 0);  // CTA-uniform branch: the window has landed (visible to every waiting thread)
            // first slot: binary search; the others gallop on from their left neighbour's answer.  Dead slots of
            // the last tile carry x = 0 and are masked out of okbits at the end (no per-slot liveness branches)
            uint32_t at;
            {
                const uint32_t need = clamp_add32(xs[0], glo);
                uint32_t lo = 0, hi = wlen;
                while (lo < hi) {
                    const uint32_t mid = (lo + hi) >> 1;
                    if (s_win[mid] < need) lo = mid + 1; else hi = mid;
                }
                at = lo;
            }
#pragma unroll
            for (int i = 0; i < JF_ITEMS; i++) {
                if (i > 0) at = smem_gallop(s_win, at, wlen, clamp_add32(xs[i], glo));
                js[i] = j_first + at;
                ls[i] = s_win[min(at, (uint32_t)JF_WIN - 1)];
                if (at < wlen && ls[i] <= clamp_add32(xs[i], ghi)) okbits |= 1u << i;
            }
            okbits &= (1u << live) - 1u;
        } else {
            // dense right list (window too large to stage): the thread's slots are searched in lock-step inside the
            // tile's window, one independent load per slot and round -- dependent global loads are the cost here,
            // and eight serial gallops per thread cost eight times the latency
            const uint32_t* W = R1 + j_first;
            uint32_t lo_[JF_ITEMS], hi_[JF_ITEMS];
#pragma unroll
            for (int i = 0; i < JF_ITEMS; i++) {
                lo_[i] = 0;
                hi_[i] = (uint32_t)i < live ? wlen : 0u;
            }
            for (uint32_t span = wlen; span; span >>= 1) {  // ceil(log2(wlen + 1)) rounds settle every slot
#pragma unroll
                for (int i = 0; i < JF_ITEMS; i++) {
                    if (lo_[i] < hi_[i]) {
                        const uint32_t mid = (lo_[i] + hi_[i]) >> 1;
                        if ((uint64_t)__ldg(W + mid) < xs[i] + glo) lo_[i] = mid + 1; else hi_[i] = mid;
                    }
                }
            }
#pragma unroll
            for (int i = 0; i < JF_ITEMS; i++) {
                if ((uint32_t)i >= live) break;
                js[i] = j_first + lo_[i];
                if (lo_[i] < wlen) {
                    ls[i] = __ldg(W + lo_[i]);
                    if ((uint64_t)ls[i] <= xs[i] + ghi) okbits |= 1u << i;
                }
            }
        }
    } else {
        // tile with segment boundaries: every slot finds its own segment and searches that segment's whole right
        // list in place -- all slots of the thread in lock-step, so that their dependent loads overlap
        const uint32_t* Rs[JF_ITEMS];
        uint32_t lo_[JF_ITEMS], hi_[JF_ITEMS], needs[JF_ITEMS], limits[JF_ITEMS];
        uint32_t longest = 0;
#pragma unroll
        for (int i = 0; i < JF_ITEMS; i++) {
            const uint32_t slot = sbase + i;
            segs[i] = seg_lo;
            xs[i] = 0;
            Rs[i] = nullptr;
            lo_[i] = hi_[i] = needs[i] = limits[i] = 0;
            if (slot >= slot0 + nslots) continue;
            const uint32_t seg = find_segment(d.slot_off, seg_lo, seg_hi, slot);
            segs[i] = seg;
            if (d.right_list) {
                Rs[i] = d.right_list[seg];
                hi_[i] = d.right_len[seg];
            } else {
                Rs[i] = d.prev_vpos + __ldg(d.prev_vseg + seg);
                hi_[i] = __ldg(d.prev_vlen + seg);
            }
            xs[i] = __ldg(d.left[seg] + (slot - __ldg(d.slot_off + seg)));
            const uint64_t cap = 1ull << 33;
            needs[i] = clamp_add32(xs[i], d.lo[seg] < cap ? d.lo[seg] : cap);
            limits[i] = clamp_add32(xs[i], d.hi[seg] < cap ? d.hi[seg] : cap);
            longest = max(longest, hi_[i]);
        }
        uint32_t rl[JF_ITEMS];
#pragma unroll
        for (int i = 0; i < JF_ITEMS; i++) rl[i] = hi_[i];
        for (uint32_t span = longest; span; span >>= 1) {
#pragma unroll
            for (int i = 0; i < JF_ITEMS; i++) {
                if (lo_[i] < hi_[i]) {
                    const uint32_t mid = (lo_[i] + hi_[i]) >> 1;
                    if (__ldg(Rs[i] + mid) < needs[i]) lo_[i] = mid + 1; else hi_[i] = mid;
                }
            }
        }
#pragma unroll
        for (int i = 0; i < JF_ITEMS; i++) {
            if (sbase + i >= slot0 + nslots) break;
            js[i] = lo_[i];
            if (lo_[i] < rl[i]) {
                ls[i] = __ldg(Rs[i] + lo_[i]);
                if (ls[i] <= limits[i]) okbits |= 1u << i;
            }
        }
    }

    // ---- in-tile ranks (order kept), staged records, segment starts
    const uint32_t mine = __popc(okbits);
    uint32_t inc = mine;
#pragma unroll
    for (int dd = 1; dd < 32; dd <<= 1) {
        const uint32_t o = __shfl_up_sync(0xffffffffu, inc, dd);
        if (lane >= (uint32_t)dd) inc += o;
    }
    if (lane == 31) s_wsum[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        const uint32_t w = lane < (uint32_t)JF_WARPS ? s_wsum[lane] : 0u;
        uint32_t winc = w;
#pragma unroll
        for (int dd = 1; dd < 32; dd <<= 1) {
            const uint32_t o = __shfl_up_sync(0xffffffffu, winc, dd);
            if (lane >= (uint32_t)dd) winc += o;
        }
        s_wsum[lane] = winc - w;
        if (lane == 31) d.tile_count[tile] = winc;
    }
    __syncthreads();
    uint32_t rank = s_wsum[warp] + inc - mine;  // viable slots of the tile before this thread's first slot
    // segment starts: seg_rank[seg] = viable slots of this tile before the segment's first slot; also covers the
    // empty segments that start at the same slot.  In a single-segment tile only the first slot can be one.
    if (single) {
        if (tid == 0 && slot0 == __ldg(d.slot_off + seg_lo)) {
            for (uint32_t sg = seg_lo;; sg--) {
                d.seg_rank[sg] = 0;
                if (sg == 0 || __ldg(d.slot_off + sg - 1) != slot0) break;
            }
        }
    } else {
        uint32_t r2 = rank;
#pragma unroll
        for (int i = 0; i < JF_ITEMS; i++) {
            const uint32_t slot = sbase + i;
            if (slot >= slot0 + nslots) break;
            uint32_t sg = segs[i];
            if (slot == __ldg(d.slot_off + sg)) {
                for (;;) {
                    d.seg_rank[sg] = r2;
                    if (sg == 0 || __ldg(d.slot_off + sg - 1) != slot) break;
                    sg--;
                }
            }
            r2 += (okbits >> i) & 1u;
        }
    }
    const bool keep_last = d.right_list != nullptr;  // deeper levels gather the tuple's end later
#pragma unroll
    for (int i = 0; i < JF_ITEMS; i++) {
        if ((okbits >> i) & 1u) {  // dead slots never set their bit
            d.stage_x[slot0 + rank] = xs[i];
            d.stage_j[slot0 + rank] = js[i];
            if (keep_last) d.stage_last[slot0 + rank] = ls[i];
            rank++;
        }
    }
    if (d.dbg && tid == 0) {
        unsigned long long t1;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t1));
        const int kind = !single ? 2 : (use_win ? 0 : 1);
        atomicAdd(d.dbg + kind * 2, t1 - dbg_t0);
        atomicAdd(d.dbg + kind * 2 + 1, 1ull);
    }
}

// Moves the staged records of every tile to their final place; tile_prefix = exclusive scan of tile_count.
//   vpos = x, vnxt = index of the partner in the right list, vlast = last position of the completed tuple
constexpr int JC_THREADS = 256;
__global__ void __launch_bounds__(JC_THREADS) join_compact_kernel(DepthDesc d)
{
    const uint32_t tile = blockIdx.x;
    const uint32_t slot0 = tile * JF_TILE;
    const uint32_t base = __ldg(d.tile_prefix + tile);
    const uint32_t cnt = __ldg(d.tile_prefix + tile + 1) - base;
    if (d.right_list) {  // depth 1: the partner's position IS the tuple's last position
        for (uint32_t i = threadIdx.x; i < cnt; i += JC_THREADS) {
            d.vpos[base + i] = d.stage_x[slot0 + i];
            d.vnxt[base + i] = d.stage_j[slot0 + i];
            d.vlast[base + i] = d.stage_last[slot0 + i];
        }
    } else {
        // deeper levels: last position of the partner's tuple; the tile's records belong to the segments
        // [seg_lo, seg_hi], found per record from the final vseg (join_vseg_kernel runs first)
        const uint32_t seg_lo = d.tiles[tile].seg, seg_hi = d.tiles[tile + 1].seg;
        for (uint32_t i = threadIdx.x; i < cnt; i += JC_THREADS) {
            const uint32_t o = base + i;
            const uint32_t j = d.stage_j[slot0 + i];
            // segment of record o: largest sg with vseg[sg] <= o among the non-empty ones -- vseg is ascending
            uint32_t lo_s = seg_lo, hi_s = seg_hi;
            while (lo_s < hi_s) {
                const uint32_t mid = (lo_s + hi_s + 1) >> 1;
                if (d.vseg[mid] <= o) lo_s = mid; else hi_s = mid - 1;
            }
            d.vpos[o] = d.stage_x[slot0 + i];
            d.vnxt[o] = j;
            d.vlast[o] = __ldg(d.prev_vlast + __ldg(d.prev_vseg + lo_s) + j);
        }
    }
}

// vseg[sg] = rank of the segment's first slot in the compacted arrays; vseg[nseg] = *total_out = viable slots
__global__ void join_vseg_kernel(DepthDesc d, uint32_t ntiles)
{
    const uint32_t sg = blockIdx.x * blockDim.x + threadIdx.x;
    if (sg > d.nseg) return;
    const uint32_t all = __ldg(d.tile_prefix + ntiles);
    const uint32_t slot = sg < d.nseg ? __ldg(d.slot_off + sg) : d.total_slots;
    d.vseg[sg] = slot >= d.total_slots ? all : __ldg(d.tile_prefix + slot / JF_TILE) + d.seg_rank[sg];
    if (sg == d.nseg) *d.total_out = all;
}

// The segments of a pass are a subset of the chunk's queries (seg_map[i] = query).  Publishes where every
// query's viable list starts and how long it is: gseg / glen are indexed by QUERY and are what the next depth
// and the selection read.  `base` = viable slots other passes of the depth have already written.
__global__ void join_publish_kernel(const uint32_t* vseg_local, const uint32_t* seg_map, uint32_t nseg, const uint32_t* base,
                                    uint32_t* gseg, uint32_t* glen, uint32_t* total_out)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t b = base ? *base : 0u;
    if (i < nseg) {
        const uint32_t q = seg_map[i];
        gseg[q] = b + vseg_local[i];
        glen[q] = vseg_local[i + 1] - vseg_local[i];
    }
    if (i == nseg) *total_out = b + vseg_local[nseg];
}

// ---- right-driven join (depth 1 only) ---------------------------------------------------------------------
// When the right list is far shorter than the left one (a frequent first word followed by a rare one: four
// fifths of the join slots of BASELINE config 2), almost no left element is viable and visiting them all is
// wasted work.  Driven from the right: the left elements whose partner is y_j are exactly
//     L  intersected with  [ max(y_j - hi, y_{j-1} - lo + 1),  y_j - lo ]
// (x + lo <= y_j picks y_j or an earlier partner, x + lo > y_{j-1} excludes the earlier ones, y_j <= x + hi is
// the gap bound).  The ranges are disjoint and ascending in j, so the viable list comes out sorted and already
// compact: two binary searches in L per right element, a prefix sum over the counts, then the elements are copied.
struct RevDesc {
    uint32_t nseg;                 // right-driven segments
    uint32_t total_slots;          // slot_off[nseg] = right elements of all of them
    const uint32_t* slot_off;      // [nseg+1]
    const uint32_t* const* left;   // [nseg]
    const uint32_t* left_len;      // [nseg]
    const uint32_t* const* right;  // [nseg]
    const uint64_t* lo;            // [nseg] start-to-start bounds
    const uint64_t* hi;
    uint32_t* first;               // [total_slots] index in L of the first element paired with this right element
    uint32_t* count;               // [total_slots] how many
    const uint32_t* prefix;        // [total_slots+1] exclusive scan of count
    const uint32_t* base;          // viable slots written by the left-driven pass of the depth
    uint32_t* vpos;
    uint32_t* vnxt;
    uint32_t* vlast;
};

__global__ void __launch_bounds__(256) join_rev_count_kernel(RevDesc r)
{
    const uint32_t slot = blockIdx.x * 256 + threadIdx.x;
    if (slot >= r.total_slots) return;
    const uint32_t seg = find_segment(r.slot_off, 0, r.nseg - 1, slot);
    const uint32_t j = slot - __ldg(r.slot_off + seg);
    const uint32_t* R = r.right[seg];
    const uint32_t* L = r.left[seg];
    const uint32_t nl = r.left_len[seg];
    const uint64_t lo = r.lo[seg], hi = r.hi[seg];
    const uint64_t y = __ldg(R + j);
    uint32_t a = 0, n = 0;
    if (y >= lo) {
        const uint64_t x_hi = y - lo;                       // largest x with x + lo <= y
        uint64_t x_lo = y > hi ? y - hi : 0;                // smallest x with y <= x + hi
        if (j > 0) {
            const uint64_t yp = __ldg(R + j - 1);
            if (yp >= lo) x_lo = max(x_lo, yp - lo + 1);    // x + lo > y_{j-1}
        }
        if (x_lo <= x_hi) {
            a = full_lower_bound(L, nl, x_lo);
            const uint32_t b = full_lower_bound(L, nl, x_hi + 1);
            n = b - a;
        }
    }
    r.first[slot] = a;
    r.count[slot] = n;
}

// one warp per 32 right elements; the lanes copy each element's run of left positions together
__global__ void __launch_bounds__(256) join_rev_expand_kernel(RevDesc r)
{
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t slot = blockIdx.x * 256 + threadIdx.x;
    uint32_t a = 0, n = 0, o = 0, j = 0, y = 0;
    const uint32_t* L = nullptr;
    if (slot < r.total_slots) {
        n = r.count[slot];
        if (n) {
            const uint32_t seg = find_segment(r.slot_off, 0, r.nseg - 1, slot);
            j = slot - __ldg(r.slot_off + seg);
            a = r.first[slot];
            o = *r.base + __ldg(r.prefix + slot);
            L = r.left[seg];
            y = __ldg(r.right[seg] + j);
        }
    }
    uint32_t todo = __ballot_sync(0xffffffffu, n != 0);
    while (todo) {
        const int src = __ffs(todo) - 1;
        todo &= todo - 1;
        const uint32_t sn = __shfl_sync(0xffffffffu, n, src), sa = __shfl_sync(0xffffffffu, a, src),
                       so = __shfl_sync(0xffffffffu, o, src), sj = __shfl_sync(0xffffffffu, j, src),
                       sy = __shfl_sync(0xffffffffu, y, src);
        const uint32_t* sL = reinterpret_cast<const uint32_t*>(__shfl_sync(0xffffffffu, reinterpret_cast<unsigned long long>(L), src));
        for (uint32_t t = lane; t < sn; t += 32) {
            r.vpos[so + t] = __ldg(sL + sa + t);
            r.vnxt[so + t] = sj;
            r.vlast[so + t] = sy;
        }
    }
}

// local vseg of the right-driven pass: viable slots before each segment's first right element
__global__ void join_rev_vseg_kernel(RevDesc r, uint32_t* vseg_local)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i <= r.nseg) vseg_local[i] = __ldg(r.prefix + __ldg(r.slot_off + i));
}

}  // namespace vlgb
