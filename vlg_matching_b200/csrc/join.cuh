// join.cuh — gap-bounded join of sorted occurrence lists (north-star (3)).
//
// Replaces vlg_iterator::relax()/next() (include/sdsl/vlg_index.hpp:227-291) for a whole chunk of
// queries at once, in the list formulation of SURVEY.md A.2: for level i = k-2 .. 0 of every query,
// an occurrence x of P_i is VIABLE iff the smallest viable occurrence y of P_{i+1} with
// y >= x + lo_i exists and y <= x + hi_i   (lo_i = a_i + |P_i|, hi_i = b_i + |P_i|, vlg_index.hpp:95).
//
// One "depth" d = k-1-i is processed per launch for ALL queries of the chunk that have such a level.
// The slots of a depth are the concatenation of the left lists (one segment per query), cut into tiles of
// 4096 slots:
//   probe   : raw[slot] = j = lower_bound(V_right, x + lo) if V_right[j] <= x + hi, else NONE; viable count per tile
//   prefix  : exclusive scan of the tile counts (prefix.cuh; ~1 value per 4096 slots)
//   compact : in-tile ranks by warp ballots; vpos / vnxt / vlast [rank] = x / j / last position of the tuple
// (A single-pass variant with a decoupled look-back was measured slower: probe times per tile vary widely and
// the look-back makes every wave wait for its slowest tile.)
// vseg[seg] = rank of the segment's first slot gives every query its segment inside the compacted arrays, so
// no size ever has to travel to the host between depths.
//
// Both lists are ascending and x + lo is monotone in x, so a tile of consecutive slots maps to one contiguous
// window of the right list (a merge-path partition): the window is staged in shared memory when it is small,
// else the lanes gallop from their neighbours' answers.
// Algorithmic bytes per depth: 4 B per slot (x) + 4 B per right-list element + 8 B per slot (raw out / in) +
// 16 B per viable slot (x again, vlast gather; 12 written... counted as 4 + 4 + 12).
#pragma once
#include "common.cuh"

namespace vlgb {

constexpr int JOIN_THREADS = 256;
constexpr int MAX_K = 64;  // subpatterns per query (reference grid goes to 32, num_subpatterns.config:1-5)

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
    const uint32_t* prev_vseg;   // [prev nseg + 1]
    const uint64_t* lo;          // start-to-start lower bound of this level (saturated)
    const uint64_t* hi;
    // outputs
    uint32_t* raw;               // [total_slots] index into the right list, NONE32 = not viable
    uint32_t* tile_count;        // [tiles] viable slots per tile
    const uint32_t* tile_prefix; // [tiles+1] exclusive scan of tile_count
    const uint32_t* tile_seg;    // [tiles] segment of each tile's first slot (join_tileseg_kernel)
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

constexpr int JF_THREADS = 1024;
constexpr int JF_ITEMS = 4;                          // slots per thread
constexpr int JF_TILE = JF_THREADS * JF_ITEMS;       // slots per tile: large tiles keep the look-back short
constexpr int JF_WIN = 8192;                         // right-list window staged in shared memory (elements)
constexpr int JF_WARPS = JF_THREADS / 32;

// A tile = JF_TILE consecutive slots.  Inside a tile of one segment the part of the right list that can be
// hit, R[lb(need of first slot) .. lb(need of last slot)], is staged in shared memory when it is at most
// JF_WIN elements (left and right lists of similar density) and searched there; otherwise lane 0 of every
// warp gallops from the tile's first answer and the other lanes gallop from lane 0's.
// tile_seg[t] = segment that holds slot t * JF_TILE
__global__ void join_tileseg_kernel(const uint32_t* slot_off, uint32_t nseg, uint32_t ntiles, uint32_t* tile_seg)
{
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < ntiles) tile_seg[t] = find_segment(slot_off, 0, nseg - 1, t * (uint32_t)JF_TILE);
}

__global__ void __launch_bounds__(JF_THREADS) join_probe_kernel(DepthDesc d)
{
    __shared__ uint32_t s_win[JF_WIN];
    __shared__ uint32_t s_j[2], s_count;
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_count = 0;
    const uint32_t tile = blockIdx.x;
    const uint32_t slot0 = tile * JF_TILE;
    const uint32_t nslots = min((uint32_t)JF_TILE, d.total_slots - slot0);
    const uint32_t seg_lo = __ldg(d.tile_seg + tile);  // segment of the tile's first slot (join_tileseg_kernel)
    // upper bound for the segment searches of this tile's slots: the segment of the next tile's first slot
    const uint32_t seg_hi = slot0 + nslots == d.total_slots ? d.nseg - 1 : __ldg(d.tile_seg + tile + 1);
    const bool single = __ldg(d.slot_off + seg_lo + 1) >= slot0 + nslots;  // segment seg_lo covers the whole tile

    // ---- single-segment tile: bracket the reachable part of the right list
    const uint32_t* R1 = nullptr;
    uint32_t rlen1 = 0, vbase1 = 0;
    if (single) {
        if (d.right_list) {
            R1 = d.right_list[seg_lo];
            rlen1 = d.right_len[seg_lo];
        } else {
            vbase1 = __ldg(d.prev_vseg + seg_lo);
            R1 = d.prev_vpos + vbase1;
            rlen1 = __ldg(d.prev_vseg + seg_lo + 1) - vbase1;
        }
        if (warp < 2) {  // two warps bracket the window in parallel
            const uint32_t s = warp == 0 ? slot0 : slot0 + nslots - 1;
            const uint32_t x = __ldg(d.left[seg_lo] + (s - __ldg(d.slot_off + seg_lo)));
            const uint32_t j = warp_lower_bound(R1, rlen1, sat_add64(x, d.lo[seg_lo]), lane);
            if (lane == 0) s_j[warp] = j;
        }
    }
    __syncthreads();
    const uint32_t j_first = single ? s_j[0] : 0;
    const uint32_t wlen = single ? min(rlen1, s_j[1] + 1) - j_first : 0;
    const bool use_win = single && wlen <= (uint32_t)JF_WIN;
    if (use_win) {
        for (uint32_t i = tid; i < wlen; i += JF_THREADS) s_win[i] = __ldg(R1 + j_first + i);
        __syncthreads();
    }

    // ---- probe: JF_ITEMS rounds of JOIN_THREADS consecutive slots
    uint32_t viable = 0;
#pragma unroll
    for (int it = 0; it < JF_ITEMS; it++) {
        const uint32_t slot = slot0 + it * JF_THREADS + tid;
        const bool live = slot < slot0 + nslots;
        uint32_t seg = seg_lo, x = 0, j = 0, last = 0;
        bool ok = false;
        if (use_win) {
            if (live) {
                x = __ldg(d.left[seg] + (slot - __ldg(d.slot_off + seg)));
                const uint64_t need = sat_add64(x, d.lo[seg]), limit = sat_add64(x, d.hi[seg]);
                uint32_t lo = 0, hi = wlen;
                while (lo < hi) {
                    const uint32_t mid = (lo + hi) >> 1;
                    if ((uint64_t)s_win[mid] < need) lo = mid + 1; else hi = mid;
                }
                j = j_first + lo;
                if (lo < wlen) {
                    last = s_win[lo];
                    ok = (uint64_t)last <= limit;
                }
            }
        } else {
            const uint32_t* R = R1;
            uint32_t rlen = rlen1;
            uint64_t need = 0, limit = 0;
            if (live) {
                if (!single) {
                    seg = find_segment(d.slot_off, seg_lo, seg_hi, slot);
                    if (d.right_list) {
                        R = d.right_list[seg];
                        rlen = d.right_len[seg];
                    } else {
                        const uint32_t b = __ldg(d.prev_vseg + seg);
                        R = d.prev_vpos + b;
                        rlen = __ldg(d.prev_vseg + seg + 1) - b;
                    }
                }
                x = __ldg(d.left[seg] + (slot - __ldg(d.slot_off + seg)));
                need = sat_add64(x, d.lo[seg]);
                limit = sat_add64(x, d.hi[seg]);
            }
            if (single) {  // warp-uniform branch: lane 0 gallops from the tile's first answer, the rest from lane 0's
                uint32_t j0 = 0;
                if (lane == 0 && live) j0 = gallop_lower_bound(R, j_first, rlen, need);
                j0 = __shfl_sync(0xffffffffu, j0, 0);
                if (live) j = lane == 0 ? j0 : gallop_lower_bound(R, j0, rlen, need);
            } else if (live) {
                j = full_lower_bound(R, rlen, need);
            }
            if (live && j < rlen) {
                last = __ldg(R + j);
                ok = (uint64_t)last <= limit;
            }
        }
        if (live) d.raw[slot] = ok ? j : NONE32;
        viable += __popc(__ballot_sync(0xffffffffu, ok));
    }
    if (lane == 0 && viable) atomicAdd(&s_count, viable);
    __syncthreads();
    if (tid == 0) d.tile_count[tile] = s_count;
}

// Compaction of the viable slots of a tile; tile_prefix = exclusive scan of tile_count (prefix.cuh).
__global__ void __launch_bounds__(JF_THREADS) join_compact_kernel(DepthDesc d)
{
    __shared__ uint32_t s_cnt[JF_ITEMS * JF_WARPS];  // flags per (round, warp), then their exclusive prefix
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t tile = blockIdx.x;
    const uint32_t slot0 = tile * JF_TILE;
    const uint32_t nslots = min((uint32_t)JF_TILE, d.total_slots - slot0);
    const uint32_t seg_lo = __ldg(d.tile_seg + tile);
    const uint32_t seg_hi = slot0 + nslots == d.total_slots ? d.nseg - 1 : __ldg(d.tile_seg + tile + 1);
    const bool single = __ldg(d.slot_off + seg_lo + 1) >= slot0 + nslots;
    uint32_t jv[JF_ITEMS], fl[JF_ITEMS];
#pragma unroll
    for (int it = 0; it < JF_ITEMS; it++) {
        const uint32_t slot = slot0 + it * JF_THREADS + tid;
        jv[it] = slot < slot0 + nslots ? d.raw[slot] : NONE32;
        fl[it] = __ballot_sync(0xffffffffu, jv[it] != NONE32);
        if (lane == 0) s_cnt[it * JF_WARPS + warp] = __popc(fl[it]);
    }
    __syncthreads();
    if (warp == 0) {
        constexpr int PER = JF_ITEMS * JF_WARPS / 32;  // counts per lane, consecutive
        uint32_t c[PER], v = 0;
#pragma unroll
        for (int i = 0; i < PER; i++) {
            c[i] = s_cnt[lane * PER + i];
            v += c[i];
        }
        uint32_t inc = v;
#pragma unroll
        for (int dd = 1; dd < 32; dd <<= 1) {
            const uint32_t o = __shfl_up_sync(0xffffffffu, inc, dd);
            if (lane >= (uint32_t)dd) inc += o;
        }
        uint32_t run = inc - v;
#pragma unroll
        for (int i = 0; i < PER; i++) {
            s_cnt[lane * PER + i] = run;
            run += c[i];
        }
    }
    __syncthreads();
    const uint32_t prefix = __ldg(d.tile_prefix + tile);
#pragma unroll
    for (int it = 0; it < JF_ITEMS; it++) {
        const uint32_t slot = slot0 + it * JF_THREADS + tid;
        if (slot >= slot0 + nslots) break;
        const uint32_t o = prefix + s_cnt[it * JF_WARPS + warp] + __popc(fl[it] & ((1u << lane) - 1u));
        uint32_t sg = single ? seg_lo : find_segment(d.slot_off, seg_lo, seg_hi, slot);
        const uint32_t so = __ldg(d.slot_off + sg);
        if (jv[it] != NONE32) {
            const uint32_t j = jv[it];
            d.vpos[o] = __ldg(d.left[sg] + (slot - so));
            d.vnxt[o] = j;
            d.vlast[o] = d.right_list ? __ldg(d.right_list[sg] + j) : __ldg(d.prev_vlast + __ldg(d.prev_vseg + sg) + j);
        }
        if (slot == so) {  // segment start; also covers empty segments that start at the same slot
            for (;;) {
                d.vseg[sg] = o;
                if (sg == 0 || __ldg(d.slot_off + sg - 1) != slot) break;
                sg--;
            }
        }
    }
    if (slot0 + nslots == d.total_slots && tid == 0) {  // last tile: trailing empty segments and the grand total
        const uint32_t all = __ldg(d.tile_prefix + tile + 1);
        for (uint32_t sg = d.nseg;; sg--) {
            d.vseg[sg] = all;
            if (sg == 0 || __ldg(d.slot_off + sg - 1) != d.total_slots) break;
        }
        *d.total_out = all;
    }
}

}  // namespace vlgb
