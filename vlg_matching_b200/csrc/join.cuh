// join.cuh — gap-bounded join of sorted occurrence lists (north-star (3)).
//
// Replaces vlg_iterator::relax()/next() (include/sdsl/vlg_index.hpp:227-291) for a whole chunk of
// queries at once, in the list formulation of SURVEY.md A.2: for level i = k-2 .. 0 of every query,
// an occurrence x of P_i is VIABLE iff the smallest viable occurrence y of P_{i+1} with
// y >= x + lo_i exists and y <= x + hi_i   (lo_i = a_i + |P_i|, hi_i = b_i + |P_i|, vlg_index.hpp:95).
//
// One "depth" d = k-1-i is processed per launch for ALL queries of the chunk that have such a level.
// The slots of a depth are the concatenation of the left lists (one segment per query):
//   probe   : raw[slot]  = lower_bound(V_right, x + lo)  if V_right[.] <= x + hi   else NONE
//   prefix  : rank[slot] = exclusive prefix of (raw != NONE)                       (prefix.cuh)
//   compact : vpos / vnxt / vlast [rank[slot]] = x / raw / last position of the completed tuple
// vseg[seg] = rank[slot_off[seg]] gives every query its segment inside the compacted arrays, so no size
// ever has to travel to the host between depths.
//
// Both lists are ascending and x + lo is monotone in x, so neighbouring lanes probe neighbouring
// places of the right list: lane 0 of a warp does the full binary search, the other lanes gallop from
// lane 0's answer (a warp-cooperative merge step; the loads of a warp fall into one or two sectors).
// Algorithmic bytes per slot: probe 4 (x) + 4 (right) + 4 (raw); compact 4+4+4+4 read, 12 written.
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
    uint32_t* raw;               // [total_slots]
    uint32_t* rank;              // [total_slots+1]
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

__global__ void __launch_bounds__(JOIN_THREADS) join_probe_kernel(DepthDesc d)
{
    __shared__ uint32_t s_seg[2];
    const uint32_t slot0 = blockIdx.x * JOIN_THREADS;
    if (threadIdx.x < 2) {
        uint32_t s = threadIdx.x == 0 ? slot0 : min(slot0 + JOIN_THREADS - 1, d.total_slots - 1);
        s_seg[threadIdx.x] = find_segment(d.slot_off, 0, d.nseg - 1, s);
    }
    __syncthreads();
    const uint32_t slot = slot0 + threadIdx.x;
    const bool live = slot < d.total_slots;
    const uint32_t lane = threadIdx.x & 31;

    uint32_t seg = 0, rlen = 0;
    const uint32_t* R = nullptr;
    uint64_t need = 0, limit = 0;
    if (live) {
        seg = find_segment(d.slot_off, s_seg[0], s_seg[1], slot);
        const uint32_t x = __ldg(d.left[seg] + (slot - __ldg(d.slot_off + seg)));
        need = sat_add64(x, d.lo[seg]);
        limit = sat_add64(x, d.hi[seg]);
        if (d.right_list) {
            R = d.right_list[seg];
            rlen = d.right_len[seg];
        } else {
            uint32_t b = __ldg(d.prev_vseg + seg);
            R = d.prev_vpos + b;
            rlen = __ldg(d.prev_vseg + seg + 1) - b;
        }
    }
    // warp-cooperative narrowing: when the whole warp works on one segment, lane 0 searches the full
    // list and everybody else gallops from its answer.
    const uint32_t act = __ballot_sync(0xffffffffu, live);
    const uint32_t seg0 = __shfl_sync(0xffffffffu, seg, 0);
    const bool uniform = __all_sync(0xffffffffu, !live || seg == seg0) && (act & 1u);
    uint32_t j = 0;
    if (uniform) {
        uint32_t j0 = 0;
        if (lane == 0) j0 = full_lower_bound(R, rlen, need);
        j0 = __shfl_sync(0xffffffffu, j0, 0);
        if (live) j = lane == 0 ? j0 : gallop_lower_bound(R, j0, rlen, need);
    } else if (live) {
        j = full_lower_bound(R, rlen, need);
    }
    if (live) {
        bool ok = j < rlen && (uint64_t)__ldg(R + j) <= limit;
        d.raw[slot] = ok ? j : NONE32;
    }
}

__global__ void __launch_bounds__(JOIN_THREADS) join_compact_kernel(DepthDesc d)
{
    __shared__ uint32_t s_seg[2];
    const uint32_t slot0 = blockIdx.x * JOIN_THREADS;
    if (threadIdx.x < 2) {
        uint32_t s = threadIdx.x == 0 ? slot0 : min(slot0 + JOIN_THREADS - 1, d.total_slots - 1);
        s_seg[threadIdx.x] = find_segment(d.slot_off, 0, d.nseg - 1, s);
    }
    __syncthreads();
    const uint32_t slot = slot0 + threadIdx.x;
    if (slot >= d.total_slots) return;
    const uint32_t j = d.raw[slot];
    if (j == NONE32) return;
    const uint32_t seg = find_segment(d.slot_off, s_seg[0], s_seg[1], slot);
    const uint32_t o = d.rank[slot];
    d.vpos[o] = __ldg(d.left[seg] + (slot - __ldg(d.slot_off + seg)));
    d.vnxt[o] = j;
    d.vlast[o] = d.right_list ? __ldg(d.right_list[seg] + j) : __ldg(d.prev_vlast + __ldg(d.prev_vseg + seg) + j);
}

// vseg[s] = rank[slot_off[s]] for s = 0 .. nseg
__global__ void join_vseg_kernel(DepthDesc d)
{
    uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s <= d.nseg) d.vseg[s] = d.rank[d.slot_off[s]];
}

}  // namespace vlgb
