// select.cuh — greedy, non-overlapping selection of matches and result emission (north-star (4)).
//
// Replaces vlg_iterator::pull_forward()/operator++ (include/sdsl/vlg_index.hpp:253-264,349-356) in the
// formulation of SURVEY.md A.3/A.5: with V_0 the ascending viable starts of a query and
//   E(x) = last position of x's completed tuple + |P_k|        (the cursor pull_forward leaves)
// the reported matches are the orbit of  t0 = lower_bound(V_0, entry cursor)  under
//   succ(t) = lower_bound(V_0, E(V_0[t]))    (monotone, succ(t) > t).
// k = 1 quirk (wt_helper.hpp:777-784): |P| = 1 and an even (whole-text) position p gives
// E = p + 2.
//
// Slots: every query owns a TILE-aligned segment of upper-bound size (|L_0| owned part), so a tile
// never straddles two queries.  Kernels:
//   resolve : per query, where V_0 / last live (scan list for k = 1, else the deepest join level)
//   succ    : succ[slot], gallop from t+1, one thread per slot
//   tile    : exitp[slot] = first orbit element of slot that lies outside its tile (one warp per tile, segments of
//             32 resolved from the right)
//   spec    : one thread per CHUNK of 64 tiles hops tile to tile from the chunk's first element (a guess)
//   walk    : one thread per query hops from the true start; orbits of a monotone successor coalesce, so
//             inside a chunk it soon lands on an element the speculative hop also visited and jumps to that
//             chunk's recorded exit: ~chunks x (a few hops) dependent loads instead of one per tile
//   entry   : entry[tile] = first selected element in the tile (true hops before the coalescing point,
//             speculative ones after it)
//   mark    : per entered tile, one bit per slot for the orbit of entry[tile] (one warp per tile walks the staged
//             succ values) + the tile's count
//   prefix  : exclusive scan of the tile counts                             (prefix.cuh)
//   emit    : starts / tuples / exit cursor
#pragma once
#include "join.cuh"

namespace vlgb {

constexpr int SEL_TILE = 1024;  // = threads per CTA of the tile kernels
constexpr int SEL_LOG = 10;
constexpr int SEL_CHUNK_TILES = 64;  // tiles per speculative chunk
constexpr int MARK_WARPS = 8;        // tiles per CTA of the warp-per-tile kernels

struct SelDesc {
    uint32_t nq;                 // queries of the chunk
    // the selection layout is built ON THE DEVICE from the V_0 sizes the join left there (sel_layout_kernel,
    // sel_fill_kernel): no size travels to the host between join and selection
    const uint32_t* lay;         // [3]: slots (= sel_off[nq], multiple of SEL_TILE), tiles, speculative chunks
    uint32_t* sel_off;           // [nq+1], TILE-aligned prefix of the V_0 sizes
    uint32_t* tile_q;            // [tiles] query of every tile
    const uint32_t* kq;          // subpatterns of each query
    const uint32_t* last_len;    // |P_k|
    const uint64_t* entry_local; // entry cursor relative to the shard start (saturated at 0)
    // k = 1: V_0 is the owned prefix of the scan list
    const uint32_t* const* list0;
    const uint32_t* list0_len;
    // k >= 2: V_0 is the compacted level of depth k-1
    const uint32_t* const* depth_vpos;   // [MAX_K] by depth
    const uint32_t* const* depth_vnxt;
    const uint32_t* const* depth_vlast;
    const uint32_t* const* depth_vseg;   // [MAX_K][queries] first index of the query's segment at that depth
    const uint32_t* const* depth_vlen;   // [MAX_K][queries] its length
    const uint32_t* const* last_list;    // scan list of the last subpattern (tuple resolve at depth 1)
    uint64_t global_offset;      // of the shard, in BYTES (symbol offset << sym_shift)
    uint32_t sym_shift;          // log2(bytes per symbol): 0 for byte texts; reported positions are >> sym_shift
    // resolved per query
    const uint32_t** v0;
    const uint32_t** vlast;      // nullptr entry for k = 1
    uint32_t* v0_len;
    uint32_t* t0;                // first selectable index
    // per slot
    uint32_t* succ;
    uint32_t* exitp;
    uint32_t* flagbits;          // [tiles * 32] selected-element bits of every warp of an entered tile
    uint32_t* tile_count;        // [tiles] selected elements per tile (0 for tiles never entered)
    const uint32_t* tile_prefix; // [tiles + 1] exclusive scan of tile_count
    uint32_t* entry;             // [total_slots / SEL_TILE]
    // speculative chunks (SEL_CHUNK_TILES tiles of one query each)
    uint32_t* chunk_q;           // [chunks] query of the chunk
    uint32_t* chunk_tile0;       // [chunks] first tile (global tile index)
    uint32_t* q_chunk0;          // [nq] first chunk of the query
    uint32_t* spec;              // [tiles] element the speculative hop entered the tile with (NONE32: not visited)
    uint32_t* spec_exit;         // [nchunks] where the speculative hop left the chunk
    uint32_t* coalesce;          // [nchunks] first tile from which the speculative entries are the true ones
    // emission
    const uint64_t* tuple_off;   // [nq] first tuple word of each query (host prefix over count*k)
    uint64_t* starts;            // [matches of the chunk]
    uint64_t* tuples;            // nullable
    uint64_t* exit_cursor;       // [nq] whole-text cursor, pre-set to the entry cursor
    uint64_t* first_start;       // [nq] whole-text position of the first reported match, pre-set to ~0
    uint64_t* counts;            // [nq]
};

// ---- selection layout on the device ----------------------------------------------------------------------------
// Every query owns a TILE-aligned range of slots for its V_0 (so a tile never straddles two queries), cut into
// speculative chunks of SEL_CHUNK_TILES tiles.  One CTA: exclusive scans of tiles and chunks per query.
__global__ void __launch_bounds__(1024) sel_layout_kernel(SelDesc s, uint32_t* lay)
{
    __shared__ uint64_t wsum[32];
    __shared__ uint64_t carry_s;
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) carry_s = 0;
    __syncthreads();
    for (uint32_t q0 = 0; q0 < s.nq; q0 += 1024) {
        const uint32_t q = q0 + tid;
        const uint32_t tiles = q < s.nq ? (s.v0_len[q] + SEL_TILE - 1) / SEL_TILE : 0u;
        const uint32_t chunks = (tiles + SEL_CHUNK_TILES - 1) / SEL_CHUNK_TILES;
        const uint64_t mine = ((uint64_t)tiles << 32) | chunks;  // both sums stay below 2^32
        uint64_t inc = mine;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint64_t o = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= (uint32_t)d) inc += o;
        }
        if (lane == 31) wsum[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            const uint64_t w = wsum[lane];
            uint64_t winc = w;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint64_t o = __shfl_up_sync(0xffffffffu, winc, d);
                if (lane >= (uint32_t)d) winc += o;
            }
            wsum[lane] = winc - w;
        }
        __syncthreads();
        const uint64_t excl = carry_s + wsum[warp] + inc - mine;
        if (q < s.nq) {
            s.sel_off[q] = (uint32_t)(excl >> 32) * SEL_TILE;
            s.q_chunk0[q] = (uint32_t)excl;
        }
        __syncthreads();
        if (tid == 1023) carry_s = excl + mine;
        __syncthreads();
    }
    if (tid == 0) {
        const uint64_t all = carry_s;
        s.sel_off[s.nq] = (uint32_t)(all >> 32) * SEL_TILE;
        lay[0] = (uint32_t)(all >> 32) * SEL_TILE;
        lay[1] = (uint32_t)(all >> 32);
        lay[2] = (uint32_t)all;
    }
}

// tile -> query map and the chunk tables
__global__ void sel_fill_kernel(SelDesc s)
{
    const uint32_t ntiles = s.lay[1];
    for (uint32_t tile = blockIdx.x * blockDim.x + threadIdx.x; tile < ntiles; tile += gridDim.x * blockDim.x) {
        const uint32_t slot = tile * SEL_TILE;
        uint32_t lo = 0, hi = s.nq;  // last q with sel_off[q] <= slot (empty queries share their offset with the next one)
        while (lo < hi) {
            const uint32_t mid = (lo + hi + 1) >> 1;
            if (s.sel_off[mid] <= slot) lo = mid; else hi = mid - 1;
        }
        s.tile_q[tile] = lo;
        const uint32_t rel = tile - s.sel_off[lo] / SEL_TILE;
        if (rel % SEL_CHUNK_TILES == 0) {
            const uint32_t c = s.q_chunk0[lo] + rel / SEL_CHUNK_TILES;
            s.chunk_q[c] = lo;
            s.chunk_tile0[c] = tile;
        }
    }
}

__global__ void sel_resolve_kernel(SelDesc s)
{
    uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= s.nq) return;
    const uint32_t k = s.kq[q];
    const uint32_t* v;
    const uint32_t* vl = nullptr;
    uint32_t len;
    if (k == 1) {
        v = s.list0[q];
        len = s.list0_len[q];
    } else {
        const uint32_t d = k - 1;
        const uint32_t first = s.depth_vseg[d][q];
        v = s.depth_vpos[d] + first;
        vl = s.depth_vlast[d] + first;
        len = s.depth_vlen[d][q];
    }
    s.v0[q] = v;
    s.vlast[q] = vl;
    s.v0_len[q] = len;
    s.t0[q] = full_lower_bound(v, len, s.entry_local[q]);
}

__device__ __forceinline__ uint32_t sel_query_of_tile(const SelDesc& s, uint32_t slot0)
{
    return __ldg(s.tile_q + slot0 / SEL_TILE);  // host-built map: one load instead of a binary search over sel_off
}

// Pointer doubling inside one tile.  ptr[i] (shared) holds a V_0 index; indices inside
// [base, base + SEL_TILE) are followed, anything else is final.
__device__ __forceinline__ uint32_t sel_hop(const uint32_t* ptr, uint32_t p, uint32_t base)
{
    uint32_t rel = p - base;
    return rel < (uint32_t)SEL_TILE ? ptr[rel] : p;
}

// succ[] and exitp[] of one tile in ONE kernel (before: one thread per viable start galloping in global memory, L2
// hit rate 13 %, then a second kernel that read the tile's succ values back).  Persistent CTAs take tiles in turn:
// the number of tiles is known on the device only.
// E(x) is monotone in x and a match spans a few hundred bytes, so the successors of a tile's 1024 starts lie in the
// tile itself or just behind it: the CTA stages V_0[base, base + SELT_WIN) in shared memory with coalesced loads and
// every thread searches its 4 consecutive starts there (gallop from the previous answer); only a successor beyond the
// window is searched in global memory.  Then the exits: every warp resolves its block of 128 slots from the right
// (shuffle pointer jumping inside a segment of 32, as before), and a second step follows the at most 7 block exits.
constexpr int SELT_THREADS = 256;
constexpr int SELT_ITEMS = SEL_TILE / SELT_THREADS;  // consecutive slots per thread
constexpr int SELT_WIN = 3 * SEL_TILE;               // staged elements of V_0
constexpr int SELT_BLOCK = SEL_TILE / (SELT_THREADS / 32);  // slots per warp in the exit step
static_assert(SELT_BLOCK % 32 == 0, "a warp's block is whole segments");

__global__ void __launch_bounds__(SELT_THREADS) sel_succtile_kernel(SelDesc s)
{
    __shared__ uint32_t s_v[SELT_WIN];
    __shared__ uint32_t s_ptr[SEL_TILE];
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t ntiles = s.lay[1];
    for (uint32_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    __syncthreads();  // the previous tile's shared arrays are no longer in use
    const uint32_t slot0 = tile * SEL_TILE;
    const uint32_t q = sel_query_of_tile(s, slot0);
    const uint32_t base = slot0 - s.sel_off[q];  // V_0 index of the tile's first element
    const uint32_t len = s.v0_len[q];
    if (base >= len) continue;  // CTA-uniform
    const uint32_t* v = s.v0[q];
    const uint32_t* vl = s.vlast[q];
    const uint32_t ll = s.last_len[q];
    const uint32_t wn = min((uint32_t)SELT_WIN, len - base);
    for (uint32_t i = tid; i < wn; i += SELT_THREADS) s_v[i] = __ldg(v + base + i);
    const uint32_t rel0 = tid * SELT_ITEMS;
    uint32_t last[SELT_ITEMS];
#pragma unroll
    for (int i = 0; i < SELT_ITEMS; i++) last[i] = (vl && base + rel0 + i < len) ? __ldg(vl + base + rel0 + i) : 0u;
    __syncthreads();
    const bool quirk = !vl && ll == (1u << s.sym_shift);  // k = 1, one symbol: sibling skip after an even position
    const bool window_is_all = base + wn == len;          // the window reaches the end of V_0
    uint32_t from = rel0 + 1;
#pragma unroll
    for (int i = 0; i < SELT_ITEMS; i++) {
        const uint32_t rel = rel0 + i;
        uint32_t sc = 0xFFFFFFF0u;
        if (base + rel < len) {
            const uint32_t x = s_v[rel];
            uint64_t e = (uint64_t)(vl ? last[i] : x) + ll;
            if (quirk && ((((uint64_t)x + s.global_offset) >> s.sym_shift) & 1ull) == 0) e = (uint64_t)x + (2u << s.sym_shift);
            const uint32_t need = e > 0xFFFFFFFFull ? 0xFFFFFFFFu : (uint32_t)e;
            from = max(from, rel + 1);
            const uint32_t at = smem_gallop(s_v, from, wn, need);  // in [from, wn]
            from = at;
            if (at < wn || window_is_all) sc = base + at;
            else sc = gallop_lower_bound(v, base + wn, len, e);    // beyond the window (dense V_0, long matches)
        }
        s_ptr[rel] = sc;
        if (base + rel < len) s.succ[slot0 + rel] = sc;
    }
    __syncthreads();
    // ---- exits, step 1: inside the warp's block, segments of 32 from the right
    const uint32_t brel = warp * SELT_BLOCK;
    for (int seg = SELT_BLOCK / 32 - 1; seg >= 0; seg--) {
        const uint32_t segrel = brel + seg * 32;
        uint32_t p = s_ptr[segrel + lane];
#pragma unroll
        for (int r = 0; r < 5; r++) {
            const uint32_t rr = p - (base + segrel);
            const uint32_t o = __shfl_sync(0xffffffffu, p, rr & 31);
            if (rr < 32u) p = o;
        }
        const uint32_t rb = p - (base + brel);
        if (rb < (uint32_t)SELT_BLOCK) p = s_ptr[brel + rb];  // a later segment of the block: final already
        __syncwarp();
        s_ptr[segrel + lane] = p;
        __syncwarp();
    }
    __syncthreads();
    // ---- step 2: from block exit to block exit until the orbit leaves the tile (at most 7 hops)
#pragma unroll
    for (int i = 0; i < SELT_ITEMS; i++) {
        const uint32_t rel = tid + i * SELT_THREADS;  // coalesced stores
        if (base + rel >= len) continue;
        uint32_t p = s_ptr[rel];
        while (p - base < (uint32_t)SEL_TILE) p = s_ptr[p - base];
        s.exitp[slot0 + rel] = p;
    }
    }
}

__global__ void sel_spec_kernel(SelDesc s)
{
    const uint32_t nchunks = s.lay[2];
    for (uint32_t c = blockIdx.x * blockDim.x + threadIdx.x; c < nchunks; c += gridDim.x * blockDim.x) {
    const uint32_t q = s.chunk_q[c];
    const uint32_t off = s.sel_off[q];
    const uint32_t len = s.v0_len[q];
    const uint32_t first = s.chunk_tile0[c] * SEL_TILE - off;  // V_0 index of the chunk's first element
    const uint32_t end = min(len, first + SEL_CHUNK_TILES * SEL_TILE);
    uint32_t cur = max(first, s.t0[q]);  // the chunk that holds t0 starts on the true orbit
    while (cur < end) {
        s.spec[(off + cur) / SEL_TILE] = cur;
        cur = s.exitp[off + cur];
    }
    s.spec_exit[c] = cur;
    }
}

__global__ void sel_walk_kernel(SelDesc s)
{
    // one thread per query; spread over warps so that long walks do not serialise 32 short ones
    uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= s.nq) return;
    const uint32_t len = s.v0_len[q];
    const uint32_t off = s.sel_off[q];
    const uint32_t tile_first = off / SEL_TILE, c0 = s.q_chunk0[q];
    uint32_t cur = s.t0[q];
    while (cur < len) {
        const uint32_t tile = (off + cur) / SEL_TILE;
        const uint32_t c = c0 + (tile - tile_first) / SEL_CHUNK_TILES;
        if (s.spec[tile] == cur) {  // from here on the speculative hops of this chunk are the true ones
            s.coalesce[c] = tile;
            cur = s.spec_exit[c];
            continue;
        }
        s.entry[tile] = cur;
        cur = s.exitp[off + cur];
    }
}

__global__ void sel_entry_kernel(SelDesc s)
{
    const uint32_t ntiles = s.lay[1];
    for (uint32_t tile = blockIdx.x * blockDim.x + threadIdx.x; tile < ntiles; tile += gridDim.x * blockDim.x) {
        const uint32_t q = sel_query_of_tile(s, tile * SEL_TILE);
        const uint32_t c = s.q_chunk0[q] + (tile - s.sel_off[q] / SEL_TILE) / SEL_CHUNK_TILES;
        const uint32_t from = s.coalesce[c];
        if (from != NONE32 && tile >= from) s.entry[tile] = s.spec[tile];
    }
}

// One warp per entered tile: the tile's succ[] values are staged in shared memory (coalesced) and the orbit of
// entry[tile] is marked segment by segment (32 slots = one flag word).  Inside a segment the orbit is the set reachable
// from its first element under j(i) = succ(i): five rounds of "add the image of the set under j, square j" with one
// warp-wide OR-reduction and one shuffle each; the successor of the segment's last orbit element enters the next
// segment.  (Lane 0 walking the orbit alone -- one dependent shared-memory load per selected element, 3 of 32 lanes
// active -- cost 110 M warp instructions, profiles/r02p.)
// The orbit of element `e` (a V_0 index inside the tile, or NONE32: nobody enters) marked in the tile's 32 flag words;
// sc = the warp's staging block of SEL_TILE words.  All lanes of one warp; e is warp-uniform.
__device__ __forceinline__ void sel_mark_tile(const SelDesc& s, uint32_t tile, uint32_t e, uint32_t* sc, uint32_t lane)
{
    if (e == NONE32) {  // tile not entered: no selected element (the flag words stay meaningful for re-entry)
        if (lane == 0) s.tile_count[tile] = 0;
        s.flagbits[tile * 32 + lane] = 0u;
        return;
    }
    const uint32_t slot0 = tile * SEL_TILE;
    const uint32_t q = sel_query_of_tile(s, slot0);
    const uint32_t base = slot0 - s.sel_off[q];
    const uint32_t len = s.v0_len[q];
#pragma unroll 8
    for (uint32_t i = lane; i < (uint32_t)SEL_TILE; i += 32) sc[i] = base + i < len ? s.succ[slot0 + i] : 0xFFFFFFF0u;
    __syncwarp();
    const uint32_t end = min((uint32_t)SEL_TILE, len - base);  // succ = len ("no successor") may fall inside the tile
    uint32_t cur = e - base;  // warp-uniform: the orbit element that enters the next segment
    uint32_t cnt = 0, my_bits = 0;
    while (cur < end) {
        const uint32_t w = cur >> 5, segbase = w * 32;
        uint32_t j = sc[segbase + lane] - base - segbase;  // >= 32: the successor lies outside the segment (wraps to huge)
        uint32_t R = 1u << (cur & 31);
#pragma unroll
        for (int r = 0; r < 5; r++) {
            const uint32_t img = __reduce_or_sync(0xffffffffu, (((R >> lane) & 1u) && j < 32u) ? (1u << j) : 0u);
            R |= img;
            const uint32_t j2 = __shfl_sync(0xffffffffu, j, j & 31);
            if (j < 32u) j = j2;
        }
        // "no successor" is index len, which may fall inside the segment: not an element
        if (end - segbase < 32u) R &= (1u << (end - segbase)) - 1u;
        if (lane == w) my_bits = R;
        cnt += __popc(R);
        cur = sc[segbase + 31 - __clz(R)] - base;  // successor of the segment's last orbit element
    }
    s.flagbits[tile * 32 + lane] = my_bits;
    if (lane == 0) s.tile_count[tile] = cnt;
}

__global__ void __launch_bounds__(MARK_WARPS * 32) sel_mark_kernel(SelDesc s)
{
    __shared__ uint32_t s_succ[MARK_WARPS][SEL_TILE];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t ntiles = s.lay[1];
    for (uint32_t tile = blockIdx.x * MARK_WARPS + warp; tile < ntiles; tile += gridDim.x * MARK_WARPS) {
        __syncwarp();
        sel_mark_tile(s, tile, s.entry[tile], s_succ[warp], lane);
    }
}

// counts[q] = selected elements in the query's tiles
__global__ void sel_counts_kernel(SelDesc s)
{
    uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q < s.nq) s.counts[q] = (uint64_t)(s.tile_prefix[s.sel_off[q + 1] / SEL_TILE] - s.tile_prefix[s.sel_off[q] / SEL_TILE]);
}

// One WARP per entered tile, one LANE per flag word (32 slots): the lane's rank is the tile's prefix plus a warp scan
// over the words' popcounts, then every lane walks the set bits of its word.  (One thread per slot -- 1024-thread
// CTAs, every warp redoing the scan, 29 % of the lanes with anything to write -- cost 92 M warp instructions for 7 M
// matches, profiles/r02p.)  The tile's V_0 values are staged in shared memory with coalesced loads.
constexpr int EMIT_WARPS = 8;
__global__ void __launch_bounds__(EMIT_WARPS * 32) sel_emit_kernel(SelDesc s)
{
    __shared__ uint32_t s_x[EMIT_WARPS][SEL_TILE];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t tile = blockIdx.x * EMIT_WARPS + warp;
    if (tile >= s.lay[1]) return;
    // independent loads first, then what depends on q
    const uint32_t entered = s.entry[tile];
    const uint32_t slot0 = tile * SEL_TILE;
    const uint32_t q = sel_query_of_tile(s, slot0);
    uint32_t bits = s.flagbits[tile * 32 + lane];
    const uint32_t tile_base = s.tile_prefix[tile];
    if (entered == NONE32) return;  // warp-uniform
    const uint32_t off = s.sel_off[q], off_end = s.sel_off[q + 1];
    const uint32_t* v0 = s.v0[q];
    const uint32_t len = s.v0_len[q];
    const uint32_t tbase = slot0 - off;  // V_0 index of the tile's first slot
    uint32_t* sx = s_x[warp];
#pragma unroll 8
    for (uint32_t i = lane; i < (uint32_t)SEL_TILE; i += 32) sx[i] = tbase + i < len ? __ldg(v0 + tbase + i) : 0u;
    uint32_t inc = __popc(bits);
    const uint32_t mine = inc;
#pragma unroll
    for (int dd = 1; dd < 32; dd <<= 1) {
        const uint32_t up = __shfl_up_sync(0xffffffffu, inc, dd);
        if (lane >= (uint32_t)dd) inc += up;
    }
    uint32_t o = tile_base + inc - mine;  // rank of the lane's first selected element among the chunk's matches
    const uint32_t qfirst = s.tile_prefix[off / SEL_TILE];
    const uint32_t qend = s.tile_prefix[off_end / SEL_TILE];
    const uint32_t k = s.kq[q];
    const uint64_t g = s.global_offset;
    const uint32_t sh = s.sym_shift;
    const uint32_t* vl = s.vlast[q];
    __syncwarp();
    while (bits) {
        const uint32_t b = __ffs(bits) - 1;
        bits &= bits - 1;
        const uint32_t rel = lane * 32 + b, t = tbase + rel;
        const uint32_t x = sx[rel];
        if (s.starts) s.starts[o] = (g + x) >> sh;
        if (o == qfirst) s.first_start[q] = (g + x) >> sh;
        if (o + 1 == qend) {  // last match of the query: the cursor it leaves behind
            const uint32_t ll = s.last_len[q];
            uint64_t e = (uint64_t)(vl ? __ldg(vl + t) : x) + ll;
            if (!vl && ll == (1u << sh) && ((((uint64_t)x + g) >> sh) & 1ull) == 0) e = (uint64_t)x + (2u << sh);
            s.exit_cursor[q] = (g + e) >> sh;
        }
        if (s.tuples) {
            uint64_t* out = s.tuples + s.tuple_off[q] + (uint64_t)(o - qfirst) * k;
            out[0] = (g + x) >> sh;
            uint32_t idx = t;
            for (uint32_t d = k - 1, i = 1; d >= 1; d--, i++) {
                idx = __ldg(s.depth_vnxt[d] + s.depth_vseg[d][q] + idx);
                uint32_t pos = d > 1 ? __ldg(s.depth_vpos[d - 1] + s.depth_vseg[d - 1][q] + idx) : __ldg(s.last_list[q] + idx);
                out[i] = (g + pos) >> sh;
            }
        }
        o++;
    }
}

// ---- re-entry with new cursors (vlgb_result_fixup, multi-shard exactness) -----------------------------------
// The selection state of the last chunk stays on the device: succ[] / exitp[] for every viable start and one flag bit
// per start for the reported orbit.  A query that re-enters with another cursor starts its orbit at another element;
// orbits of the monotone successor coalesce, so only the stretch before the first common element changes.
//   refix  : one thread per re-entered query goes TILE BY TILE (exitp: one hop per tile of 1024 starts, where a walk
//            along succ[] takes one dependent load per MATCH) from the tile of the earlier of the two starts: the old
//            orbit's entry into a tile is the tile's first flagged start, the new one comes from the hops.  A tile both
//            orbits enter at the same element ends the walk (from there on they are the same orbit); a tile they enter
//            differently (or only one of them enters) is recorded as dirty with its new entry.
//   remark : one warp per tile of the chunk: dirty tiles are marked afresh from their new entry (sel_mark_tile), the
//            other tiles of re-entered queries are recounted from their flag words, all other tiles count 0, so that
//            the prefix scan + emit kernels produce exactly the re-entered queries' matches, back to back.
struct RefixDesc {
    uint32_t n;                  // queries that re-enter
    const uint32_t* q;           // [n] their index in the chunk
    const uint64_t* entry_local; // [n] new cursor relative to the shard start (bytes, saturated at 0)
    uint32_t* redo_flag;         // [nq] set for re-entered queries
    uint32_t* dirty;             // [tiles] NONE32 = as it was; REFIX_EMPTY = no orbit element any more; else the new entry
};
constexpr uint32_t REFIX_EMPTY = 0xFFFFFFFEu;

__global__ void sel_refix_kernel(SelDesc s, RefixDesc r)
{
    // one query per WARP (lane 0 works): 32 walks of different lengths in one warp would run at the pace of the longest
    const uint32_t i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (i >= r.n || (threadIdx.x & 31u)) return;
    const uint32_t q = r.q[i];
    r.redo_flag[q] = 1u;
    const uint32_t len = s.v0_len[q], off = s.sel_off[q];
    const uint32_t t_old = s.t0[q];
    const uint32_t t_new = full_lower_bound(s.v0[q], len, r.entry_local[i]);
    s.t0[q] = t_new;
    const uint32_t first_el = min(t_old, t_new);
    if (first_el >= len) return;  // neither orbit has an element
    const uint32_t tile0 = off / SEL_TILE, ntq = (len + SEL_TILE - 1) / SEL_TILE;
    uint32_t cur = t_new;
    for (uint32_t t = first_el / SEL_TILE; t < ntq; t++) {
        const uint32_t gt = tile0 + t;
        // independent loads: the tile's flag words and the hop out of the tile
        const bool here = cur < len && cur / SEL_TILE == t;
        const uint32_t nxt = here ? s.exitp[off + cur] : 0u;
        const uint4* fw = reinterpret_cast<const uint4*>(s.flagbits + (size_t)gt * 32);
        uint32_t old = NONE32;
#pragma unroll
        for (int w4 = 7; w4 >= 0; w4--) {
            const uint4 v = fw[w4];
            if (v.w) old = t * SEL_TILE + (w4 * 4 + 3) * 32 + (__ffs(v.w) - 1);
            if (v.z) old = t * SEL_TILE + (w4 * 4 + 2) * 32 + (__ffs(v.z) - 1);
            if (v.y) old = t * SEL_TILE + (w4 * 4 + 1) * 32 + (__ffs(v.y) - 1);
            if (v.x) old = t * SEL_TILE + (w4 * 4 + 0) * 32 + (__ffs(v.x) - 1);
        }
        const uint32_t now = here ? cur : NONE32;
        if (now == old) {
            if (now != NONE32) break;  // both orbits enter the tile at the same element: identical from here on
        } else {
            r.dirty[gt] = here ? cur : REFIX_EMPTY;
        }
        if (here) cur = nxt;
    }
}

__global__ void __launch_bounds__(MARK_WARPS * 32) sel_remark_kernel(SelDesc s, RefixDesc r, uint32_t ntiles)
{
    __shared__ uint32_t s_succ[MARK_WARPS][SEL_TILE];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t tile = blockIdx.x * MARK_WARPS + warp;
    if (tile >= ntiles) return;
    const uint32_t q = sel_query_of_tile(s, tile * SEL_TILE);
    const uint32_t d = r.dirty[tile];
    if (!r.redo_flag[q]) {  // warp-uniform
        if (lane == 0) {
            s.tile_count[tile] = 0;
            s.entry[tile] = NONE32;
        }
        return;
    }
    uint32_t c;
    if (d == NONE32) {
        c = __reduce_add_sync(0xffffffffu, __popc(s.flagbits[tile * 32 + lane]));
        if (lane == 0) s.tile_count[tile] = c;
    } else {
        sel_mark_tile(s, tile, d == REFIX_EMPTY ? NONE32 : d, s_succ[warp], lane);
        __syncwarp();
        c = __shfl_sync(0xffffffffu, lane == 0 ? s.tile_count[tile] : 0u, 0);
    }
    if (lane == 0) s.entry[tile] = c ? 0u : NONE32;  // emit skips tiles without a selected element
}

}  // namespace vlgb
