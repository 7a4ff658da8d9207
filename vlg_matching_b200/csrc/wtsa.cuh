// wtsa.cuh — the index the reference actually stores, re-laid-out for B200: a wavelet tree over the suffix
// array in flat 64-byte rank blocks (north-star (1)), and the two query kernels on top of it (north-star (2)):
//
//   reference                                                        here
//   bit_vector_il<512> + rank_support_il::rank1                      RankBlock / wt_rank1: ONE 64-byte block per rank:
//     (bit_vector_il.hpp:114-154, 264-278): 72-byte blocks,            u64 ones-before + 448 payload bits, 64-byte
//     not line aligned, 1 cumulative + 8 payload words                 aligned, read with four 128-bit loads
//   wt_int::operator[]  (wt_int.hpp:339-361)                         wt_access: SA[i] through the levels, one block
//                                                                      per level (node offsets are closed-form)
//   wt_int::expand / value_range + wt_range_walker                   wt_quantile: the k-th smallest text position
//     (wt_int.hpp:836-852,913-939; wt_helper.hpp:726-785):             inside an SA range, one thread per output
//     DFS that enumerates an SA range in text order                    element -> the list comes out ASCENDING with no
//                                                                      sort, top levels shared through L1/L2
//   forward_search(text, sa) (suffix_array_algorithm.hpp:48-115):    wtsa_range_kernel: one warp per (subpattern,
//     two binary searches over the SA, probes compare text             bound), 32 probes per round (33-ary search),
//                                                                      SA[i] from the plain SA when kept, else wt_access
//
// The tree is over a PERMUTATION of [0, n) (the suffix array), which makes every node boundary closed-form:
// the node of value prefix `lo` at level l (bit b = L-1-l decides the child) occupies indices [min(lo,n), ...) of the
// level's array, exactly lo/2 of the values below `lo` have bit b set, and the right child starts at lo + 2^b.
// Hence no per-level header has to be read at all (the reference keeps m_zero-counts per level in rank reads,
// wt_int.hpp:836-852: 5 rank1 per expansion; here: 2 per level, 1 when the range fits one block).
#pragma once
#include "common.cuh"

namespace vlgb {

constexpr uint32_t WT_BLOCK_BITS = 448;  // 7 payload words of 64 bits + one u64 header = 64 bytes

struct __align__(64) RankBlock {
    uint64_t ones_before;  // ones of this level in all earlier blocks
    uint64_t w[7];
};
static_assert(sizeof(RankBlock) == 64, "one rank block is one 64-byte line half");

struct WtSa {                 // device-resident, read-only at query time
    uint32_t n = 0;           // suffixes (= text bytes of the shard)
    uint32_t levels = 0;      // L = bits of n-1
    uint32_t blocks_per_level = 0;
    const uint32_t* sa = nullptr;       // plain suffix array (nullptr when the index was built compact)
    const RankBlock* blocks = nullptr;  // [levels][blocks_per_level]
};

struct LoadedBlock {
    uint4 q[4];
    __device__ __forceinline__ uint64_t word(int j) const  // j = 0: header, 1..7: payload
    {
        const uint4& v = q[j >> 1];
        return (j & 1) ? ((uint64_t)v.w << 32 | v.z) : ((uint64_t)v.y << 32 | v.x);
    }
};

__device__ __forceinline__ LoadedBlock wt_load_block(const RankBlock* level, uint32_t blk)
{
    const uint4* p = reinterpret_cast<const uint4*>(level + blk);
    LoadedBlock b;
    b.q[0] = __ldg(p);
    b.q[1] = __ldg(p + 1);
    b.q[2] = __ldg(p + 2);
    b.q[3] = __ldg(p + 3);
    return b;
}

// ones in [0, blk*448 + off) of the level, from a loaded block (off in [0, 448])
__device__ __forceinline__ uint32_t wt_block_rank(const LoadedBlock& b, uint32_t off)
{
    uint32_t cnt = (uint32_t)b.word(0);
    const uint32_t wq = off >> 6, r = off & 63u;
#pragma unroll
    for (int j = 0; j < 7; j++) {
        const uint64_t w = b.word(j + 1);
        const uint64_t m = (uint32_t)j < wq ? ~0ull : ((uint32_t)j == wq ? ((1ull << r) - 1ull) : 0ull);
        cnt += __popcll(w & m);
    }
    return cnt;
}

__device__ __forceinline__ uint32_t wt_block_bit(const LoadedBlock& b, uint32_t off)
{
    uint32_t bit = 0;
    const uint32_t wq = off >> 6, r = off & 63u;
#pragma unroll
    for (int j = 0; j < 7; j++)
        if ((uint32_t)j == wq) bit = (uint32_t)(b.word(j + 1) >> r) & 1u;
    return bit;
}

__device__ __forceinline__ uint32_t wt_rank1(const RankBlock* level, uint32_t i)
{
    const uint32_t blk = i / WT_BLOCK_BITS;
    return wt_block_rank(wt_load_block(level, blk), i - blk * WT_BLOCK_BITS);
}

// SA[i] read through the tree (wt_int::operator[], wt_int.hpp:339-361): one rank block per level.
__device__ __forceinline__ uint32_t wt_access(const WtSa& W, uint32_t i)
{
    uint32_t lo = 0, idx = i;
    for (uint32_t l = 0; l < W.levels; l++) {
        const uint32_t b = W.levels - 1 - l;
        const RankBlock* level = W.blocks + (size_t)l * W.blocks_per_level;
        const uint32_t blk = idx / WT_BLOCK_BITS, off = idx - blk * WT_BLOCK_BITS;
        const LoadedBlock B = wt_load_block(level, blk);
        const uint32_t a = wt_block_rank(B, off) - (lo >> 1);  // ones of the node before idx
        if (wt_block_bit(B, off)) {
            idx = lo + (1u << b) + a;
            lo += 1u << b;
        } else {
            idx -= a;
        }
    }
    return lo;
}

// k-th smallest value (0-based) among SA[l .. r): the walk of wt_range_walker (wt_helper.hpp:726-785) for ONE output
// element.  Two rank blocks per level while the range spans several blocks, one as soon as it fits a block (always
// once it has shrunk to a single element, where this is wt_access).
__device__ __forceinline__ uint32_t wt_quantile(const WtSa& W, uint32_t l, uint32_t r, uint32_t k)
{
    uint32_t lo = 0;
    for (uint32_t lev = 0; lev < W.levels; lev++) {
        const uint32_t b = W.levels - 1 - lev;
        const RankBlock* level = W.blocks + (size_t)lev * W.blocks_per_level;
        const uint32_t bl = l / WT_BLOCK_BITS, br = r / WT_BLOCK_BITS;
        const LoadedBlock BL = wt_load_block(level, bl);
        const uint32_t half = lo >> 1;
        const uint32_t rl = wt_block_rank(BL, l - bl * WT_BLOCK_BITS) - half;
        uint32_t rr;
        if (br == bl) rr = wt_block_rank(BL, r - br * WT_BLOCK_BITS) - half;
        else rr = wt_block_rank(wt_load_block(level, br), r - br * WT_BLOCK_BITS) - half;
        const uint32_t nz = (r - l) - (rr - rl);  // zeros of the range: the elements that go left
        if (k < nz) {
            l -= rl;
            r -= rr;
        } else {
            k -= nz;
            l = lo + (1u << b) + rl;
            r = lo + (1u << b) + rr;
            lo += 1u << b;
        }
    }
    return lo;
}

#ifndef VLGB_WTSA_DEVICE_FUNCTIONS_ONLY  // the query kernels live in ONE translation unit (engine.cu)
// ---- SA range of every distinct subpattern (forward_search, suffix_array_algorithm.hpp:48-115) ------------------
struct WtsaRangeArgs {
    WtSa W;
    const uint8_t* text;       // the shard (n bytes)
    uint32_t n_pat;
    const uint8_t* pat_bytes;  // all subpatterns back to back
    const uint32_t* pat_off;   // [n_pat + 1]
    uint32_t* bound;           // [2 * n_pat]: lower bound, upper bound (first suffix whose m-prefix is >= / > the pattern)
};

// -1 / 0 / +1: the suffix at text position t compared with the pattern as the reference's comparison does -- only
// the first m symbols count, a suffix that ends early is smaller (suffix_array_algorithm.hpp:76-88).
__device__ __forceinline__ int wtsa_compare(const uint8_t* __restrict__ text, uint32_t n, uint32_t t,
                                            const uint8_t* __restrict__ pat, uint32_t m)
{
    for (uint32_t j = 0; j < m; j++) {
        if (t + j >= n) return -1;
        const uint32_t c = __ldg(text + t + j), p = __ldg(pat + j);
        if (c != p) return c < p ? -1 : 1;
    }
    return 0;
}

// One warp per (subpattern, bound): 32 probes per round cut the interval 33-fold -- six rounds for 2^28 suffixes
// instead of 28 dependent probes; the first rounds of every warp touch the same SA entries (L1/L2 hits).
__global__ void __launch_bounds__(128) wtsa_range_kernel(WtsaRangeArgs a)
{
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (w >= 2 * a.n_pat) return;
    const uint32_t p = w >> 1, upper = w & 1u;
    const uint8_t* pat = a.pat_bytes + a.pat_off[p];
    const uint32_t m = a.pat_off[p + 1] - a.pat_off[p];
    const uint32_t n = a.W.n;
    uint32_t lo = 0, hi = n;  // the answer lies in [lo, hi]
    while (hi > lo) {
        const uint32_t span = hi - lo;
        const uint32_t step = span > 32 ? (span + 32) / 33 : 1;
        const uint32_t idx = span > 32 ? lo + (lane + 1) * step - 1 : lo + lane;
        bool ge = true;  // probes beyond the interval count as "already past the bound"
        if (idx < hi) {
            const uint32_t t = a.W.sa ? __ldg(a.W.sa + idx) : wt_access(a.W, idx);
            const int c = wtsa_compare(a.text, n, t, pat, m);
            ge = upper ? c > 0 : c >= 0;
        }
        const uint32_t bal = __ballot_sync(0xffffffffu, ge);
        const uint32_t f = bal ? (uint32_t)__ffs(bal) - 1u : 32u;  // first probe at or past the bound
        if (span > 32) {
            const uint32_t nlo = lo + f * step;
            if (f < 32) hi = min(hi, lo + (f + 1) * step - 1);
            lo = nlo;
        } else {
            lo = min(hi, lo + f);
            hi = lo;
        }
    }
    if (lane == 0) a.bound[w] = lo;
}

// ---- in-order enumeration of the SA ranges (wt_range_walker, wt_helper.hpp:726-785) -----------------------------
struct WtsaEnumArgs {
    WtSa W;
    uint32_t n_pat;
    const uint32_t* sp;         // [n_pat] first SA index of every subpattern's range
    const uint64_t* out_first;  // [n_pat + 1] running total of the range sizes (element e belongs to the last p with
                                //   out_first[p] <= e)
    const uint64_t* list_base;  // [n_pat] first word of every list inside `out`
    uint32_t* out;              // occurrence lists (u32 shard-local positions, ascending per subpattern)
    uint64_t total;
};

__global__ void __launch_bounds__(256) wtsa_enum_kernel(WtsaEnumArgs a)
{
    const uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= a.total) return;
    uint32_t plo = 0, phi = a.n_pat - 1;  // largest p with out_first[p] <= e
    while (plo < phi) {
        const uint32_t mid = (plo + phi + 1) >> 1;
        if (__ldg(a.out_first + mid) <= e) plo = mid; else phi = mid - 1;
    }
    const uint32_t k = (uint32_t)(e - __ldg(a.out_first + plo));
    const uint32_t cnt = (uint32_t)(__ldg(a.out_first + plo + 1) - __ldg(a.out_first + plo));
    const uint32_t l = __ldg(a.sp + plo);
    a.out[__ldg(a.list_base + plo) + k] = wt_quantile(a.W, l, l + cnt, k);
}

#endif  // VLGB_WTSA_DEVICE_FUNCTIONS_ONLY

// host side of the construction (sa_build.cu: device prefix-doubling suffix sort + level-by-level tree build)
struct WtSaOwned {
    WtSa view;
    uint32_t* d_sa = nullptr;
    RankBlock* d_blocks = nullptr;
    double build_ms = 0, sa_ms = 0;
    uint32_t doubling_rounds = 0;
    size_t bytes() const { return (view.sa ? (size_t)view.n * 4 : 0) + (size_t)view.levels * view.blocks_per_level * sizeof(RankBlock); }
};
// Throws CudaError.  keep_sa = false frees the plain suffix array after the tree is built (compact index: SA probes
// then go through wt_access).
void wtsa_build(cudaStream_t stream, const uint8_t* d_text, uint64_t n, bool keep_sa, WtSaOwned& out);
void wtsa_free(WtSaOwned& w);

}  // namespace vlgb
