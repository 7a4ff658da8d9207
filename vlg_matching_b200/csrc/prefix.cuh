// prefix.cuh — device exclusive prefix sums over 0/1 flags (result compaction, north-star (4)).
//
// Three-phase scan (tile reduce -> scan of tile sums (recursive) -> tile downsweep).  Each thread owns
// 4 consecutive items read with one 16-byte (or 4-byte for u8 flags) load; a tile is 1024 threads x 4.
// Algorithmic bytes per item: flag read twice + 4 B rank written.
#pragma once
#include "common.cuh"

namespace vlgb {

constexpr int PFX_THREADS = 1024;
constexpr int PFX_ITEMS = 4;
constexpr int PFX_TILE = PFX_THREADS * PFX_ITEMS;

struct U4 {
    uint32_t v[4];
};

// flag sources -------------------------------------------------------------------------------------
struct FlagNotNone {  // u32 array, flag = (value != NONE32)
    const uint32_t* a;
    __device__ U4 operator()(uint64_t i4, uint64_t n) const
    {
        U4 r;
        uint64_t i = i4 * 4;
        if (i + 4 <= n) {
            uint4 q = *reinterpret_cast<const uint4*>(a + i);
            r.v[0] = q.x != NONE32; r.v[1] = q.y != NONE32; r.v[2] = q.z != NONE32; r.v[3] = q.w != NONE32;
        } else {
#pragma unroll
            for (int j = 0; j < 4; j++) r.v[j] = (i + j < n) ? (a[i + j] != NONE32) : 0u;
        }
        return r;
    }
};
struct FlagByte {  // u8 array of 0/1
    const uint8_t* a;
    __device__ U4 operator()(uint64_t i4, uint64_t n) const
    {
        U4 r;
        uint64_t i = i4 * 4;
        if (i + 4 <= n) {
            uchar4 q = *reinterpret_cast<const uchar4*>(a + i);
            r.v[0] = q.x; r.v[1] = q.y; r.v[2] = q.z; r.v[3] = q.w;
        } else {
#pragma unroll
            for (int j = 0; j < 4; j++) r.v[j] = (i + j < n) ? a[i + j] : 0u;
        }
        return r;
    }
};
struct ValueU32 {  // plain u32 values (used for the recursion over tile sums)
    const uint32_t* a;
    __device__ U4 operator()(uint64_t i4, uint64_t n) const
    {
        U4 r;
        uint64_t i = i4 * 4;
        if (i + 4 <= n) {
            uint4 q = *reinterpret_cast<const uint4*>(a + i);
            r.v[0] = q.x; r.v[1] = q.y; r.v[2] = q.z; r.v[3] = q.w;
        } else {
#pragma unroll
            for (int j = 0; j < 4; j++) r.v[j] = (i + j < n) ? a[i + j] : 0u;
        }
        return r;
    }
};

// block-wide exclusive scan of one value per thread; returns the exclusive prefix, *total = block sum
__device__ inline uint32_t block_exclusive_scan(uint32_t v, uint32_t* total)
{
    __shared__ uint32_t warp_sums[32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t o = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += o;
    }
    if (lane == 31) warp_sums[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = (lane < (blockDim.x >> 5)) ? warp_sums[lane] : 0u;
        uint32_t winc = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            uint32_t o = __shfl_up_sync(0xffffffffu, winc, d);
            if (lane >= d) winc += o;
        }
        warp_sums[lane] = winc - w;  // exclusive prefix of the warp sums
        if (lane == 31) *total = winc;
    }
    __syncthreads();
    uint32_t r = warp_sums[warp] + inc - v;
    __syncthreads();  // warp_sums may be reused by the next call
    return r;
}

template <class F>
__global__ void __launch_bounds__(PFX_THREADS) pfx_reduce_kernel(F in, uint64_t n, uint32_t* tile_sums)
{
    __shared__ uint32_t total;
    uint64_t i4 = (uint64_t)blockIdx.x * PFX_THREADS + threadIdx.x;
    U4 f = in(i4, n);
    uint32_t s = f.v[0] + f.v[1] + f.v[2] + f.v[3];
    block_exclusive_scan(s, &total);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}

// out has n+1 entries; out[n] = grand total (written by the last tile)
template <class F>
__global__ void __launch_bounds__(PFX_THREADS)
pfx_downsweep_kernel(F in, uint64_t n, const uint32_t* tile_prefix, uint32_t* out)
{
    __shared__ uint32_t total;
    uint64_t i4 = (uint64_t)blockIdx.x * PFX_THREADS + threadIdx.x;
    U4 f = in(i4, n);
    uint32_t s = f.v[0] + f.v[1] + f.v[2] + f.v[3];
    uint32_t base = block_exclusive_scan(s, &total) + (tile_prefix ? tile_prefix[blockIdx.x] : 0u);
    uint64_t i = i4 * 4;
    uint32_t r0 = base, r1 = r0 + f.v[0], r2 = r1 + f.v[1], r3 = r2 + f.v[2];
    if (i + 4 <= n) {
        *reinterpret_cast<uint4*>(out + i) = make_uint4(r0, r1, r2, r3);
    } else {
        if (i < n) out[i] = r0;
        if (i + 1 < n) out[i + 1] = r1;
        if (i + 2 < n) out[i + 2] = r2;
    }
    if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) out[n] = (tile_prefix ? tile_prefix[blockIdx.x] : 0u) + total;
}

// Scratch needed (in u32 words) for a scan over n items.
inline size_t pfx_scratch_words(uint64_t n)
{
    size_t words = 64;
    uint64_t m = (n + PFX_TILE - 1) / PFX_TILE;
    while (m > 1) {
        words += ((m + 63) & ~uint64_t(63)) + ((m + 1 + 63) & ~uint64_t(63));
        m = (m + PFX_TILE - 1) / PFX_TILE;
    }
    return words;
}

// Exclusive scan of flags `in` over n items into out[0..n] (out 16-byte aligned).  Returns launches.
template <class F>
inline uint32_t exclusive_scan(F in, uint64_t n, uint32_t* out, uint32_t* scratch, cudaStream_t st)
{
    if (n == 0) {
        cudaMemsetAsync(out, 0, sizeof(uint32_t), st);
        return 0;
    }
    uint64_t tiles = (n + PFX_TILE - 1) / PFX_TILE;
    uint32_t launches = 0;
    if (tiles == 1) {
        pfx_downsweep_kernel<F><<<1, PFX_THREADS, 0, st>>>(in, n, nullptr, out);
        return 1;
    }
    uint32_t* sums = scratch;                                   // tiles entries
    uint32_t* sums_scanned = scratch + ((tiles + 63) & ~uint64_t(63));  // tiles+1 entries
    uint32_t* rest = sums_scanned + ((tiles + 1 + 63) & ~uint64_t(63));
    pfx_reduce_kernel<F><<<(unsigned)tiles, PFX_THREADS, 0, st>>>(in, n, sums);
    launches++;
    launches += exclusive_scan(ValueU32{sums}, tiles, sums_scanned, rest, st);
    pfx_downsweep_kernel<F><<<(unsigned)tiles, PFX_THREADS, 0, st>>>(in, n, sums_scanned, out);
    launches++;
    return launches;
}

}  // namespace vlgb
