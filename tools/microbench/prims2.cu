// Second round of primitive timings with a cheap baseline (xorshift, power-of-two masks) so that costs of a few cycles show.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t rng(uint32_t& s) { s ^= s << 13; s ^= s >> 17; s ^= s << 5; return s; }
// Zipf-ish value: half of the draws fall on 4 hot values
__device__ __forceinline__ uint32_t draw(uint32_t& s, uint32_t mask, int zipf)
{
    uint32_t x = rng(s);
    if (zipf && (x & 0x100000u)) return (x >> 8) & 3u;
    return (x >> 8) & mask;
}
template <int MODE>
__global__ void __launch_bounds__(1024) k(uint32_t* out, uint32_t* g, int iters, uint32_t mask, int zipf, uint32_t* bad)
{
    __shared__ uint32_t tab[8192];
    for (int i = threadIdx.x; i < 8192; i += blockDim.x) tab[i] = 0;
    __syncthreads();
    uint32_t s = threadIdx.x * 9781u + blockIdx.x * 7919u + 1u, acc = 0;
    const uint32_t lane = threadIdx.x & 31, lt = (1u << lane) - 1u, warp = threadIdx.x >> 5;
    uint32_t* mine = tab + warp * 256;  // warp-private 256 counters
    for (int it = 0; it < iters; it++) {
        uint32_t v = draw(s, mask, zipf);
        if (MODE == 0) acc += v;
        else if (MODE == 1) acc += __popc(__match_any_sync(0xffffffffu, v) & lt);
        else if (MODE == 2) { atomicAdd(&tab[v & 8191], 1u); }
        else if (MODE == 3) { acc += atomicAdd(&tab[v & 8191], 1u); }
        else if (MODE == 4) { acc += tab[v & 8191]; }
        else if (MODE == 5) { acc += __ballot_sync(0xffffffffu, v & 1); }
        else if (MODE == 6) { acc += __shfl_sync(0xffffffffu, v, v & 31); }
        else if (MODE == 7) {  // is the order of same-address ATOMS inside one warp instruction the lane order?
            const uint32_t a = v & 255;
            const uint32_t old = atomicAdd(&mine[a], 1u);
            const uint32_t peers = __match_any_sync(0xffffffffu, a);
            const uint32_t base = __shfl_sync(0xffffffffu, old, __ffs(peers) - 1);
            if (old != base + __popc(peers & lt)) atomicAdd(bad, 1u);
            __syncwarp();
        } else if (MODE == 8) {  // marker protocol: ATOMS ret + LDS + any
            const uint32_t a = v & 255;
            const uint32_t r = atomicAdd(&mine[a], 1u);
            __syncwarp();
            const uint32_t now = mine[a];
            bool dup = now - r > 1;
            if (__any_sync(0xffffffffu, dup)) {
                if (dup) atomicOr(&mine[a], 0x80000000u);
                __syncwarp();
                const bool last = !dup && (mine[a] & 0x80000000u);
                __syncwarp();
                if (dup) atomicAnd(&mine[a], 0x7fffffffu);
                dup |= last;
                const uint32_t m = __ballot_sync(0xffffffffu, dup);
                if (dup) {
                    const uint32_t grp = __match_any_sync(m, a);
                    acc += (now & 0x7fffffffu) - __popc(grp) + __popc(grp & lt);
                }
            }
            acc += r;
            __syncwarp();
        } else if (MODE == 9) {  // REDG into a warp-private 8 KB row
            atomicAdd(g + (size_t)((blockIdx.x * 32 + warp) % 4736) * 2048 + (v & 2047), 1u);
        }
    }
    if (acc == 0xdeadbeef) out[0] = acc + tab[threadIdx.x];
}
template <int MODE>
void run(const char* name, uint32_t* out, uint32_t* g, uint32_t mask, int zipf, float base_ms, uint32_t* bad, float* ret = nullptr)
{
    const int iters = 4000;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaMemset(bad, 0, 4);
    k<MODE><<<148, 1024>>>(out, g, 10, mask, zipf, bad);
    cudaEventRecord(e0);
    k<MODE><<<148, 1024>>>(out, g, iters, mask, zipf, bad);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    uint32_t hb = 0; cudaMemcpy(&hb, bad, 4, cudaMemcpyDeviceToHost);
    printf("%-40s mask=%4x zipf=%d %8.3f ms -> %6.2f cyc/warp-op/SM net  bad=%u\n", name, mask, zipf, ms, (ms - base_ms) * 1e-3 * 1.965e9 / (iters * 32.0), hb);
    if (ret) *ret = ms;
}
int main()
{
    uint32_t *out, *g, *bad;
    cudaMalloc(&out, 4096); cudaMalloc(&g, (size_t)64 << 20); cudaMalloc(&bad, 4);
    cudaMemset(g, 0, (size_t)64 << 20);
    for (int zipf = 0; zipf < 2; zipf++)
        for (uint32_t mask : {2047u, 63u}) {
            float base;
            run<0>("baseline", out, g, mask, zipf, 0, bad, &base);
            run<1>("MATCH.ANY", out, g, mask, zipf, base, bad);
            run<2>("ATOMS add noret", out, g, mask, zipf, base, bad);
            run<3>("ATOMS add ret", out, g, mask, zipf, base, bad);
            run<4>("LDS.32", out, g, mask, zipf, base, bad);
            run<5>("VOTE.BALLOT", out, g, mask, zipf, base, bad);
            run<6>("SHFL.IDX", out, g, mask, zipf, base, bad);
            run<7>("ATOMS order check (bad = out of lane order)", out, g, mask, zipf, base, bad);
            run<8>("marker protocol rank", out, g, mask, zipf, base, bad);
            run<9>("REDG warp-private row", out, g, mask, zipf, base, bad);
        }
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
