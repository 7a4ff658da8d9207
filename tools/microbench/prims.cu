// Micro-benchmarks of the primitives the scan/scatter kernels are built from (B200): cycles per warp-instruction per SM
// when every resident warp issues the primitive back to back.  nvcc -arch=sm_100a -O3 -o prims prims.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t rng(uint32_t& s) { s = s * 1664525u + 1013904223u; return s >> 8; }

template <int MODE>
__global__ void __launch_bounds__(1024) k(uint32_t* out, uint32_t* gcnt, int iters, uint32_t ndistinct)
{
    __shared__ uint32_t tab[8192];
    for (int i = threadIdx.x; i < 8192; i += blockDim.x) tab[i] = i * 2654435761u;
    __syncthreads();
    uint32_t s = threadIdx.x * 9781u + blockIdx.x * 7919u + 1u, acc = 0;
    const uint32_t lane = threadIdx.x & 31, lt = (1u << lane) - 1u;
    for (int it = 0; it < iters; it++) {
        uint32_t v = rng(s) % ndistinct;  // "pid"
        if (MODE == 0) {  // baseline: rng only
            acc += v;
        } else if (MODE == 1) {  // MATCH.ANY
            acc += __popc(__match_any_sync(0xffffffffu, v) & lt);
        } else if (MODE == 2) {  // ballot peers, 11 bits
            uint32_t peers = 0xffffffffu;
#pragma unroll
            for (int b = 0; b < 11; b++) {
                const uint32_t t = v & (1u << b);
                const uint32_t bal = __ballot_sync(0xffffffffu, t != 0);
                peers &= bal ^ (t ? 0u : 0xffffffffu);
            }
            acc += __popc(peers & lt);
        } else if (MODE == 3) {  // shared atomic add, spread addresses
            atomicAdd(&tab[v & 8191], 1u);
        } else if (MODE == 4) {  // global RED, spread addresses in a 32 MB region (row-private rows of 2048)
            atomicAdd(gcnt + (size_t)((blockIdx.x * 32 + (threadIdx.x >> 5)) % 4096) * 2048 + (v & 2047), 1u);
        } else if (MODE == 5) {  // LDS random word
            acc += tab[v & 8191];
        } else if (MODE == 6) {  // LDS random byte (64 KB table emulated in 32 KB)
            acc += reinterpret_cast<uint8_t*>(tab)[v & 32767];
        } else if (MODE == 7) {  // shared atomic add returning (rank)
            acc += atomicAdd(&tab[v & 8191], 1u);
        } else if (MODE == 8) {  // plain RMW by lane (no dedupe; cost reference)
            uint32_t c = tab[v & 8191];
            tab[v & 8191] = c + 1;
            acc += c;
        } else if (MODE == 9) {  // scattered global store of 4 bytes into a 200 MB region
            gcnt[(v * 2654435761u) % (50u << 20)] = v;
        } else if (MODE == 10) { // shfl
            acc += __shfl_sync(0xffffffffu, v, v & 31);
        }
    }
    if (acc == 0xdeadbeef) out[0] = acc + tab[threadIdx.x];
}

template <int MODE>
void run(const char* name, uint32_t* out, uint32_t* g, uint32_t nd, float base_ms, float* ret = nullptr)
{
    const int iters = 2000, blocks = 148, threads = 1024;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    k<MODE><<<blocks, threads>>>(out, g, 10, nd);
    cudaEventRecord(e0);
    k<MODE><<<blocks, threads>>>(out, g, iters, nd);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    const double warp_instr_per_sm = (double)iters * 32;  // 32 warps per SM
    const double cyc = (ms - base_ms) * 1e-3 * 1.965e9 / warp_instr_per_sm;
    printf("%-44s nd=%5u  %8.3f ms  -> %7.2f cycles per warp-op per SM (net of rng baseline)\n", name, nd, ms, cyc);
    if (ret) *ret = ms;
}

int main()
{
    uint32_t *out, *g;
    cudaMalloc(&out, 4096);
    cudaMalloc(&g, (size_t)256 << 20);
    cudaMemset(g, 0, (size_t)256 << 20);
    for (uint32_t nd : {2048u, 64u, 8u}) {
        float base = 0;
        run<0>("baseline rng", out, g, nd, 0, &base);
        run<1>("MATCH.ANY + popc", out, g, nd, base);
        run<2>("11 ballots peers + popc", out, g, nd, base);
        run<3>("ATOMS add (no return)", out, g, nd, base);
        run<7>("ATOMS add (return)", out, g, nd, base);
        run<4>("REDG add, row-private 8 KB rows", out, g, nd, base);
        run<5>("LDS.32 random", out, g, nd, base);
        run<6>("LDS.U8 random", out, g, nd, base);
        run<8>("LDS+STS plain RMW", out, g, nd, base);
        run<9>("STG.32 scattered over 200 MB", out, g, nd, base);
        run<10>("SHFL idx", out, g, nd, base);
    }
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
