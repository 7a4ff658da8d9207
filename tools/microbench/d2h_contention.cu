// How long does a SMALL device-to-host read take while a 58 MB device-to-host copy is in flight on another stream?
// (Question behind VLGB_ASYNC_COPY: the search of batch i+1 needs a few KB of counts on the host while the result
// lists of batch i travel.)  Variants of the small read: cudaMemcpyAsync on a second stream; a kernel writing into
// pinned (mapped) host memory; the same with the big copy cut into pieces of 1 / 4 MB.
// build: nvcc -O2 -gencode arch=compute_100a,code=sm_100a -o d2h_contention d2h_contention.cu
#include <cuda_runtime.h>
#include <chrono>
#include <cstdio>
#include <vector>
#include <algorithm>
static __global__ void small_d2h_kernel(unsigned* dst, const unsigned* src, unsigned words)
{
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < words; i += gridDim.x * blockDim.x) dst[i] = src[i];
}
static double now() { return std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
int main()
{
    const size_t BIG = 58u << 20, SMALL = 8192;
    char *dbig, *hbig, *dsmall, *hsmall;
    cudaMalloc(&dbig, BIG); cudaMallocHost(&hbig, BIG); cudaMalloc(&dsmall, SMALL); cudaMallocHost(&hsmall, SMALL);
    cudaStream_t a, b; cudaStreamCreateWithFlags(&a, cudaStreamNonBlocking); cudaStreamCreateWithFlags(&b, cudaStreamNonBlocking);
    int lo, hi; cudaDeviceGetStreamPriorityRange(&lo, &hi);
    cudaStream_t ahi; cudaStreamCreateWithPriority(&ahi, cudaStreamNonBlocking, hi);
    auto big = [&](size_t piece) { for (size_t o = 0; o < BIG; o += piece) cudaMemcpyAsync(hbig + o, dbig + o, std::min(piece, BIG - o), cudaMemcpyDeviceToHost, b); };
    for (int variant = 0; variant < 4; variant++) {            // 0: memcpy on a, 1: kernel on a, 2: memcpy on high-priority stream, 3: kernel, 256 B only
        for (size_t piece : {BIG, (size_t)4 << 20, (size_t)1 << 20, (size_t)0}) {   // 0 = no big copy in flight
            std::vector<double> us;
            for (int it = 0; it < 12; it++) {
                cudaDeviceSynchronize();
                if (piece) big(piece);
                // 300 us into the big copy
                const double t0 = now(); while (now() - t0 < 300) {}
                const double t1 = now();
                cudaStream_t s = variant == 2 ? ahi : a;
                if (variant == 0 || variant == 2) cudaMemcpyAsync(hsmall, dsmall, SMALL, cudaMemcpyDeviceToHost, s);
                else small_d2h_kernel<<<1, 256, 0, s>>>((unsigned*)hsmall, (const unsigned*)dsmall, variant == 3 ? 64 : SMALL / 4);
                cudaStreamSynchronize(s);
                us.push_back(now() - t1);
            }
            std::sort(us.begin(), us.end());
            printf("variant %d (%s) big copy in pieces of %5zu KB: small read median %7.1f us  min %7.1f  max %7.1f\n", variant,
                   variant == 0 ? "cudaMemcpyAsync, 2nd stream" : variant == 1 ? "kernel -> pinned, 8 KB" : variant == 2 ? "cudaMemcpyAsync, high-priority stream" : "kernel -> pinned, 256 B",
                   piece >> 10, us[us.size() / 2], us.front(), us.back());
        }
    }
    double t0 = now(); big(BIG); cudaStreamSynchronize(b); printf("big copy alone: %.1f us\n", now() - t0);
    return 0;
}
