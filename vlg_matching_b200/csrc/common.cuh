// common.cuh — error handling, per-thread context (stream, arena, stage timers) for libvlgb200.
#pragma once
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <ctime>
#include <string>
#include <vector>

#include "../../include/vlgb200.h"

namespace vlgb {

// ---------------------------------------------------------------------------------------------
// error state (thread local; never throws across the ABI)
// ---------------------------------------------------------------------------------------------
inline std::string& last_error()
{
    static thread_local std::string e;
    return e;
}

inline int fail(int code, const char* fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    last_error() = buf;
    return code;
}

struct CudaError {
    cudaError_t e;
    const char* file;
    int line;
};

#define VLGB_CUDA(call)                                                        \
    do {                                                                       \
        cudaError_t e__ = (call);                                              \
        if (e__ != cudaSuccess) throw ::vlgb::CudaError{e__, __FILE__, __LINE__}; \
    } while (0)

// ---------------------------------------------------------------------------------------------
// stages (vlgb_stats)
// ---------------------------------------------------------------------------------------------
enum Stage : uint32_t {
    ST_SCAN_COUNT = 0,  // text scan, pass 1: per-(chunk, pattern) occurrence counts
    ST_SCAN_OFFSETS,    // count-matrix column scans
    ST_SCAN_FILL,       // text scan, pass 2: write sorted occurrence lists
    ST_JOIN_PROBE,      // gap join: lower_bound of x+lo in the right-hand viable list
    ST_JOIN_COMPACT,    // gap join: compaction of viable elements (+ prefix sums)
    ST_SELECT_SUCC,     // non-overlap successor of every viable start
    ST_SELECT_TILE,     // in-tile pointer doubling
    ST_SELECT_WALK,     // tile-to-tile walk + marking
    ST_EMIT,            // compaction of selected matches + tuple chain resolve
    ST_PREFIX,          // stand-alone device prefix sums
    ST_RANK_RANGE,      // rank path: SA range of every distinct subpattern (warp-cooperative search)
    ST_RANK_ENUM,       // rank path: in-order enumeration of the SA ranges through the wavelet tree
    ST_COUNT
};

inline const char* stage_name(uint32_t s)
{
    static const char* names[ST_COUNT] = {"scan_count", "scan_offsets", "scan_fill", "join_probe", "join_compact",
                                          "select_succ", "select_tile", "select_walk", "emit", "prefix", "rank_range", "rank_enum"};
    return s < ST_COUNT ? names[s] : "";
}

// ---------------------------------------------------------------------------------------------
// device arena: one growing allocation per purpose, reused across calls (no cudaMalloc in steady state)
// ---------------------------------------------------------------------------------------------
struct DeviceBuffer {
    void* p = nullptr;
    size_t cap = 0;
    void* get(size_t bytes)
    {
        if (bytes > cap) {
            if (p) cudaFree(p);
            p = nullptr;
            size_t want = bytes + bytes / 8 + 256;
            cudaError_t e = cudaMalloc(&p, want);
            if (e != cudaSuccess) {
                p = nullptr;
                cap = 0;
                throw CudaError{e, __FILE__, __LINE__};
            }
            cap = want;
        }
        return p;
    }
    void release()
    {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
};

// Bump allocator over one DeviceBuffer.
struct Arena {
    DeviceBuffer buf;
    size_t used = 0;
    size_t planned = 0;
    void reserve(size_t bytes)
    {
        buf.get(bytes);
        used = 0;
    }
    template <class T>
    T* take(size_t n)
    {
        size_t bytes = (n * sizeof(T) + 255) & ~size_t(255);
        if (used + bytes > buf.cap) throw CudaError{cudaErrorMemoryAllocation, __FILE__, __LINE__};
        T* r = reinterpret_cast<T*>(static_cast<char*>(buf.p) + used);
        used += bytes;
        return r;
    }
    static size_t padded(size_t n, size_t elem) { return (n * elem + 255) & ~size_t(255); }
};

struct PinnedBuffer {
    void* p = nullptr;
    size_t cap = 0;
    void* get(size_t bytes)
    {
        if (bytes > cap) {
            if (p) cudaFreeHost(p);
            p = nullptr;
            size_t want = bytes + bytes / 8 + 256;
            cudaError_t e = cudaMallocHost(&p, want);
            if (e != cudaSuccess) {
                p = nullptr;
                cap = 0;
                throw CudaError{e, __FILE__, __LINE__};
            }
            cap = want;
        }
        return p;
    }
};

// ---------------------------------------------------------------------------------------------
// per-thread context
// ---------------------------------------------------------------------------------------------
struct Context {
    int device = -1;
    bool ready = false;
    cudaStream_t stream = nullptr;
    int sm_count = 148;
    Arena scan_arena;    // count matrix, pattern tables
    Arena list_arena;    // occurrence lists of the current segment
    Arena join_arena;    // per-chunk join / selection state
    DeviceBuffer out_buf; // starts / tuples of the current chunk before they go to the host
    PinnedBuffer pinned; // small host staging: control blobs on their way up, sizes / counters on their way down
    size_t stage_used = 0;
    struct PendingD2H {
        void* host;
        const void* pinned;
        size_t bytes;
    };
    std::vector<PendingD2H> pending;
    static constexpr size_t STAGE_BYTES = 8u << 20;
    // bump allocation inside the pinned block; nullptr when the block is full (the caller then copies directly)
    void* stage_take(size_t bytes)
    {
        if (!pinned.p) pinned.get(STAGE_BYTES);
        const size_t at = (stage_used + 63) & ~size_t(63);
        if (at + bytes > pinned.cap) return nullptr;
        stage_used = at + bytes;
        return static_cast<char*>(pinned.p) + at;
    }
    // timers
    cudaEvent_t ev[2 * 512] = {};
    bool ev_ready = false;
    struct Span {
        uint32_t stage;
        int e0, e1;
    };
    std::vector<Span> spans;
    int ev_used = 0;
    vlgb_stats stats;
    uint64_t stream_budget = 0;  // 0 = default (a third of the free memory, <= 24 GiB); tiny = no hit stream
    uint64_t join_budget = 0;    // 0 = default
    uint64_t serial = 0;         // searches run on this thread (validity of the state kept for vlgb_result_fixup)
    int list_path = 0;           // VLGB_PATH_AUTO / _SCAN / _RANK (vlgb_set_path)
    cudaStream_t copy_stream = nullptr;  // result copies of VLGB_ASYNC_COPY searches (created on first use)
    cudaEvent_t emit_done = nullptr;     // orders a copy on copy_stream behind the emit kernel on `stream`
    bool copy_issued = false;            // the current search put something on copy_stream

    bool stage_timing = true;    // vlgb_set_stage_timing: CUDA events around every stage (vlgb_last_stats().ms)
    void begin(uint32_t stage)
    {
        if (!stage_timing) return;
        if (!ev_ready) {
            for (auto& e : ev) cudaEventCreate(&e);
            ev_ready = true;
        }
        if (ev_used + 2 > 2 * 512) flush_timers();
        cudaEventRecord(ev[ev_used], stream);
        spans.push_back({stage, ev_used, ev_used + 1});
        ev_used += 2;
    }
    void end(uint32_t stage, uint32_t launches, uint64_t bytes)
    {
        if (stage_timing) cudaEventRecord(ev[spans.back().e1], stream);
        stats.launches[stage] += launches;
        stats.bytes[stage] += bytes;
    }
    void flush_timers()
    {
        if (spans.empty()) return;
        cudaStreamSynchronize(stream);
        for (auto& s : spans) {
            float ms = 0;
            cudaEventElapsedTime(&ms, ev[s.e0], ev[s.e1]);
            stats.ms[s.stage] += ms;
        }
        spans.clear();
        ev_used = 0;
    }
};

inline Context& ctx()
{
    static thread_local Context c;
    return c;
}

inline void ensure_ready()
{
    Context& c = ctx();
    if (c.ready) return;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) throw CudaError{e == cudaSuccess ? cudaErrorNoDevice : e, __FILE__, __LINE__};
    if (c.device < 0) c.device = 0;
    VLGB_CUDA(cudaSetDevice(c.device));
    VLGB_CUDA(cudaStreamCreateWithFlags(&c.stream, cudaStreamNonBlocking));
    cudaDeviceProp prop;
    VLGB_CUDA(cudaGetDeviceProperties(&prop, c.device));
    c.sm_count = prop.multiProcessorCount;
    c.ready = true;
}

__host__ __device__ inline uint64_t sat_add64(uint64_t a, uint64_t b)
{
    return a > ~uint64_t(0) - b ? ~uint64_t(0) : a + b;
}

__host__ __device__ inline uint64_t sat_shl64(uint64_t a, uint32_t sh)
{
    return sh && (a >> (64 - sh)) ? ~uint64_t(0) : a << sh;
}

constexpr uint32_t NONE32 = 0xFFFFFFFFu;

// Small host<->device copies go through pinned memory: a copy from/to pageable memory is staged by the driver
// synchronously (tens of microseconds each, a dozen per search).
inline void h2d(Context& c, void* dst, const void* src, size_t bytes)
{
    if (!bytes) return;
    void* p = c.stage_take(bytes);
    if (p) {
        memcpy(p, src, bytes);
        src = p;
    }
    VLGB_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, c.stream));
}
// Small results (list sizes, counts, cursors: a few KB) are WRITTEN into the pinned staging block by a kernel (the block
// is mapped into the device's address space, unified addressing) instead of going through the copy engine: the
// device-to-host engine serves its requests in order, and with VLGB_ASYNC_COPY the 58 MB result copy of the previous
// search would hold up every one of these reads of the next search by up to a millisecond.
static __global__ void small_d2h_kernel(uint32_t* __restrict__ dst, const uint32_t* __restrict__ src, uint32_t words)
{
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < words; i += gridDim.x * blockDim.x) dst[i] = src[i];
}
inline void d2h(Context& c, void* host_dst, const void* dev_src, size_t bytes)  // lands in host_dst at the next sync_stream()
{
    if (!bytes) return;
    void* p = c.stage_take(bytes);
    if (p) {
        if (bytes <= (64u << 10) && bytes % 4 == 0 && reinterpret_cast<uintptr_t>(dev_src) % 4 == 0) {
            small_d2h_kernel<<<1, 256, 0, c.stream>>>(static_cast<uint32_t*>(p), static_cast<const uint32_t*>(dev_src), (uint32_t)(bytes / 4));
        } else {
            VLGB_CUDA(cudaMemcpyAsync(p, dev_src, bytes, cudaMemcpyDeviceToHost, c.stream));
        }
        c.pending.push_back({host_dst, p, bytes});
    } else {
        VLGB_CUDA(cudaMemcpyAsync(host_dst, dev_src, bytes, cudaMemcpyDeviceToHost, c.stream));
    }
}
// A result copy on the copy stream goes out in pieces: while ONE large device-to-host transfer is in flight, a kernel
// writing a few KB into pinned memory (small_d2h_kernel: the next search's counts) waits 60-80 us in the median and up
// to the whole transfer; behind pieces of 2 MB it waits ~20-30 us (idle link: 13 us; tools/microbench/d2h_contention.cu).
inline void d2h_in_pieces(void* host_dst, const void* dev_src, size_t bytes, cudaStream_t st, bool pieces)
{
    const size_t piece = pieces ? (size_t)2 << 20 : bytes;
    for (size_t o = 0; o < bytes; o += piece)
        VLGB_CUDA(cudaMemcpyAsync(static_cast<char*>(host_dst) + o, static_cast<const char*>(dev_src) + o, std::min(piece, bytes - o),
                                  cudaMemcpyDeviceToHost, st));
}
inline void sync_stream(Context& c)
{
    VLGB_CUDA(cudaStreamSynchronize(c.stream));
    c.stats.host_syncs++;
    for (auto& d : c.pending) memcpy(d.host, d.pinned, d.bytes);
    c.pending.clear();
}

// Host wall-clock trace of one search call (VLGB_TRACE=1 in the environment; stderr).  Measurement only.
struct HostTrace {
    bool on;
    double t0, last;
    static double now()
    {
        timespec ts;
        clock_gettime(CLOCK_MONOTONIC, &ts);
        return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6;
    }
    HostTrace()
    {
        static const bool env = getenv("VLGB_TRACE") != nullptr;
        on = env;
        t0 = last = on ? now() : 0;
    }
    void mark(const char* what)
    {
        if (!on) return;
        const double t = now();
        fprintf(stderr, "[vlgb trace] %-28s +%.3f ms (at %.3f)\n", what, t - last, t - t0);
        last = t;
    }
};
inline HostTrace& trace()
{
    static thread_local HostTrace t;
    return t;
}

}  // namespace vlgb
