// sa_build.cu — index construction on the device: suffix array by prefix doubling, then the
// wavelet tree over it, level by level, straight into the flat rank blocks of wtsa.cuh.
//
// Replaces construct(vlg_index&) (include/sdsl/vlg_index.hpp:375-392): construct_sa -> qsufsort
// (construct_sa.hpp:145-171), wt_int ctor (wt_int.hpp:180-269: level-wise stable partition of the SA),
// bit_vector_il ctor (bit_vector_il.hpp:114-154: interleaving of the rank words).  The reference needs 84 s for a
// 32 MiB text on one host core; this is a one-time cost reported separately (index_build_ms).
//
// Suffix sort: rank_0 = the first 8 bytes of every suffix (big-endian, zero padded: the end of the text sorts first,
// like the reference's appended 0 symbol); round h sorts all suffixes by (rank_h[i], rank_h[i + h]) with an LSD radix
// sort over exactly the key bits in use and re-ranks; h doubles until every rank is unique.  The sort primitive is
// cub::DeviceRadixSort (CUDA toolkit library, like cuBLAS for a GEMM): construction is not the query hot path.
// Tree: the SA is a permutation of [0, n), so a node's place inside its level is closed-form (wtsa.cuh); level l is
// written from the level-l order of the values, then the values are stably partitioned by bit L-1-l INSIDE every node
// using the rank blocks just written (destination = node start + rank inside the node).
#include <cub/cub.cuh>

#include <chrono>

#define VLGB_WTSA_DEVICE_FUNCTIONS_ONLY
#include "wtsa.cuh"

namespace vlgb {

namespace {

struct Tmp {  // scratch that lives for one build
    std::vector<void*> ptrs;
    template <class T>
    T* get(size_t n)
    {
        void* p = nullptr;
        cudaError_t e = cudaMalloc(&p, n * sizeof(T) + 256);
        if (e != cudaSuccess) throw CudaError{e, __FILE__, __LINE__};
        ptrs.push_back(p);
        return static_cast<T*>(p);
    }
    ~Tmp()
    {
        for (void* p : ptrs) cudaFree(p);
    }
};

__global__ void sa_first_keys_kernel(const uint8_t* __restrict__ text, uint32_t n, uint64_t* key, uint32_t* val)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint64_t k = 0;
#pragma unroll
    for (uint32_t j = 0; j < 8; j++) k = (k << 8) | (i + j < n ? (uint64_t)__ldg(text + i + j) : 0ull);
    key[i] = k;
    val[i] = i;
}

// head[j] = j if sorted key j differs from its predecessor, else 0 (inclusive max-scan -> first index of the group)
__global__ void sa_heads_kernel(const uint64_t* __restrict__ key, uint32_t n, uint32_t* head, unsigned long long* groups)
{
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    bool is_head = false;
    if (j < n) {
        is_head = j == 0 || key[j] != key[j - 1];
        head[j] = is_head ? j : 0u;
    }
    const uint32_t cnt = __syncthreads_count(is_head);
    if (threadIdx.x == 0 && cnt) atomicAdd(groups, (unsigned long long)cnt);
}

__global__ void sa_scatter_rank_kernel(const uint32_t* __restrict__ sa, const uint32_t* __restrict__ grp, uint32_t n, uint32_t* rank)
{
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < n) rank[sa[j]] = grp[j];
}

__global__ void sa_pair_keys_kernel(const uint32_t* __restrict__ rank, uint32_t n, uint32_t h, uint32_t bits, uint64_t* key, uint32_t* val)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint64_t second = (uint64_t)i + h < n ? (uint64_t)rank[i + h] + 1ull : 0ull;  // a suffix that ends sorts first
    key[i] = ((uint64_t)rank[i] << bits) | second;
    val[i] = i;
}

struct MaxOp {
    __device__ __forceinline__ uint32_t operator()(uint32_t a, uint32_t b) const { return a > b ? a : b; }
};

// ---- wavelet tree levels ------------------------------------------------------------------------------------------
// One warp per rank block: 14 ballots of the level's bit over 448 consecutive values.
__global__ void __launch_bounds__(256) wt_pack_level_kernel(const uint32_t* __restrict__ cur, uint32_t n, uint32_t bit, RankBlock* level,
                                                            uint32_t nblocks, uint32_t* block_ones)
{
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t blk = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (blk >= nblocks) return;
    const uint64_t base = (uint64_t)blk * WT_BLOCK_BITS;
    uint32_t mine = 0;  // lane t < 14 keeps ballot t
#pragma unroll
    for (uint32_t t = 0; t < 14; t++) {
        const uint64_t i = base + t * 32 + lane;
        const uint32_t b = i < n ? (__ldg(cur + i) >> bit) & 1u : 0u;
        const uint32_t word = __ballot_sync(0xffffffffu, b);
        if (lane == t) mine = word;
    }
    const uint32_t ones = __reduce_add_sync(0xffffffffu, lane < 14 ? (uint32_t)__popc(mine) : 0u);
    const uint32_t hi = __shfl_down_sync(0xffffffffu, mine, 1);
    if (lane < 14 && !(lane & 1)) level[blk].w[lane >> 1] = (uint64_t)hi << 32 | mine;
    if (lane == 0) block_ones[blk] = ones;
}

__global__ void wt_headers_kernel(RankBlock* level, uint32_t nblocks, const uint32_t* __restrict__ ones_prefix)
{
    const uint32_t blk = blockIdx.x * blockDim.x + threadIdx.x;
    if (blk < nblocks) level[blk].ones_before = ones_prefix[blk];
}

// stable partition of every node by `bit`: the level-(l+1) order of the values
__global__ void __launch_bounds__(256) wt_partition_kernel(const uint32_t* __restrict__ cur, uint32_t n, uint32_t bit, const RankBlock* __restrict__ level,
                                                           uint32_t* __restrict__ nxt)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t v = cur[i];
    const uint32_t lo = bit >= 31 ? 0u : (v >> (bit + 1)) << (bit + 1);
    const uint32_t a = wt_rank1(level, i) - (lo >> 1);  // ones of the node before i
    nxt[(v >> bit) & 1u ? lo + (1u << bit) + a : i - a] = v;
}

double now_ms()
{
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

}  // namespace

void wtsa_free(WtSaOwned& w)
{
    if (w.d_sa) cudaFree(w.d_sa);
    if (w.d_blocks) cudaFree(w.d_blocks);
    w = WtSaOwned();
}

void wtsa_build(cudaStream_t st, const uint8_t* d_text, uint64_t n64, bool keep_sa, WtSaOwned& out)
{
    out = WtSaOwned();
    if (n64 < 2 || n64 > 0xFFFFF000ull) throw CudaError{cudaErrorInvalidValue, __FILE__, __LINE__};
    const uint32_t n = (uint32_t)n64;
    const double t_begin = now_ms();
    const uint32_t T = 256, G = (n + T - 1) / T;
    uint32_t bits = 1;
    while ((1ull << bits) <= n) bits++;  // ranks and rank+1 are <= n < 2^bits
    uint32_t levels = 1;
    while ((1ull << levels) < n) levels++;  // values are <= n-1 < 2^levels

    VLGB_CUDA(cudaMalloc(&out.d_sa, (size_t)n * 4 + 256));
    {
        Tmp tmp;
        uint64_t* key_a = tmp.get<uint64_t>(n);
        uint64_t* key_b = tmp.get<uint64_t>(n);
        uint32_t* val_a = tmp.get<uint32_t>(n);
        uint32_t* val_b = tmp.get<uint32_t>(n);
        uint32_t* rank = tmp.get<uint32_t>(n);
        uint32_t* head = tmp.get<uint32_t>(n);
        unsigned long long* d_groups = tmp.get<unsigned long long>(1);
        size_t sort_bytes = 0, scan_bytes = 0;
        {
            cub::DoubleBuffer<uint64_t> dk(key_a, key_b);
            cub::DoubleBuffer<uint32_t> dv(val_a, val_b);
            VLGB_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, dk, dv, (uint64_t)n, 0, 64, st));
            VLGB_CUDA(cub::DeviceScan::InclusiveScan(nullptr, scan_bytes, head, head, MaxOp(), (uint64_t)n, st));
        }
        void* cub_tmp = tmp.get<uint8_t>(std::max(sort_bytes, scan_bytes));
        size_t cub_bytes = std::max(sort_bytes, scan_bytes);

        sa_first_keys_kernel<<<G, T, 0, st>>>(d_text, n, key_a, val_a);
        uint32_t h = 8;
        int end_bit = 64;
        for (uint32_t round = 0;; round++) {
            cub::DoubleBuffer<uint64_t> dk(key_a, key_b);
            cub::DoubleBuffer<uint32_t> dv(val_a, val_b);
            size_t tb = cub_bytes;
            VLGB_CUDA(cub::DeviceRadixSort::SortPairs(cub_tmp, tb, dk, dv, (uint64_t)n, 0, end_bit, st));
            const uint64_t* sorted_key = dk.Current();
            const uint32_t* sorted_val = dv.Current();
            VLGB_CUDA(cudaMemsetAsync(d_groups, 0, 8, st));
            sa_heads_kernel<<<G, T, 0, st>>>(sorted_key, n, head, d_groups);
            unsigned long long groups = 0;
            VLGB_CUDA(cudaMemcpyAsync(&groups, d_groups, 8, cudaMemcpyDeviceToHost, st));
            VLGB_CUDA(cudaStreamSynchronize(st));
            out.doubling_rounds = round;
            if (groups == n) {  // every suffix has its own rank: sorted_val is the suffix array
                VLGB_CUDA(cudaMemcpyAsync(out.d_sa, sorted_val, (size_t)n * 4, cudaMemcpyDeviceToDevice, st));
                break;
            }
            tb = cub_bytes;
            VLGB_CUDA(cub::DeviceScan::InclusiveScan(cub_tmp, tb, head, head, MaxOp(), (uint64_t)n, st));
            sa_scatter_rank_kernel<<<G, T, 0, st>>>(sorted_val, head, n, rank);
            sa_pair_keys_kernel<<<G, T, 0, st>>>(rank, n, h, bits, key_a, val_a);
            end_bit = (int)(2 * bits);
            if (h >= 0x80000000u) throw CudaError{cudaErrorUnknown, __FILE__, __LINE__};  // cannot happen: ranks are unique once h >= n
            h *= 2;
        }
        VLGB_CUDA(cudaStreamSynchronize(st));
    }
    out.sa_ms = now_ms() - t_begin;

    // ---- the tree
    const uint32_t nblocks = n / WT_BLOCK_BITS + 1;  // rank1(n) is addressable
    VLGB_CUDA(cudaMalloc(&out.d_blocks, (size_t)levels * nblocks * sizeof(RankBlock)));
    {
        Tmp tmp;
        uint32_t* cur = tmp.get<uint32_t>(n);
        uint32_t* nxt = tmp.get<uint32_t>(n);
        uint32_t* ones = tmp.get<uint32_t>(nblocks + 1);
        uint32_t* ones_prefix = tmp.get<uint32_t>(nblocks + 1);
        size_t scan_bytes = 0;
        VLGB_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, ones, ones_prefix, (uint64_t)nblocks, st));
        void* cub_tmp = tmp.get<uint8_t>(scan_bytes);
        VLGB_CUDA(cudaMemcpyAsync(cur, out.d_sa, (size_t)n * 4, cudaMemcpyDeviceToDevice, st));
        const uint32_t GB = (uint32_t)(((uint64_t)nblocks * 32 + T - 1) / T);
        for (uint32_t l = 0; l < levels; l++) {
            const uint32_t bit = levels - 1 - l;
            RankBlock* level = out.d_blocks + (size_t)l * nblocks;
            wt_pack_level_kernel<<<GB, T, 0, st>>>(cur, n, bit, level, nblocks, ones);
            size_t tb = scan_bytes;
            VLGB_CUDA(cub::DeviceScan::ExclusiveSum(cub_tmp, tb, ones, ones_prefix, (uint64_t)nblocks, st));
            wt_headers_kernel<<<(nblocks + T - 1) / T, T, 0, st>>>(level, nblocks, ones_prefix);
            if (l + 1 < levels) {
                wt_partition_kernel<<<G, T, 0, st>>>(cur, n, bit, level, nxt);
                std::swap(cur, nxt);
            }
        }
        VLGB_CUDA(cudaGetLastError());
        VLGB_CUDA(cudaStreamSynchronize(st));
    }
    out.view.n = n;
    out.view.levels = levels;
    out.view.blocks_per_level = nblocks;
    out.view.blocks = out.d_blocks;
    if (!keep_sa) {
        cudaFree(out.d_sa);
        out.d_sa = nullptr;
    }
    out.view.sa = out.d_sa;
    out.build_ms = now_ms() - t_begin;
}

}  // namespace vlgb
