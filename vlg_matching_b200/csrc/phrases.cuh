// phrases.cuh — frequent-phrase extraction on the device.
//
// Replaces examples/generate-pattern.cpp:1-120, which builds a compressed suffix TREE (cst_sct3<csa_wt<wt_int>>) of the
// text and walks it for the `count` most frequent substrings of `length` symbols from a charset -- the patterns of the
// reference's experiment grid (benchmark/gapped-matching/Makefile:90-106).  Here the suffix array of the rank path
// (sa_build.cu) already lists the suffixes in lexicographic order, so the distinct `length`-grams are the runs of
// equal `length`-prefixes: one pass flags the run starts, a prefix sum + scatter compacts them, run lengths are the
// occurrence counts.  Only runs that can make the final list travel to the host (a power-of-two threshold taken from
// a histogram of the counts); the host then replays the reference's min-heap rule on them in lexicographic order, so
// ties are broken exactly as the reference breaks them.
#pragma once
#include "prefix.cuh"
#include "wtsa.cuh"

namespace vlgb {

// charset 0: isalnum or '_' (generate-pattern.cpp:41); 1: + ' ' (:42); 2: any symbol (human_readable = 0)
__device__ __forceinline__ bool phrase_char_ok(uint32_t c, uint32_t charset)
{
    if (charset >= 2) return true;
    const bool alnum = (c >= '0' && c <= '9') || (c >= 'A' && c <= 'Z') || (c >= 'a' && c <= 'z') || c == '_';
    return alnum || (charset == 1 && c == ' ');
}

__device__ __forceinline__ uint32_t phrase_sa(const WtSa& W, uint32_t i) { return W.sa ? __ldg(W.sa + i) : wt_access(W, i); }

// brk[i] = 1 when suffix SA[i] starts a new run: its `len`-prefix differs from SA[i-1]'s (a suffix shorter than `len`
// is a run of its own)
__global__ void __launch_bounds__(256) phrase_breaks_kernel(WtSa W, const uint8_t* __restrict__ text, uint32_t len, uint8_t* brk)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= W.n) return;
    bool b = true;
    if (i > 0) {
        const uint32_t t = phrase_sa(W, i), tp = phrase_sa(W, i - 1);
        if ((uint64_t)t + len <= W.n && (uint64_t)tp + len <= W.n) {
            b = false;
            for (uint32_t j = 0; j < len; j++)
                if (__ldg(text + t + j) != __ldg(text + tp + j)) { b = true; break; }
        }
    }
    brk[i] = b;
}

__global__ void __launch_bounds__(256) phrase_run_starts_kernel(const uint8_t* __restrict__ brk, const uint32_t* __restrict__ rank, uint32_t n,
                                                                 uint32_t* run_start)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && brk[i]) run_start[rank[i]] = i;
}

// occ[r] = occurrences of run r's gram if it is a full-length gram inside the charset, else 0; hist[b] counts the
// runs with 2^b <= occ < 2^(b+1)
__global__ void __launch_bounds__(256) phrase_run_counts_kernel(WtSa W, const uint8_t* __restrict__ text, uint32_t len, uint32_t charset,
                                                                 const uint32_t* __restrict__ run_start, uint32_t n_runs, uint32_t* occ,
                                                                 uint32_t* hist)
{
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_runs) return;
    const uint32_t i = run_start[r];
    const uint32_t end = r + 1 < n_runs ? run_start[r + 1] : W.n;
    const uint32_t t = phrase_sa(W, i);
    bool ok = (uint64_t)t + len <= W.n;
    for (uint32_t j = 0; ok && j < len; j++) ok = phrase_char_ok(__ldg(text + t + j), charset);
    const uint32_t c = ok ? end - i : 0u;
    occ[r] = c;
    if (c) atomicAdd(hist + (31 - __clz(c)), 1u);
}

__global__ void __launch_bounds__(256) phrase_select_flags_kernel(const uint32_t* __restrict__ occ, uint32_t n_runs, uint32_t min_occ, uint8_t* flag)
{
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r < n_runs) flag[r] = occ[r] >= min_occ;
}

__global__ void __launch_bounds__(256) phrase_gather_kernel(WtSa W, const uint8_t* __restrict__ text, uint32_t len, const uint8_t* __restrict__ flag,
                                                             const uint32_t* __restrict__ rank, const uint32_t* __restrict__ run_start,
                                                             const uint32_t* __restrict__ occ, uint32_t n_runs, uint8_t* grams, uint32_t* out_occ)
{
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_runs || !flag[r]) return;
    const uint32_t o = rank[r];
    const uint32_t t = phrase_sa(W, run_start[r]);
    for (uint32_t j = 0; j < len; j++) grams[(size_t)o * len + j] = __ldg(text + t + j);
    out_occ[o] = occ[r];
}

struct PhraseRuns {  // selected runs in lexicographic order
    std::vector<uint8_t> grams;  // [n][len]
    std::vector<uint32_t> occ;   // [n]
};

// want: how many phrases the caller will keep.  Throws CudaError.
inline void frequent_runs(Context& c, const WtSa& W, const uint8_t* d_text, uint32_t len, uint32_t charset, uint32_t want, PhraseRuns& out)
{
    const uint32_t n = W.n, T = 256;
    out.grams.clear();
    out.occ.clear();
    DeviceBuffer b_brk, b_rank, b_runs, b_occ, b_scratch, b_hist, b_grams, b_oocc;
    struct Free {
        DeviceBuffer* b[8];
        ~Free() { for (auto* x : b) x->release(); }
    } guard{{&b_brk, &b_rank, &b_runs, &b_occ, &b_scratch, &b_hist, &b_grams, &b_oocc}};
    uint8_t* brk = (uint8_t*)b_brk.get((size_t)n + 16);
    uint32_t* rank = (uint32_t*)b_rank.get(((size_t)n + 8) * 4);
    uint32_t* scratch = (uint32_t*)b_scratch.get(pfx_scratch_words(n) * 4);
    phrase_breaks_kernel<<<(n + T - 1) / T, T, 0, c.stream>>>(W, d_text, len, brk);
    exclusive_scan(FlagByte{brk}, n, rank, scratch, c.stream);
    uint32_t n_runs = 0;
    VLGB_CUDA(cudaMemcpyAsync(&n_runs, rank + n, 4, cudaMemcpyDeviceToHost, c.stream));
    VLGB_CUDA(cudaStreamSynchronize(c.stream));
    if (!n_runs) return;
    uint32_t* run_start = (uint32_t*)b_runs.get((size_t)n_runs * 4);
    uint32_t* occ = (uint32_t*)b_occ.get((size_t)n_runs * 4);
    uint32_t* hist = (uint32_t*)b_hist.get(32 * 4);
    VLGB_CUDA(cudaMemsetAsync(hist, 0, 32 * 4, c.stream));
    phrase_run_starts_kernel<<<(n + T - 1) / T, T, 0, c.stream>>>(brk, rank, n, run_start);
    phrase_run_counts_kernel<<<(n_runs + T - 1) / T, T, 0, c.stream>>>(W, d_text, len, charset, run_start, n_runs, occ, hist);
    uint32_t h[32];
    VLGB_CUDA(cudaMemcpyAsync(h, hist, sizeof h, cudaMemcpyDeviceToHost, c.stream));
    VLGB_CUDA(cudaStreamSynchronize(c.stream));
    // largest power of two that still leaves >= want runs: every run that can end in the heap has at least the
    // want-th largest count, which is >= this threshold
    uint32_t b = 31;
    uint64_t acc = 0;
    for (;; b--) {
        acc += h[b];
        if (acc >= want || b == 0) break;
    }
    const uint32_t min_occ = 1u << b;
    // brk / rank are reused for the selection flags and their prefix (n_runs <= n)
    phrase_select_flags_kernel<<<(n_runs + T - 1) / T, T, 0, c.stream>>>(occ, n_runs, min_occ, brk);
    exclusive_scan(FlagByte{brk}, n_runs, rank, scratch, c.stream);
    uint32_t n_sel = 0;
    VLGB_CUDA(cudaMemcpyAsync(&n_sel, rank + n_runs, 4, cudaMemcpyDeviceToHost, c.stream));
    VLGB_CUDA(cudaStreamSynchronize(c.stream));
    if (!n_sel) return;
    uint8_t* grams = (uint8_t*)b_grams.get((size_t)n_sel * len);
    uint32_t* oocc = (uint32_t*)b_oocc.get((size_t)n_sel * 4);
    phrase_gather_kernel<<<(n_runs + T - 1) / T, T, 0, c.stream>>>(W, d_text, len, brk, rank, run_start, occ, n_runs, grams, oocc);
    out.grams.resize((size_t)n_sel * len);
    out.occ.resize(n_sel);
    VLGB_CUDA(cudaMemcpyAsync(out.grams.data(), grams, out.grams.size(), cudaMemcpyDeviceToHost, c.stream));
    VLGB_CUDA(cudaMemcpyAsync(out.occ.data(), oocc, (size_t)n_sel * 4, cudaMemcpyDeviceToHost, c.stream));
    VLGB_CUDA(cudaGetLastError());
    VLGB_CUDA(cudaStreamSynchronize(c.stream));
}

}  // namespace vlgb
