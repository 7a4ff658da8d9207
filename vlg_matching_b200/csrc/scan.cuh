// scan.cuh — multi-pattern text scan: text segment -> one sorted occurrence list per distinct
// subpattern (north-star (2), "coalesced text-scan path").  Replaces, for a whole batch at once, what
// the reference obtains per subpattern from forward_search (suffix_array_algorithm.hpp:48-115) plus the
// in-order WT enumeration of the SA range (wt_helper.hpp:726-785): the ascending list of text
// positions where the subpattern occurs.
//
// Layout / algorithm
//  * up to SCAN_S patterns per pass; patterns are keyed by (class, first min(len,8) bytes) where
//    class = min(len, 9).  Class <= 8 keys are exact; class 9 entries are verified byte by byte.
//  * shared memory per CTA: open-addressing hash table of the pass (SoA, SCAN_H slots), three
//    prefilters (first byte / hashed first 2 bytes / hashed first 3 bytes -> class mask) and one
//    private cursor row per warp.
//  * every warp owns one contiguous chunk of the segment ("row") and walks it 32 positions per step,
//    lane = position, so warp-order = text order.  Pass 1 counts occurrences per (row, pattern); the
//    column-wise exclusive scan of that count matrix gives every row its write cursor per pattern;
//    pass 2 repeats the walk and writes positions at cursor + rank (rank among equal patterns in the
//    step by __match_any_sync), which makes every list sorted and the result deterministic.
//  * algorithmic bytes per pass: 2 x segment bytes read + 4 B per occurrence written (+ count matrix).
#pragma once
#include <algorithm>
#include <string>
#include <unordered_map>

#include "common.cuh"

namespace vlgb {

constexpr int SCAN_S = 1024;       // patterns per pass
constexpr int SCAN_H = 2048;       // hash slots
constexpr int SCAN_WARPS = 8;      // warps per CTA
constexpr int SCAN_F3 = 8192;      // 3-byte prefilter entries (u16 class masks)
constexpr int SCAN_F2_WORDS = 64;  // 2048 bits
constexpr int SCAN_F1_WORDS = 8;   // 256 bits
constexpr int SCAN_RB = 64;        // rows per row-block in the count-matrix scan

struct ScanTables {  // one pass; same layout on host, in global memory and in shared memory
    uint32_t klo[SCAN_H];
    uint32_t khi[SCAN_H];
    uint32_t meta[SCAN_H];  // 0 = empty, else 0x80000000 | class << 16 | local pid
    uint16_t f3[SCAN_F3];   // bit c set: some pattern of class c (3..9) starts with these 3 bytes (hashed)
    uint32_t f2[SCAN_F2_WORDS];
    uint32_t f1[SCAN_F1_WORDS];
};

__host__ __device__ inline uint32_t scan_hash(uint32_t klo, uint32_t khi, uint32_t c)
{
    uint32_t h = (klo * 0x9E3779B1u) ^ ((khi + c) * 0x85EBCA77u);
    h ^= h >> 15;
    h *= 0x2C1B3C6Du;
    h ^= h >> 12;
    return h & (SCAN_H - 1);
}
__host__ __device__ inline uint32_t scan_h3(uint32_t w) { return ((w & 0xFFFFFFu) * 0x9E3779B1u) >> 19; }
__host__ __device__ inline uint32_t scan_h2(uint32_t w) { return ((w & 0xFFFFu) * 0x9E3779B1u) >> 21; }

struct ScanArgs {
    const uint8_t* text;     // segment start (device), >= 16 readable bytes after valid_len
    uint32_t valid_len;      // bytes of the segment that may be matched (owned + halo)
    uint32_t chunk;          // positions per row (multiple of 32)
    uint32_t rows;
    uint32_t s_act;          // patterns in this pass
    uint32_t s_stride;       // row stride of the count matrix (multiple of 32)
    const ScanTables* tables;
    const uint32_t* long_len;  // [s_act] full length (class 9 only)
    const uint32_t* long_off;  // [s_act] offset into long_blob
    const uint8_t* long_blob;
    uint32_t* matrix;        // [rows][s_stride]: counts (pass 1 out) / write cursors inside each list (pass 2 in)
    const uint64_t* list_base;  // pass 2: [s_act] first word of every list inside `out`
    uint32_t* out;           // pass 2: positions (u32, shard-local)
};

__device__ __forceinline__ uint64_t load_window(const uint8_t* p)
{
    uintptr_t a = reinterpret_cast<uintptr_t>(p);
    const uint64_t* q = reinterpret_cast<const uint64_t*>(a & ~uintptr_t(7));
    uint32_t sh = (uint32_t)(a & 7) * 8;
    uint64_t lo = __ldg(q), hi = __ldg(q + 1);
    return sh ? ((lo >> sh) | (hi << (64 - sh))) : lo;
}

template <bool FILL>
__global__ void __launch_bounds__(SCAN_WARPS * 32) scan_kernel(ScanArgs a)
{
    extern __shared__ __align__(16) uint8_t smem_raw[];
    ScanTables* T = reinterpret_cast<ScanTables*>(smem_raw);
    uint32_t* cursors = reinterpret_cast<uint32_t*>(smem_raw + sizeof(ScanTables));

    {  // stage the pass tables
        const uint4* src = reinterpret_cast<const uint4*>(a.tables);
        uint4* dst = reinterpret_cast<uint4*>(T);
        for (uint32_t i = threadIdx.x; i < sizeof(ScanTables) / 16; i += blockDim.x) dst[i] = __ldg(src + i);
    }
    __syncthreads();

    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t row = blockIdx.x * SCAN_WARPS + warp;
    if (row >= a.rows) return;
    uint32_t* cur = cursors + warp * SCAN_S;
    uint32_t* mrow = a.matrix + (size_t)row * a.s_stride;
    for (uint32_t i = lane; i < a.s_act; i += 32) cur[i] = FILL ? mrow[i] : 0u;
    __syncwarp();

    const uint32_t lo = row * a.chunk;
    const uint32_t hi = min(lo + a.chunk, a.valid_len);
    const uint32_t lt_mask = (1u << lane) - 1u;

    for (uint32_t base = lo; base < hi; base += 32) {
        const uint32_t p = base + lane;
        uint64_t w = 0;
        uint32_t m = 0;
        if (p < hi) {
            w = load_window(a.text + p);
            const uint32_t w32 = (uint32_t)w;
            uint32_t b0 = w32 & 0xFFu;
            m = ((T->f1[b0 >> 5] >> (b0 & 31)) & 1u) << 1;
            uint32_t i2 = scan_h2(w32);
            m |= ((T->f2[i2 >> 5] >> (i2 & 31)) & 1u) << 2;
            m |= T->f3[scan_h3(w32)];
        }
        uint32_t allm = __reduce_or_sync(0xffffffffu, m);
        while (allm) {
            const uint32_t c = __ffs(allm) - 1;
            allm &= allm - 1;
            const bool act = (m >> c) & 1u;
            uint32_t klo = 0, khi = 0, slot = 0;
            if (act) {
                uint64_t key = c >= 8 ? w : (w & ((1ull << (8 * c)) - 1ull));
                klo = (uint32_t)key;
                khi = (uint32_t)(key >> 32);
                slot = scan_hash(klo, khi, c);
            }
            const uint32_t want = 0x80000000u | (c << 16);
            bool more = act;
            // classes 1..8 have at most one entry per key; class 9 may have several (same first 8 bytes)
            do {
                uint32_t pid = NONE32;
                if (more) {
                    for (;;) {
                        uint32_t meta = T->meta[slot];
                        if (meta == 0) { more = false; break; }
                        bool eq = ((meta & 0xFFFF0000u) == want) && T->klo[slot] == klo && T->khi[slot] == khi;
                        slot = (slot + 1) & (SCAN_H - 1);
                        if (eq) {
                            uint32_t cand = meta & 0xFFFFu;
                            bool ok;
                            if (c < 9) {
                                ok = (uint64_t)p + c <= a.valid_len;
                                more = false;
                            } else {
                                uint32_t L = a.long_len[cand];
                                ok = (uint64_t)p + L <= a.valid_len;
                                if (ok) {
                                    const uint8_t* pat = a.long_blob + a.long_off[cand];
                                    const uint8_t* txt = a.text + p;
                                    for (uint32_t j = 8; j < L; j++)
                                        if (__ldg(txt + j) != __ldg(pat + j)) { ok = false; break; }
                                }
                            }
                            // one candidate per round, hit or miss: lanes holding the same key walk the
                            // same slot sequence, so each pattern is emitted in one round per step and
                            // its list stays sorted
                            if (ok) pid = cand;
                            break;
                        }
                    }
                }
                // ---- warp-wide emit: all 32 lanes take part
                if (!FILL) {
                    if (pid != NONE32) atomicAdd(&cur[pid], 1u);
                } else {
                    uint32_t grp = __match_any_sync(0xffffffffu, pid);
                    int leader = __ffs(grp) - 1;
                    uint32_t rank = __popc(grp & lt_mask);
                    uint32_t old = 0;
                    if (pid != NONE32 && (int)lane == leader) {
                        old = cur[pid];
                        cur[pid] = old + __popc(grp);
                    }
                    old = __shfl_sync(0xffffffffu, old, leader);
                    if (pid != NONE32) a.out[__ldg(a.list_base + pid) + old + rank] = p;
                    __syncwarp();
                }
            } while (c == 9 && __any_sync(0xffffffffu, more));
        }
    }
    if (!FILL) {
        __syncwarp();
        for (uint32_t i = lane; i < a.s_act; i += 32) mrow[i] = cur[i];
    }
}

// ---- count-matrix column scans -----------------------------------------------------------------
// part[rb][pid] = sum of matrix[r][pid] over the rows of row-block rb
__global__ void scan_colsum_kernel(const uint32_t* matrix, uint32_t rows, uint32_t stride, uint32_t* part)
{
    uint32_t pid = blockIdx.x * blockDim.x + threadIdx.x;
    if (pid >= stride) return;
    uint32_t r0 = blockIdx.y * SCAN_RB, r1 = min(rows, r0 + SCAN_RB);
    uint32_t s = 0;
    for (uint32_t r = r0; r < r1; r++) s += matrix[(size_t)r * stride + pid];
    part[(size_t)blockIdx.y * stride + pid] = s;
}

// part -> exclusive prefix over row-blocks (in place), total[pid] = list length
__global__ void scan_colbase_kernel(uint32_t* part, uint32_t nrb, uint32_t stride, uint32_t* total)
{
    uint32_t pid = blockIdx.x * blockDim.x + threadIdx.x;
    if (pid >= stride) return;
    uint32_t run = 0;
    for (uint32_t rb = 0; rb < nrb; rb++) {
        uint32_t v = part[(size_t)rb * stride + pid];
        part[(size_t)rb * stride + pid] = run;
        run += v;
    }
    total[pid] = run;
}

// matrix[r][pid] <- part[rb][pid] + (prefix of the counts inside the row-block) = write cursor of row r
// inside list pid
__global__ void scan_rowoff_kernel(uint32_t* matrix, uint32_t rows, uint32_t stride, const uint32_t* part)
{
    uint32_t pid = blockIdx.x * blockDim.x + threadIdx.x;
    if (pid >= stride) return;
    uint32_t r0 = blockIdx.y * SCAN_RB, r1 = min(rows, r0 + SCAN_RB);
    uint32_t run = part[(size_t)blockIdx.y * stride + pid];
    for (uint32_t r = r0; r < r1; r++) {
        uint32_t c = matrix[(size_t)r * stride + pid];
        matrix[(size_t)r * stride + pid] = run;
        run += c;
    }
}

// owned[pid] = number of positions < owned_len in list pid (lists are sorted)
__global__ void scan_owned_kernel(const uint32_t* pos, const uint64_t* base, const uint32_t* total, uint32_t s_act,
                                  uint32_t owned_len, uint32_t* owned)
{
    uint32_t pid = blockIdx.x * blockDim.x + threadIdx.x;
    if (pid >= s_act) return;
    const uint32_t* L = pos + base[pid];
    uint32_t lo = 0, hi = total[pid];
    while (lo < hi) {
        uint32_t mid = lo + ((hi - lo) >> 1);
        if (L[mid] < owned_len) lo = mid + 1; else hi = mid;
    }
    owned[pid] = lo;
}

// ---- host side: pattern dictionary and pass tables ----------------------------------------------
struct PatternDict {
    std::vector<std::string> pats;               // distinct subpatterns of the batch
    std::unordered_map<std::string, uint32_t> id;
    uint32_t add(const uint8_t* p, uint64_t len)
    {
        std::string s(reinterpret_cast<const char*>(p), (size_t)len);
        auto it = id.find(s);
        if (it != id.end()) return it->second;
        uint32_t k = (uint32_t)pats.size();
        pats.push_back(s);
        id.emplace(std::move(s), k);
        return k;
    }
};

struct PassHost {
    ScanTables tables;
    std::vector<uint32_t> long_len, long_off;
    std::string long_blob;
    uint32_t first = 0, count = 0;  // patterns [first, first+count) of the dictionary
};

inline void build_pass(const PatternDict& d, uint32_t first, uint32_t count, PassHost& ph)
{
    memset(&ph.tables, 0, sizeof ph.tables);
    ph.first = first;
    ph.count = count;
    ph.long_len.assign(count, 0);
    ph.long_off.assign(count, 0);
    ph.long_blob.clear();
    for (uint32_t i = 0; i < count; i++) {
        const std::string& s = d.pats[first + i];
        uint32_t len = (uint32_t)s.size();
        uint32_t c = len < 9 ? len : 9;
        uint64_t key = 0;
        for (uint32_t j = 0; j < len && j < 8; j++) key |= (uint64_t)(uint8_t)s[j] << (8 * j);
        uint32_t klo = (uint32_t)key, khi = (uint32_t)(key >> 32);
        uint32_t slot = scan_hash(klo, khi, c);
        while (ph.tables.meta[slot]) slot = (slot + 1) & (SCAN_H - 1);
        ph.tables.klo[slot] = klo;
        ph.tables.khi[slot] = khi;
        ph.tables.meta[slot] = 0x80000000u | (c << 16) | i;
        uint32_t w32 = klo;
        if (c == 1) {
            ph.tables.f1[(w32 & 0xFF) >> 5] |= 1u << (w32 & 31);
        } else if (c == 2) {
            uint32_t i2 = scan_h2(w32);
            ph.tables.f2[i2 >> 5] |= 1u << (i2 & 31);
        } else {
            ph.tables.f3[scan_h3(w32)] |= (uint16_t)(1u << c);
        }
        if (c == 9) {
            ph.long_len[i] = len;
            ph.long_off[i] = (uint32_t)ph.long_blob.size();
            ph.long_blob += s;
        }
    }
}

}  // namespace vlgb
