// scan.cuh — multi-pattern text scan: text shard -> one sorted occurrence list per distinct
// subpattern (north-star (2), "coalesced text-scan path").  Replaces, for a whole batch at once, what
// the reference obtains per subpattern from forward_search (suffix_array_algorithm.hpp:48-115) plus the
// in-order WT enumeration of the SA range (wt_helper.hpp:726-785): the ascending list of text
// positions where the subpattern occurs.
//
// Layout / algorithm
//  * up to SCAN_S patterns per pass; patterns are keyed by (class, first min(len,8) bytes) where
//    class = min(len, 9).  Class <= 8 keys are exact; class 9 entries are verified byte by byte.
//  * shared memory per CTA: open-addressing hash table of the pass (SoA, SCAN_H slots), the prefilter
//    f3 (hashed first 3 bytes -> mask of the classes 3..9 that have a pattern with that prefix; f1/f2 for
//    1- and 2-byte patterns), one u16 cursor row and one work queue per warp.
//  * every warp owns one contiguous chunk of the shard ("row") and walks it 256 positions per step:
//    lane l holds bytes [8l, 8l+16) of the step in registers (one coalesced 8-byte load + a shuffle), tests
//    its 8 positions against the prefilter, and the hits are expanded into (position, class, window)
//    work items that are compacted IN POSITION ORDER into the warp's queue.  The queue is then drained
//    32 items at a time with all lanes busy: one hash probe per item.
//  * pass 1 counts occurrences per (row, pattern); the column-wise exclusive scan of that count matrix
//    gives every row its write cursor per pattern; pass 2 repeats the walk and writes positions at
//    cursor + rank (rank among equal patterns of the round by __match_any_sync), which makes every list
//    sorted and the result deterministic.
//  * algorithmic bytes per pass: 2 x shard bytes read + 4 B per occurrence written (+ count matrix).
#pragma once
#include <algorithm>
#include <string>
#include <unordered_map>

#include "common.cuh"

namespace vlgb {

constexpr int SCAN_S = 2048;       // patterns per pass
constexpr int SCAN_H = 4096;       // hash slots
constexpr int SCAN_WARPS = 16;     // warps per CTA
constexpr int SCAN_PPL = 8;        // positions per lane per step
constexpr int SCAN_STEP = 32 * SCAN_PPL;
constexpr int SCAN_F3 = 65536;     // 3-byte prefilter entries (u8: bit c-3 set = class c present)
constexpr int SCAN_F2_WORDS = 64;  // 2048 bits
constexpr int SCAN_F1_WORDS = 8;   // 256 bits
constexpr int SCAN_QCAP = 128;     // work items per warp queue
constexpr int SCAN_RB = 64;        // rows per row-block in the count-matrix scan
constexpr uint32_t SCAN_MAX_CHUNK = 65280;  // u16 per-row counters

struct ScanTables {  // one pass; same layout on host, in global memory and in shared memory
    uint32_t klo[SCAN_H];
    uint32_t khi[SCAN_H];
    uint32_t meta[SCAN_H];  // 0 = empty, else 0x80000000 | class << 16 | local pid
    uint8_t f3[SCAN_F3];
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
__host__ __device__ inline uint32_t scan_h3(uint32_t w) { return ((w & 0xFFFFFFu) * 0x9E3779B1u) >> 16; }
__host__ __device__ inline uint32_t scan_h2(uint32_t w) { return ((w & 0xFFFFu) * 0x9E3779B1u) >> 21; }

struct ScanArgs {
    const uint8_t* text;     // shard start (device, 16-byte aligned), >= 16 readable bytes after valid_len
    uint32_t valid_len;      // bytes of the shard that may be matched (owned + halo)
    uint32_t chunk;          // positions per row (multiple of SCAN_STEP, <= SCAN_MAX_CHUNK)
    uint32_t rows;
    uint32_t s_act;          // patterns in this pass
    uint32_t s_stride;       // row stride of the count matrix (multiple of 32)
    uint32_t has_short;      // the pass holds 1- or 2-byte patterns
    const ScanTables* tables;
    const uint32_t* long_len;  // [s_act] full length (class 9 only)
    const uint32_t* long_off;  // [s_act] offset into long_blob
    const uint8_t* long_blob;
    uint32_t* matrix;        // [rows][s_stride]: counts (pass 1 out) / write cursors inside each list (pass 2 in)
    const uint64_t* list_base;  // pass 2: [s_act] first word of every list inside `out`
    uint32_t* out;           // pass 2: positions (u32, shard-local)
};

constexpr size_t SCAN_SMEM = sizeof(ScanTables) + (size_t)SCAN_WARPS * (SCAN_S / 2) * sizeof(uint32_t) +
                             (size_t)SCAN_WARPS * SCAN_QCAP * sizeof(uint4);

template <bool FILL, bool SHORT>
__global__ void __launch_bounds__(SCAN_WARPS * 32, 1) scan_kernel(ScanArgs a)
{
    extern __shared__ __align__(16) uint8_t smem_raw[];
    ScanTables* T = reinterpret_cast<ScanTables*>(smem_raw);
    uint32_t* cursors = reinterpret_cast<uint32_t*>(smem_raw + sizeof(ScanTables));
    uint4* queues = reinterpret_cast<uint4*>(smem_raw + sizeof(ScanTables) + (size_t)SCAN_WARPS * (SCAN_S / 2) * 4);

    {  // stage the pass tables
        const uint4* src = reinterpret_cast<const uint4*>(a.tables);
        uint4* dst = reinterpret_cast<uint4*>(T);
        for (uint32_t i = threadIdx.x; i < sizeof(ScanTables) / 16; i += blockDim.x) dst[i] = __ldg(src + i);
    }
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t* cur = cursors + warp * (SCAN_S / 2);  // u16 counters, two per word
    uint4* q = queues + warp * SCAN_QCAP;
    for (uint32_t i = lane; i < SCAN_S / 2; i += 32) cur[i] = 0u;
    __syncthreads();

    const uint32_t row = blockIdx.x * SCAN_WARPS + warp;
    if (row >= a.rows) return;
    uint32_t* mrow = a.matrix + (size_t)row * a.s_stride;

    const uint32_t lo = row * a.chunk;
    const uint32_t hi = min(lo + a.chunk, a.valid_len);
    const uint32_t lt_mask = (1u << lane) - 1u;

    for (uint32_t base = lo; base < hi; base += SCAN_STEP) {
        // ---- 16 bytes per lane: own 8 + the next lane's 8
        const uint32_t off = base + 8 * lane;
        uint2 own = make_uint2(0u, 0u);
        if (off < a.valid_len) own = __ldg(reinterpret_cast<const uint2*>(a.text + off));
        uint2 nxt;
        nxt.x = __shfl_down_sync(0xffffffffu, own.x, 1);
        nxt.y = __shfl_down_sync(0xffffffffu, own.y, 1);
        if (lane == 31) {
            nxt = make_uint2(0u, 0u);
            if (base + SCAN_STEP < a.valid_len) nxt = __ldg(reinterpret_cast<const uint2*>(a.text + base + SCAN_STEP));
        }
        // ---- prefilter: class mask per position
        uint32_t wl[SCAN_PPL];  // low 32 bits of the window at position j
        wl[0] = own.x;
        wl[1] = __funnelshift_r(own.x, own.y, 8);
        wl[2] = __funnelshift_r(own.x, own.y, 16);
        wl[3] = __funnelshift_r(own.x, own.y, 24);
        wl[4] = own.y;
        wl[5] = __funnelshift_r(own.y, nxt.x, 8);
        wl[6] = __funnelshift_r(own.y, nxt.x, 16);
        wl[7] = __funnelshift_r(own.y, nxt.x, 24);
        uint32_t m[SCAN_PPL];
        uint32_t cnt = 0;
#pragma unroll
        for (int j = 0; j < SCAN_PPL; j++) {
            uint32_t mj = (uint32_t)T->f3[scan_h3(wl[j])] << 3;
            if (SHORT) {
                uint32_t b0 = wl[j] & 0xFFu;
                mj |= ((T->f1[b0 >> 5] >> (b0 & 31)) & 1u) << 1;
                uint32_t i2 = scan_h2(wl[j]);
                mj |= ((T->f2[i2 >> 5] >> (i2 & 31)) & 1u) << 2;
            }
            if (off + j >= hi) mj = 0;
            m[j] = mj;
            cnt += __popc(mj);
        }
        // ---- position-ordered offsets of the lanes' work items
        uint32_t incl = cnt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= (uint32_t)d) incl += o;
        }
        const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
        const uint32_t first = incl - cnt;

        for (uint32_t done = 0; done < total; done += SCAN_QCAP) {
            // ---- enqueue the items whose index falls into [done, done + QCAP)
            if (cnt && first < done + SCAN_QCAP && first + cnt > done) {
                uint32_t idx = first;
#pragma unroll
                for (int j = 0; j < SCAN_PPL; j++) {
                    uint32_t mm = m[j];
                    if (mm) {
                        uint32_t whi;
                        if (j < 4) whi = j == 0 ? own.y : __funnelshift_r(own.y, nxt.x, 8 * j);
                        else whi = j == 4 ? nxt.x : __funnelshift_r(nxt.x, nxt.y, 8 * (j - 4));
                        while (mm) {
                            uint32_t c = __ffs(mm) - 1;
                            mm &= mm - 1;
                            uint32_t rel = idx - done;
                            if (rel < (uint32_t)SCAN_QCAP) q[rel] = make_uint4(off + j, c, wl[j], whi);
                            idx++;
                        }
                    }
                }
            }
            __syncwarp();
            // ---- drain: one hash probe per item, 32 items per round
            const uint32_t nitems = min((uint32_t)SCAN_QCAP, total - done);
            for (uint32_t r0 = 0; r0 < nitems; r0 += 32) {
                const bool act = r0 + lane < nitems;
                uint4 it = make_uint4(0u, 0u, 0u, 0u);
                if (act) it = q[r0 + lane];
                const uint32_t p = it.x, c = it.y;
                uint32_t klo = it.z, khi = it.w;
                if (c < 4) { klo &= (1u << (8 * c)) - 1u; khi = 0; }  // c >= 1 for live items
                else if (c < 8) khi &= (1u << (8 * (c - 4))) - 1u;
                uint32_t slot = scan_hash(klo, khi, c);
                const uint32_t want = 0x80000000u | (c << 16);
                bool more = act;
                do {
                    uint32_t pid = NONE32;
                    if (more) {
                        for (;;) {
                            const uint32_t meta = T->meta[slot];
                            if (meta == 0) { more = false; break; }
                            const bool eq = ((meta & 0xFFFF0000u) == want) && T->klo[slot] == klo && T->khi[slot] == khi;
                            slot = (slot + 1) & (SCAN_H - 1);
                            if (eq) {
                                const uint32_t cand = meta & 0xFFFFu;
                                bool ok;
                                if (c < 9) {
                                    ok = (uint64_t)p + c <= a.valid_len;
                                    more = false;  // classes 1..8 have at most one entry per key
                                } else {
                                    const uint32_t L = a.long_len[cand];
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
                    // ---- warp-wide emit (all 32 lanes take part): rank among equal patterns = position order
                    const uint32_t grp = __match_any_sync(0xffffffffu, pid);
                    const int leader = __ffs(grp) - 1;
                    uint32_t old = 0;
                    if (pid != NONE32 && (int)lane == leader) {
                        const uint32_t sh = (pid & 1u) * 16;
                        old = (atomicAdd(&cur[pid >> 1], (uint32_t)__popc(grp) << sh) >> sh) & 0xFFFFu;
                    }
                    if (FILL) {
                        old = __shfl_sync(0xffffffffu, old, leader);
                        if (pid != NONE32)
                            a.out[__ldg(a.list_base + pid) + __ldg(mrow + pid) + old + __popc(grp & lt_mask)] = p;
                    }
                } while (__any_sync(0xffffffffu, more && c == 9));
            }
            __syncwarp();
        }
    }
    if (!FILL) {
        __syncwarp();
        for (uint32_t i = lane; i < a.s_act; i += 32) mrow[i] = (cur[i >> 1] >> ((i & 1u) * 16)) & 0xFFFFu;
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
    bool has_short = false;
};

inline void build_pass(const PatternDict& d, uint32_t first, uint32_t count, PassHost& ph)
{
    memset(&ph.tables, 0, sizeof ph.tables);
    ph.first = first;
    ph.count = count;
    ph.long_len.assign(count, 0);
    ph.long_off.assign(count, 0);
    ph.long_blob.clear();
    ph.has_short = false;
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
        if (c <= 2) ph.has_short = true;
        if (c == 1) {
            ph.tables.f1[(w32 & 0xFF) >> 5] |= 1u << (w32 & 31);
        } else if (c == 2) {
            uint32_t i2 = scan_h2(w32);
            ph.tables.f2[i2 >> 5] |= 1u << (i2 & 31);
        } else {
            ph.tables.f3[scan_h3(w32)] |= (uint8_t)(1u << (c - 3));
        }
        if (c == 9) {
            ph.long_len[i] = len;
            ph.long_off[i] = (uint32_t)ph.long_blob.size();
            ph.long_blob += s;
        }
    }
}

}  // namespace vlgb
