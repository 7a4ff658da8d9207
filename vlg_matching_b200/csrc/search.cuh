// search.cuh — host orchestration of one vlgb_search_batch call:
//   scan   (scan.cuh)    text shard -> one sorted occurrence list per distinct subpattern of the batch
//   join   (join.cuh)    right-to-left viability + next pointers, one launch set per depth
//   select (select.cuh)  greedy non-overlapping selection, prefix-sum compaction, tuple resolve
// Replaces the per-query body of locate()/count() (include/sdsl/vlg_index.hpp:395-411) and
// index_wcsearch3::search (benchmark/gapped-matching/include/index_wcsearch3.hpp:399-422) for a whole
// batch.  Host work is bookkeeping only (sizes, offsets, pointers); every position is produced on the
// device.  Two host<->device synchronisations per batch for the scan (list sizes) and one per query
// chunk (match counts) are the only places where sizes travel to the host.
#pragma once
#include <algorithm>
#include <memory>
#include <numeric>

#include "common.cuh"
#include "prefix.cuh"
#include "scan.cuh"
#include "wtsa.cuh"
#include "join.cuh"
#include "select.cuh"

namespace vlgb {

// host blob that becomes one H2D copy; offsets are 256-byte aligned
struct Packer {
    std::vector<uint8_t> host;
    template <class T>
    size_t put(const T* src, size_t n)
    {
        size_t off = (host.size() + 255) & ~size_t(255);
        host.resize(off + n * sizeof(T));
        if (n) memcpy(host.data() + off, src, n * sizeof(T));
        return off;
    }
    template <class T>
    size_t put(const std::vector<T>& v)
    {
        return put(v.data(), v.size());
    }
    size_t size() const { return (host.size() + 255) & ~size_t(255); }
};

template <class T>
inline T* dev_at(uint8_t* base, size_t off)
{
    return reinterpret_cast<T*>(base + off);
}

// ---------------------------------------------------------------------------------------------
// scan stage
// ---------------------------------------------------------------------------------------------
struct ScanOut {
    std::vector<uint32_t> tot;   // list length per distinct pattern
    std::vector<uint32_t> own;   // entries < owned_len
    std::vector<uint64_t> base;  // first word inside d_lists
    uint32_t* d_lists = nullptr;
    uint64_t total_words = 0;
};

// Rows (= warps' chunks): a whole number of waves of resident warps when the shard is large, at least
// one step per row when it is small, never more than the u16 row counters can hold.
inline uint32_t scan_chunk_for(uint64_t n, int sm_count)
{
    const uint64_t resident = (uint64_t)sm_count * SCAN_WARPS;
    uint64_t waves = (n + resident * SCAN_MAX_CHUNK - 1) / (resident * SCAN_MAX_CHUNK);
    if (waves < 1) waves = 1;
    uint64_t c = (n + resident * waves - 1) / (resident * waves);
    c = (c + SCAN_STEP - 1) / SCAN_STEP * SCAN_STEP;
    if (c < 1024) c = 1024;
    if (c > SCAN_MAX_CHUNK) c = SCAN_MAX_CHUNK;
    return (uint32_t)c;
}

inline void run_scan(Context& c, const uint8_t* d_text, uint64_t n, uint64_t owned_len, uint32_t sym_shift,
                     const PatternDict& dict, ScanOut& so)
{
    const uint32_t P = (uint32_t)dict.pats.size();
    so.tot.assign(P, 0);
    so.own.assign(P, 0);
    so.base.assign(P + 1, 0);
    so.d_lists = nullptr;
    so.total_words = 0;
    if (P == 0 || n == 0) return;

    const uint32_t chunk = scan_chunk_for(n, c.sm_count);
    const uint32_t rows = (uint32_t)((n + chunk - 1) / chunk);
    const uint32_t nrb = (rows + SCAN_RB - 1) / SCAN_RB;
    const uint32_t ngroups = (rows + SCAN_GROUP_ROWS - 1) / SCAN_GROUP_ROWS;
    const size_t smem = SCAN_SMEM;
    static thread_local bool attr_set = false;
    if (!attr_set) {
        VLGB_CUDA(cudaFuncSetAttribute(scan_kernel<SCAN_STREAM, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        VLGB_CUDA(cudaFuncSetAttribute(scan_kernel<SCAN_STREAM, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        VLGB_CUDA(cudaFuncSetAttribute(scan_kernel<SCAN_COUNT, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        VLGB_CUDA(cudaFuncSetAttribute(scan_kernel<SCAN_COUNT, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        VLGB_CUDA(cudaFuncSetAttribute(scan_kernel<SCAN_FILL, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        VLGB_CUDA(cudaFuncSetAttribute(scan_kernel<SCAN_FILL, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        VLGB_CUDA(cudaFuncSetAttribute(scan_scatter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SCAT_SMEM));
        attr_set = true;
    }

    struct PassDev {
        uint32_t first, s_act, stride, has_short, has_long;  // dictionary patterns [first, first + s_act)
        size_t o_tables, o_len, o_off, o_next, o_blob, o_base;  // inside the control blob
        uint32_t *gmatrix, *total, *halo;          // stream path: [groups][stride] counts -> group cursors; occurrences in the halo
        uint32_t *matrix = nullptr, *part = nullptr;  // legacy path: [rows][stride] counts -> row cursors, row-block sums
        uint32_t *stream = nullptr, *row_hits = nullptr, *overflow = nullptr;  // hit stream (if the budget allows one)
    };
    // passes: as many patterns each as fit the ids (SCAN_S) and the trie tables of one pass (build_pass decides).
    // The tables of the last dictionary are kept (per thread): a caller that searches the same pattern set again (the
    // same batch on the next shard of a sharded search, a repeated query file) skips the host build; the comparison
    // is on the pattern bytes themselves, not on a hash.
    struct TableCache {
        uint64_t stamp = 0;  // PatternDict::stamp it was built for (0 = none)
        std::string key;     // lengths + bytes of every pattern of that dictionary
        std::vector<PassDev> pd;
        std::vector<uint8_t> ctl;
        size_t o_allbase = 0;
        DeviceBuffer dev;    // the control blob on the device (tables, long-pattern arrays, list bases)
        int device = -1;
    };
    static thread_local TableCache cache;
    bool hit = dict.stamp && dict.stamp == cache.stamp && cache.device == c.device;
    std::string key;
    if (!hit) {
        key.reserve(P * 12);
        for (uint32_t i = 0; i < P; i++) {
            const uint32_t l = (uint32_t)dict.pats[i].size();
            key.append(reinterpret_cast<const char*>(&l), 4);
            key += dict.pats[i];
        }
        hit = key == cache.key && !cache.pd.empty() && cache.device == c.device;
    }
    std::vector<PassDev> pd;
    Packer ctl;
    size_t work_bytes = 0, o_allbase = 0;
    if (hit) {
        pd = cache.pd;
        o_allbase = cache.o_allbase;
        cache.stamp = dict.stamp;
    } else {
        std::unique_ptr<PassHost> ph(new PassHost());
        for (uint32_t first = 0; first < P;) {
            const uint32_t cnt = build_pass(dict, first, P - first, *ph);
            PassDev d;
            d.first = first;
            d.s_act = cnt;
            d.stride = (cnt + 31) & ~31u;
            d.has_short = ph->has_short;
            d.has_long = ph->has_long;
            d.o_tables = ctl.put(&ph->tables, 1);
            d.o_len = ctl.put(ph->long_len);
            d.o_off = ctl.put(ph->long_off);
            d.o_next = ctl.put(ph->long_next);
            ph->long_blob.append(16, '\0');
            d.o_blob = ctl.put(ph->long_blob.data(), ph->long_blob.size());
            std::vector<uint64_t> zero(cnt, 0);
            d.o_base = ctl.put(zero);
            pd.push_back(d);
            first += cnt;
        }
        o_allbase = ctl.put(so.base);  // P+1 u64, filled later (owned kernel)
        cache.stamp = dict.stamp;
        cache.key = key;
        cache.pd = pd;
        cache.ctl = ctl.host;
        cache.o_allbase = o_allbase;
        cache.device = -1;  // the device copy follows below
    }
    for (const PassDev& d : pd) work_bytes += Arena::padded((size_t)ngroups * d.stride, 4) + 2 * Arena::padded(d.stride, 4);
    const uint32_t npass = (uint32_t)pd.size();
    // hit streams: stream_cap entries per row (2 per text position, else 1), for as many passes as the budget
    // (a third of the free device memory, at most 24 GiB) allows; the other passes take the legacy two-scan path
    uint32_t stream_cap = 0, stream_passes = 0;
    {
        // free device memory is asked for once per thread and whenever the arena has to grow: cudaMemGetInfo costs
        // ~0.1 ms, a tenth of the scan itself
        static thread_local uint64_t cached_avail = 0;
        static thread_local size_t cached_cap = ~size_t(0);
        const size_t caps = c.scan_arena.buf.cap + c.list_arena.buf.cap + c.join_arena.buf.cap + c.out_buf.cap;
        if (cached_cap != caps) {  // some arena of this thread grew (or this is the first call)
            size_t free_b = 0, total_b = 0;
            cudaMemGetInfo(&free_b, &total_b);
            cached_avail = free_b + c.scan_arena.buf.cap;
            cached_cap = caps;
        }
        uint64_t budget = std::min<uint64_t>((uint64_t)24 << 30, cached_avail / 3);
        if (c.stream_budget) budget = std::min(budget, c.stream_budget);
        for (uint32_t per_pos = 2; per_pos >= 1 && !stream_passes; per_pos--) {
            const uint64_t per_pass = (uint64_t)rows * chunk * per_pos * 4;
            if (per_pass <= budget) {
                stream_cap = chunk * per_pos;
                stream_passes = (uint32_t)std::min<uint64_t>(npass, budget / per_pass);
            }
        }
        work_bytes += Arena::padded(npass, 4);
    }
    // legacy state: one per pass without a stream, plus ONE spare shared by the passes whose stream overflows
    // (they are redone one after the other)
    const size_t legacy_bytes = Arena::padded((size_t)rows * SCAN_S, 4) + Arena::padded((size_t)nrb * SCAN_S, 4);
    auto total_bytes = [&](uint32_t sp) {
        return work_bytes + sp * (Arena::padded((size_t)rows * stream_cap, 4) + Arena::padded(rows, 4)) +
               (size_t)(npass - sp + 1) * legacy_bytes + 4096;
    };
    trace().mark("scan: tables built");
    try {
        c.scan_arena.reserve(total_bytes(stream_passes));
    } catch (const CudaError& e) {
        // the free-memory figure above is cached and does not see what other threads / libraries took since: without
        // the streams the pass still runs (legacy two-scan path)
        if (e.e != cudaErrorMemoryAllocation || !stream_passes) throw;
        cudaGetLastError();
        stream_passes = 0;
        c.scan_arena.reserve(total_bytes(0));
    }
    // the control blob lives in the cache's own device buffer: a repeated dictionary is not uploaded again
    if (!hit) {
        uint8_t* d = (uint8_t*)cache.dev.get(ctl.size());
        h2d(c, d, ctl.host.data(), ctl.host.size());
        c.stats.h2d_bytes += ctl.host.size();
        cache.device = c.device;
    }
    uint8_t* d_ctl = (uint8_t*)cache.dev.p;
    for (uint32_t p = 0; p < npass; p++) {
        pd[p].gmatrix = c.scan_arena.take<uint32_t>((size_t)ngroups * pd[p].stride);
        pd[p].total = c.scan_arena.take<uint32_t>(pd[p].stride);
        pd[p].halo = c.scan_arena.take<uint32_t>(pd[p].stride);
        if (owned_len < n) VLGB_CUDA(cudaMemsetAsync(pd[p].halo, 0, (size_t)pd[p].stride * 4, c.stream));
    }
    uint32_t* d_overflow = c.scan_arena.take<uint32_t>(npass);
    VLGB_CUDA(cudaMemsetAsync(d_overflow, 0, (size_t)npass * 4, c.stream));
    for (uint32_t p = 0; p < stream_passes; p++) {
        pd[p].stream = c.scan_arena.take<uint32_t>((size_t)rows * stream_cap);
        pd[p].row_hits = c.scan_arena.take<uint32_t>(rows);
        pd[p].overflow = d_overflow + p;
    }
    for (uint32_t p = stream_passes; p < npass; p++) {
        pd[p].matrix = c.scan_arena.take<uint32_t>((size_t)rows * SCAN_S);
        pd[p].part = c.scan_arena.take<uint32_t>((size_t)nrb * SCAN_S);
    }
    uint32_t* spare_matrix = c.scan_arena.take<uint32_t>((size_t)rows * SCAN_S);
    uint32_t* spare_part = c.scan_arena.take<uint32_t>((size_t)nrb * SCAN_S);
    std::vector<uint32_t> overflow(npass, 0);

    auto args_for = [&](uint32_t p) {
        ScanArgs a;
        a.text = d_text;
        a.valid_len = (uint32_t)n;
        a.owned_len = (uint32_t)std::min<uint64_t>(owned_len, n);
        a.halo = pd[p].halo;
        a.chunk = chunk;
        a.rows = rows;
        a.s_act = pd[p].s_act;
        a.s_stride = pd[p].stride;
        a.has_short = pd[p].has_short;
        a.has_long = pd[p].has_long;
        static const uint32_t al[4] = {0xFFu, 0x55u, 0x11u, 0x01u};
        a.align = al[sym_shift];
        a.tables = dev_at<const ScanTables>(d_ctl, pd[p].o_tables);
        a.long_len = dev_at<const uint32_t>(d_ctl, pd[p].o_len);
        a.long_off = dev_at<const uint32_t>(d_ctl, pd[p].o_off);
        a.long_next = dev_at<const uint32_t>(d_ctl, pd[p].o_next);
        a.long_blob = dev_at<const uint8_t>(d_ctl, pd[p].o_blob);
        a.matrix = pd[p].matrix;
        a.gmatrix = pd[p].gmatrix;
        a.list_base = dev_at<const uint64_t>(d_ctl, pd[p].o_base);
        a.out = nullptr;
        a.stream = pd[p].stream;
        a.stream_cap = stream_cap;
        a.row_hits = pd[p].row_hits;
        a.overflow = pd[p].overflow;
        a.only_overflow = 0;
        return a;
    };
    const uint32_t grid = (rows + SCAN_WARPS - 1) / SCAN_WARPS;
    // legacy pass 1 of pass p into (matrix, part): per-row counts by global reductions, then per-row write cursors
    auto legacy_count = [&](uint32_t p, uint32_t* matrix, uint32_t* part, uint32_t* total) {
        ScanArgs a = args_for(p);
        a.matrix = matrix;
        c.begin(ST_SCAN_COUNT);
        VLGB_CUDA(cudaMemsetAsync(matrix, 0, (size_t)rows * pd[p].stride * 4, c.stream));
        if (a.has_short) scan_kernel<SCAN_COUNT, true><<<grid, SCAN_WARPS * 32, smem, c.stream>>>(a);
        else scan_kernel<SCAN_COUNT, false><<<grid, SCAN_WARPS * 32, smem, c.stream>>>(a);
        c.end(ST_SCAN_COUNT, 1, n);
        c.begin(ST_SCAN_OFFSETS);
        dim3 g((pd[p].stride + 127) / 128, nrb);
        scan_colsum_kernel<<<g, 128, 0, c.stream>>>(matrix, rows, pd[p].stride, part);
        scan_colbase_kernel<<<(pd[p].stride + 127) / 128, 128, 0, c.stream>>>(part, nrb, pd[p].stride, total);
        scan_rowoff_kernel<<<g, 128, 0, c.stream>>>(matrix, rows, pd[p].stride, part);
        c.end(ST_SCAN_OFFSETS, 3, (uint64_t)rows * pd[p].stride * 12);
    };

    // pass 1: occurrences counted (and recorded in the hit streams), then the write cursors
    for (uint32_t p = 0; p < npass; p++) {
        if (pd[p].stream) {
            ScanArgs a = args_for(p);
            c.begin(ST_SCAN_COUNT);
            if (a.has_short) scan_kernel<SCAN_STREAM, true><<<grid, SCAN_WARPS * 32, smem, c.stream>>>(a);
            else scan_kernel<SCAN_STREAM, false><<<grid, SCAN_WARPS * 32, smem, c.stream>>>(a);
            c.end(ST_SCAN_COUNT, 1, n);  // n read + 4 B per occurrence (added below once the counts are known)
            c.begin(ST_SCAN_OFFSETS);
            // group counts -> exclusive prefix over the groups (in place) = every group's cursor; totals = list lengths
            scan_groupbase_kernel<<<(pd[p].stride + 31) / 32, dim3(32, 32), 0, c.stream>>>(pd[p].gmatrix, ngroups, pd[p].stride, pd[p].total);
            c.end(ST_SCAN_OFFSETS, 1, (uint64_t)ngroups * pd[p].stride * 8);
        } else {
            legacy_count(p, pd[p].matrix, pd[p].part, pd[p].total);
        }
        d2h(c, so.tot.data() + pd[p].first, pd[p].total, pd[p].s_act * 4);
        c.stats.d2h_bytes += pd[p].s_act * 4;
        if (owned_len < n) {  // sharded text: the halo counts come down with the list sizes
            d2h(c, so.own.data() + pd[p].first, pd[p].halo, pd[p].s_act * 4);
            c.stats.d2h_bytes += pd[p].s_act * 4;
        }
    }
    d2h(c, overflow.data(), d_overflow, (size_t)npass * 4);
    VLGB_CUDA(cudaGetLastError());
    trace().mark("scan: pass 1 launched");
    sync_stream(c);  // sync 1: list sizes
    trace().mark("scan: sync 1");

    // owned prefix of every list = its length minus the occurrences in the halo (lists are ascending)
    if (owned_len < n) {
        for (uint32_t i = 0; i < P; i++) so.own[i] = so.tot[i] - so.own[i];
    } else {
        so.own = so.tot;
    }
    uint64_t words = 0;
    for (uint32_t i = 0; i < P; i++) {
        so.base[i] = words;
        words += ((uint64_t)so.tot[i] + 3) & ~3ull;  // 16-byte aligned list starts
    }
    so.base[P] = words;
    so.total_words = words;
    c.list_arena.reserve(words * 4 + 4096);
    so.d_lists = c.list_arena.take<uint32_t>(words + 16);
    for (uint32_t p = 0; p < npass; p++)
        h2d(c, d_ctl + pd[p].o_base, so.base.data() + pd[p].first, pd[p].s_act * 8);
    h2d(c, d_ctl + o_allbase, so.base.data(), (P + 1) * 8);
    c.stats.h2d_bytes += (uint64_t)P * 16 + 8;

    // pass 2: fill
    uint64_t occ = 0;
    for (uint32_t i = 0; i < P; i++) occ += so.tot[i];
    c.stats.total_occurrences += occ;
    for (uint32_t p = 0; p < npass; p++) {
        uint64_t pocc = 0;
        for (uint32_t i = 0; i < pd[p].s_act; i++) pocc += so.tot[pd[p].first + i];
        if (!pocc) continue;
        ScanArgs a = args_for(p);
        a.out = so.d_lists;
        if (pd[p].stream && !overflow[p]) {  // scatter the recorded occurrences
            c.begin(ST_SCAN_FILL);
            scan_scatter_kernel<<<ngroups, SCAT_WARPS * 32, scat_smem_bytes(a.s_act), c.stream>>>(a);
            c.end(ST_SCAN_FILL, 1, pocc * 8);  // 4 B per occurrence read from the stream + 4 B written to its list
            c.stats.bytes[ST_SCAN_COUNT] += pocc * 4;
            continue;
        }
        // legacy two-scan path: no stream memory for this pass, or some row found more occurrences than its stream
        // holds (> 2 per text position) -- count per row, then scan again and write at the row cursors
        if (pd[p].stream) {
            legacy_count(p, spare_matrix, spare_part, pd[p].total);
            a.matrix = spare_matrix;
        }
        c.begin(ST_SCAN_FILL);
        if (a.has_short) scan_kernel<SCAN_FILL, true><<<grid, SCAN_WARPS * 32, smem, c.stream>>>(a);
        else scan_kernel<SCAN_FILL, false><<<grid, SCAN_WARPS * 32, smem, c.stream>>>(a);
        c.end(ST_SCAN_FILL, 1, n + pocc * 4);
    }
    trace().mark("scan: pass 2 launched");
    c.stats.scan_passes += npass;
    c.stats.distinct_patterns += P;

}

// ---------------------------------------------------------------------------------------------
// rank path: occurrence lists from SA ranges (wtsa.cuh) instead of a pass over the text
// ---------------------------------------------------------------------------------------------
// Cost model (measured on B200): a scan pass costs a fixed launch/offset overhead plus the text at the
// scan's rate; the rank path costs ~0.25 us of device time per distinct subpattern for the range search (latency
// bound, hidden by parallel warps) and ~L dependent rank blocks per occurrence for the enumeration.
inline double scan_cost_ms(uint64_t n, uint32_t patterns)
{
    const uint32_t passes = (patterns + SCAN_S - 1) / SCAN_S;
    return passes * (0.12 + (double)n * 6.0e-9);
}
inline double rank_range_cost_ms(uint32_t patterns) { return 0.05 + patterns * 2.5e-4; }
inline double rank_enum_cost_ms(uint64_t occurrences, uint32_t levels) { return 0.02 + (double)occurrences * levels * 1.0e-7; }

// Returns false (and leaves `so` untouched apart from sizes) when the cost model prefers the scan.
inline bool run_rank_lists(Context& c, const uint8_t* d_text, uint64_t n, uint64_t owned_len, const WtSa& W,
                           const PatternDict& dict, ScanOut& so, bool forced)
{
    const uint32_t P = (uint32_t)dict.pats.size();
    so.tot.assign(P, 0);
    so.own.assign(P, 0);
    so.base.assign(P + 1, 0);
    so.d_lists = nullptr;
    so.total_words = 0;
    if (P == 0 || n == 0) return true;
    const double scan_ms = scan_cost_ms(n, P);
    if (!forced && rank_range_cost_ms(P) > scan_ms) return false;

    // ---- SA range of every distinct subpattern
    std::vector<uint32_t> pat_off(P + 1, 0);
    std::string blob;
    for (uint32_t i = 0; i < P; i++) {
        blob += dict.pats[i];
        pat_off[i + 1] = (uint32_t)blob.size();
    }
    Packer ctl;
    const size_t o_blob = ctl.put(blob.data(), blob.size()), o_off = ctl.put(pat_off);
    const size_t ctl_bytes = ctl.size();
    const size_t work = ctl_bytes + Arena::padded(2 * (size_t)P, 4) + 2 * Arena::padded(P + 1, 8) + 3 * Arena::padded(P, 4) + 4096;
    c.scan_arena.reserve(work);
    uint8_t* d_ctl = c.scan_arena.take<uint8_t>(ctl_bytes);
    uint32_t* d_bound = c.scan_arena.take<uint32_t>(2 * (size_t)P);
    uint64_t* d_first = c.scan_arena.take<uint64_t>(P + 1);
    uint64_t* d_base = c.scan_arena.take<uint64_t>(P + 1);
    uint32_t* d_sp = c.scan_arena.take<uint32_t>(P);
    uint32_t* d_tot = c.scan_arena.take<uint32_t>(P);
    uint32_t* d_owned = c.scan_arena.take<uint32_t>(P);
    h2d(c, d_ctl, ctl.host.data(), ctl.host.size());
    c.stats.h2d_bytes += ctl.host.size();
    WtsaRangeArgs ra;
    ra.W = W;
    ra.text = d_text;
    ra.n_pat = P;
    ra.pat_bytes = dev_at<const uint8_t>(d_ctl, o_blob);
    ra.pat_off = dev_at<const uint32_t>(d_ctl, o_off);
    ra.bound = d_bound;
    uint32_t rounds = 1;
    for (uint64_t span = n; span > 32; span /= 33) rounds++;
    c.begin(ST_RANK_RANGE);
    wtsa_range_kernel<<<(2 * P * 32 + 127) / 128, 128, 0, c.stream>>>(ra);
    // per probe: one SA entry (or `levels` rank blocks) + one 32-byte sector of text
    c.end(ST_RANK_RANGE, 1, (uint64_t)2 * P * rounds * 32 * (32 + (W.sa ? 4 : (uint64_t)W.levels * 64)));
    std::vector<uint32_t> bound(2 * (size_t)P);
    d2h(c, bound.data(), d_bound, bound.size() * 4);
    VLGB_CUDA(cudaGetLastError());
    sync_stream(c);  // sync: list sizes (the scan path pays the same one)
    c.stats.d2h_bytes += bound.size() * 4;
    uint64_t occ = 0;
    std::vector<uint32_t> sp(P);
    std::vector<uint64_t> first(P + 1, 0);
    for (uint32_t i = 0; i < P; i++) {
        sp[i] = bound[2 * i];
        so.tot[i] = bound[2 * i + 1] - bound[2 * i];
        first[i] = occ;
        occ += so.tot[i];
    }
    first[P] = occ;
    if (!forced && rank_enum_cost_ms(occ, W.levels) > scan_ms) return false;

    // ---- in-order enumeration: the lists come out ascending
    uint64_t words = 0;
    for (uint32_t i = 0; i < P; i++) {
        so.base[i] = words;
        words += ((uint64_t)so.tot[i] + 3) & ~3ull;
    }
    so.base[P] = words;
    so.total_words = words;
    c.list_arena.reserve(words * 4 + 4096);
    so.d_lists = c.list_arena.take<uint32_t>(words + 16);
    h2d(c, d_sp, sp.data(), (size_t)P * 4);
    h2d(c, d_first, first.data(), (size_t)(P + 1) * 8);
    h2d(c, d_base, so.base.data(), (size_t)(P + 1) * 8);
    c.stats.h2d_bytes += (uint64_t)P * 20 + 16;
    if (occ) {
        WtsaEnumArgs ea;
        ea.W = W;
        ea.n_pat = P;
        ea.sp = d_sp;
        ea.out_first = d_first;
        ea.list_base = d_base;
        ea.out = so.d_lists;
        ea.total = occ;
        c.begin(ST_RANK_ENUM);
        wtsa_enum_kernel<<<(unsigned)((occ + 255) / 256), 256, 0, c.stream>>>(ea);
        c.end(ST_RANK_ENUM, 1, occ * ((uint64_t)W.levels * 64 + 4));  // >= one rank block per level and element + the position
    }
    c.stats.total_occurrences += occ;
    c.stats.distinct_patterns += P;
    if (owned_len >= n) {
        so.own = so.tot;
    } else {
        h2d(c, d_tot, so.tot.data(), (size_t)P * 4);
        scan_owned_kernel<<<(P + 127) / 128, 128, 0, c.stream>>>(so.d_lists, d_base, d_tot, P, (uint32_t)owned_len, d_owned);
        d2h(c, so.own.data(), d_owned, (size_t)P * 4);
        VLGB_CUDA(cudaGetLastError());
        sync_stream(c);
        c.stats.h2d_bytes += (uint64_t)P * 4;
        c.stats.d2h_bytes += (uint64_t)P * 4;
    }
    VLGB_CUDA(cudaGetLastError());
    return true;
}

// ---------------------------------------------------------------------------------------------
// join + select stage for one chunk of queries
// ---------------------------------------------------------------------------------------------
struct PlanQuery {
    uint32_t q;        // index in the batch
    uint32_t k;
    uint32_t s0;       // first subpattern (batch numbering)
    uint64_t entry;    // whole-text entry cursor
};

struct ChunkOut {  // per query of the chunk, chunk order
    std::vector<uint64_t> counts, exit_cursor, first_start;
    uint64_t matches = 0, tuple_words = 0;
    uint64_t* h_starts = nullptr;  // pinned (inside block)
    uint64_t* h_tuples = nullptr;
    uint64_t* d_starts = nullptr;  // device copies (owned by the result when VLGB_KEEP_ON_DEVICE, else scratch)
    uint64_t* d_tuples = nullptr;
    std::vector<uint64_t> tuple_off;
};

struct BatchView {
    const vlgb_batch* b;
    const std::vector<uint32_t>* sub_pid;
    const ScanOut* so;
    uint64_t global_offset;  // bytes
    uint32_t sym_shift;      // log2(bytes per symbol)
};

inline uint64_t list_len(const BatchView& v, uint32_t sub, bool level0)
{
    uint32_t pid = (*v.sub_pid)[sub];
    return level0 ? v.so->own[pid] : v.so->tot[pid];
}

// Depth 1 of a query is joined from the right when its last list is far shorter than the list to its left: two
// binary searches per right element instead of one window search per left element (join.cuh, RevDesc).
inline uint64_t join_rev_factor()
{
    static const uint64_t f = getenv("VLGB_REV_FACTOR") ? strtoull(getenv("VLGB_REV_FACTOR"), nullptr, 10) : 16;  // tuning hook
    return f;
}
inline bool join_right_driven(const BatchView& v, const PlanQuery& pq)
{
    const uint64_t JOIN_REV_FACTOR = join_rev_factor();
    if (pq.k < 2 || JOIN_REV_FACTOR == 0) return false;
    const uint64_t nl = list_len(v, pq.s0 + pq.k - 2, pq.k == 2), nr = list_len(v, pq.s0 + pq.k - 1, false);
    return nr * JOIN_REV_FACTOR < nl;
}

// bytes of device state a query needs in the join/select stage (upper bound, used for chunking)
inline uint64_t query_state_bytes(const BatchView& v, const PlanQuery& pq)
{
    uint64_t slots = 0;
    for (uint32_t i = 0; i + 1 < pq.k; i++) slots += list_len(v, pq.s0 + i, i == 0);
    uint64_t sel = (list_len(v, pq.s0, true) + SEL_TILE - 1) / SEL_TILE * SEL_TILE;
    return slots * 36 + sel * 14 + 1024;
}

// Selection state of the last search of this thread, kept for vlgb_result_fixup (valid while no other search
// has run on the thread and the batch fitted one chunk): the join arena is left untouched, so succ[] and the
// flag bits of every query are still on the device.
struct Retained {
    bool valid = false;
    uint64_t serial = 0;
    SelDesc S;
    uint32_t nq = 0, ntiles = 0;
    std::vector<PlanQuery> qs;
    uint32_t* d_tile_count = nullptr;
    uint32_t* d_tile_prefix = nullptr;
    uint32_t* scratch = nullptr;
    uint64_t* d_counts = nullptr;
    uint64_t* d_tuple_off = nullptr;
    DeviceBuffer fix_buf;
};
inline Retained& retained()
{
    static thread_local Retained r;
    return r;
}

// `qs` sorted by k descending.  Results land in `co`; starts/tuples (if wanted) in the pinned block
// `hb` (grown by the caller-supplied allocator).
// `device_alloc(bytes)` returns device memory for the starts / tuples of the chunk: a block owned by the result when
// the caller keeps them on the device, the thread's reusable output buffer otherwise.
template <class PinnedAlloc, class DeviceAlloc>
inline void run_chunk(Context& c, const BatchView& v, const std::vector<PlanQuery>& qs, uint32_t want, bool keep_device,
                      PinnedAlloc&& pinned_alloc, DeviceAlloc&& device_alloc, ChunkOut& co)
{
    const uint32_t nq = (uint32_t)qs.size();
    const uint32_t kmax = qs[0].k;
    retained().valid = false;
    const vlgb_batch& B = *v.b;
    const ScanOut& so = *v.so;
    auto list_ptr = [&](uint32_t sub) { return (const uint32_t*)(so.d_lists + so.base[(*v.sub_pid)[sub]]); };
    auto plen = [&](uint32_t sub) { return B.pat_off[sub + 1] - B.pat_off[sub]; };

    // ---- sizes per depth
    struct DepthHost {
        uint32_t nseg = 0;                       // queries that have this depth (a prefix of the chunk)
        // left-driven pass: segments fmap[i] (query index); every array below is indexed by i
        std::vector<uint32_t> fmap, slot_off;
        std::vector<const uint32_t*> left, right_list;
        std::vector<uint32_t> right_len;
        std::vector<uint64_t> lo, hi;
        size_t o_fmap, o_slot_off, o_left, o_right_list, o_right_len, o_lo, o_hi;
        // right-driven pass (depth 1 only): segments rmap[i]
        std::vector<uint32_t> rmap, r_slot_off, r_left_len;
        std::vector<const uint32_t*> r_left, r_right;
        std::vector<uint64_t> r_lo, r_hi;
        size_t o_rmap, o_r_slot_off, o_r_left_len, o_r_left, o_r_right, o_r_lo, o_r_hi;
        uint64_t out_cap = 0;                    // upper bound of the viable slots = all left elements
        uint32_t *vpos, *vnxt, *vlast, *gseg, *glen, *vseg_f, *vseg_r, *ftotal;
    };
    std::vector<DepthHost> dh(kmax);  // index 1 .. kmax-1
    uint64_t max_slots = 0, max_rslots = 0;
    for (uint32_t d = 1; d < kmax; d++) {
        DepthHost& h = dh[d];
        uint32_t nseg = 0;
        while (nseg < nq && qs[nseg].k > d) nseg++;
        h.nseg = nseg;
        uint64_t run = 0, rrun = 0;
        for (uint32_t s = 0; s < nseg; s++) {
            const PlanQuery& pq = qs[s];
            const uint32_t lvl = pq.k - 1 - d, sub = pq.s0 + lvl;
            const uint64_t nl = list_len(v, sub, lvl == 0);
            // start-to-start bounds, vlg_index.hpp:95, in bytes (gaps are given in symbols)
            const uint64_t glo = sat_add64(sat_shl64(B.gap_min[sub], v.sym_shift), plen(sub));
            const uint64_t ghi = sat_add64(sat_shl64(B.gap_max[sub], v.sym_shift), plen(sub));
            h.out_cap += nl;
            if (d == 1 && join_right_driven(v, pq)) {
                h.rmap.push_back(s);
                h.r_slot_off.push_back((uint32_t)rrun);
                rrun += list_len(v, pq.s0 + pq.k - 1, false);
                h.r_left.push_back(list_ptr(sub));
                h.r_left_len.push_back((uint32_t)nl);
                h.r_right.push_back(list_ptr(pq.s0 + pq.k - 1));
                h.r_lo.push_back(glo);
                h.r_hi.push_back(ghi);
                continue;
            }
            h.fmap.push_back(s);
            h.slot_off.push_back((uint32_t)run);
            run += nl;
            h.left.push_back(list_ptr(sub));
            h.lo.push_back(glo);
            h.hi.push_back(ghi);
            if (d == 1) {
                h.right_list.push_back(list_ptr(pq.s0 + pq.k - 1));
                h.right_len.push_back((uint32_t)list_len(v, pq.s0 + pq.k - 1, false));
            }
        }
        if (run >= 0xFFFF0000ull || rrun >= 0xFFFF0000ull || h.out_cap >= 0xFFFF0000ull)
            throw CudaError{cudaErrorInvalidValue, __FILE__, __LINE__};  // chunking keeps this away
        h.slot_off.push_back((uint32_t)run);
        h.r_slot_off.push_back((uint32_t)rrun);
        max_slots = std::max(max_slots, run);
        max_rslots = std::max(max_rslots, rrun);
        c.stats.join_slots += run + rrun;
    }

    // ---- selection layout
    std::vector<uint32_t> sel_off(nq + 1), kq(nq), last_len(nq), list0_len(nq);
    std::vector<uint64_t> entry_local(nq), exit_init(nq), first_init(nq, ~uint64_t(0));
    std::vector<const uint32_t*> list0(nq), last_list(nq);
    uint64_t sel_run = 0;
    for (uint32_t s = 0; s < nq; s++) {
        const PlanQuery& pq = qs[s];
        sel_off[s] = (uint32_t)sel_run;
        uint64_t l0 = list_len(v, pq.s0, true);
        sel_run += (l0 + SEL_TILE - 1) / SEL_TILE * SEL_TILE;
        kq[s] = pq.k;
        uint64_t ll = plen(pq.s0 + pq.k - 1);
        last_len[s] = ll > 0xFFFFFFFFull ? 0xFFFFFFFFu : (uint32_t)ll;
        const uint64_t entry_bytes = sat_shl64(pq.entry, v.sym_shift);
        entry_local[s] = entry_bytes > v.global_offset ? entry_bytes - v.global_offset : 0;
        exit_init[s] = pq.entry;
        list0[s] = list_ptr(pq.s0);
        list0_len[s] = (uint32_t)l0;
        last_list[s] = list_ptr(pq.s0 + pq.k - 1);
    }
    if (sel_run >= 0xFFFF0000ull) throw CudaError{cudaErrorInvalidValue, __FILE__, __LINE__};
    sel_off[nq] = (uint32_t)sel_run;
    // upper bounds (every occurrence of P_1 viable); the real layout follows the join, from the V_0 sizes
    const uint32_t sel_total_ub = (uint32_t)sel_run;
    const uint32_t ntiles_ub = sel_total_ub / SEL_TILE;
    const uint32_t nchunks_ub = ntiles_ub / SEL_CHUNK_TILES + nq;

    // ---- control blob
    Packer ctl;
    for (uint32_t d = 1; d < kmax; d++) {
        DepthHost& h = dh[d];
        h.o_fmap = ctl.put(h.fmap);
        h.o_slot_off = ctl.put(h.slot_off);
        h.o_left = ctl.put(h.left);
        h.o_lo = ctl.put(h.lo);
        h.o_hi = ctl.put(h.hi);
        if (d == 1) {
            h.o_right_list = ctl.put(h.right_list);
            h.o_right_len = ctl.put(h.right_len);
            h.o_rmap = ctl.put(h.rmap);
            h.o_r_slot_off = ctl.put(h.r_slot_off);
            h.o_r_left_len = ctl.put(h.r_left_len);
            h.o_r_left = ctl.put(h.r_left);
            h.o_r_right = ctl.put(h.r_right);
            h.o_r_lo = ctl.put(h.r_lo);
            h.o_r_hi = ctl.put(h.r_hi);
        }
    }
    const size_t o_kq = ctl.put(kq), o_last_len = ctl.put(last_len),
                 o_entry = ctl.put(entry_local), o_exit = ctl.put(exit_init), o_first = ctl.put(first_init), o_list0 = ctl.put(list0),
                 o_list0_len = ctl.put(list0_len), o_last_list = ctl.put(last_list);
    std::vector<const uint32_t*> tab(5 * MAX_K, nullptr);  // depth_vpos | depth_vnxt | depth_vlast | depth_vseg | depth_vlen
    const size_t o_tab = ctl.put(tab);                     // patched below once the arena addresses are known
    std::vector<uint64_t> zeros64(nq, 0);
    const size_t o_tuple_off = ctl.put(zeros64);

    trace().mark("chunk: host layout");
    // ---- device memory
    size_t need = ctl.size() + 4096;
    const size_t max_jtiles = (max_slots + JF_TILE - 1) / JF_TILE;
    need += 3 * Arena::padded(max_slots, 4) + 2 * Arena::padded(max_jtiles + 4, 4) + Arena::padded(max_jtiles + 2, sizeof(JoinTile)) + Arena::padded(nq + 2, 4);  // staging, tile counts / prefix / segments, seg_rank
    for (uint32_t d = 1; d < kmax; d++)
        need += 3 * Arena::padded(dh[d].out_cap, 4) + 4 * Arena::padded(dh[d].nseg + 2, 4) + 256;
    need += 3 * Arena::padded(max_rslots + 4, 4);                                       // right-driven pass: first, count, prefix
    need += 2 * Arena::padded(nq, 8) + 2 * Arena::padded(nq, 4);                       // v0, vlast, v0_len, t0
    need += 2 * Arena::padded(sel_total_ub, 4) + Arena::padded((size_t)ntiles_ub * 32, 4);  // succ, exitp, flag bits
    need += 4 * Arena::padded(ntiles_ub + 2, 4);                                       // entry, spec, tile count / prefix
    need += 2 * Arena::padded(nchunks_ub + 1, 4);                                      // spec_exit, coalesce
    const size_t ctl2_bytes = 256 + Arena::padded(nq + 1, 4) + Arena::padded(nq, 4) + 2 * Arena::padded(nchunks_ub, 4) + Arena::padded(ntiles_ub + 1, 4);
    need += ctl2_bytes + 1024;                                                         // selection layout (second blob)
    need += Arena::padded(nq, 8) + Arena::padded(kmax + 1, 4);                         // counts, viable totals
    const size_t scratch_words = pfx_scratch_words(std::max<uint64_t>(std::max<uint64_t>(max_slots, max_rslots), sel_total_ub));
    need += Arena::padded(scratch_words, 4);
    c.join_arena.reserve(need);
    Arena& A = c.join_arena;
    uint8_t* d_ctl = A.take<uint8_t>(ctl.size());
    uint32_t* stage_x = A.take<uint32_t>(max_slots);
    uint32_t* stage_j = A.take<uint32_t>(max_slots);
    uint32_t* stage_last = A.take<uint32_t>(max_slots);
    uint32_t* seg_rank = A.take<uint32_t>(nq + 2);
    uint32_t* jtile_count = A.take<uint32_t>(max_jtiles + 4);
    uint32_t* jtile_prefix = A.take<uint32_t>(max_jtiles + 4);
    JoinTile* jtile_rec = A.take<JoinTile>(max_jtiles + 2);
    uint32_t* rev_first = A.take<uint32_t>(max_rslots + 4);
    uint32_t* rev_count = A.take<uint32_t>(max_rslots + 4);
    uint32_t* rev_prefix = A.take<uint32_t>(max_rslots + 4);
    for (uint32_t d = 1; d < kmax; d++) {
        const size_t t = dh[d].out_cap;
        dh[d].vpos = A.take<uint32_t>(t);
        dh[d].vnxt = A.take<uint32_t>(t);
        dh[d].vlast = A.take<uint32_t>(t);
        dh[d].gseg = A.take<uint32_t>(dh[d].nseg + 2);
        dh[d].glen = A.take<uint32_t>(dh[d].nseg + 2);
        dh[d].vseg_f = A.take<uint32_t>(dh[d].nseg + 2);
        dh[d].vseg_r = A.take<uint32_t>(dh[d].nseg + 2);
        dh[d].ftotal = A.take<uint32_t>(1);
        tab[0 * MAX_K + d] = dh[d].vpos;
        tab[1 * MAX_K + d] = dh[d].vnxt;
        tab[2 * MAX_K + d] = dh[d].vlast;
        tab[3 * MAX_K + d] = dh[d].gseg;
        tab[4 * MAX_K + d] = dh[d].glen;
    }
    memcpy(ctl.host.data() + o_tab, tab.data(), tab.size() * sizeof(tab[0]));
    const uint32_t** d_v0 = A.take<const uint32_t*>(nq);
    const uint32_t** d_vlast = A.take<const uint32_t*>(nq);
    uint32_t* d_v0_len = A.take<uint32_t>(nq);
    uint32_t* d_t0 = A.take<uint32_t>(nq);
    uint32_t* d_succ = A.take<uint32_t>(sel_total_ub);
    uint32_t* d_exitp = A.take<uint32_t>(sel_total_ub);
    uint32_t* d_flagbits = A.take<uint32_t>((size_t)ntiles_ub * 32);
    uint32_t* d_tile_count = A.take<uint32_t>(ntiles_ub + 2);
    uint32_t* d_tile_prefix = A.take<uint32_t>(ntiles_ub + 2);
    uint32_t* d_entry = A.take<uint32_t>(ntiles_ub + 1);
    uint32_t* d_spec = A.take<uint32_t>(ntiles_ub + 1);
    uint32_t* d_spec_exit = A.take<uint32_t>(nchunks_ub + 1);
    uint32_t* d_coalesce = A.take<uint32_t>(nchunks_ub + 1);
    uint8_t* d_ctl2 = A.take<uint8_t>(ctl2_bytes + 1024);
    uint64_t* d_counts = A.take<uint64_t>(nq);
    uint32_t* d_viable = A.take<uint32_t>(kmax + 1);
    uint32_t* scratch = A.take<uint32_t>(scratch_words);

    h2d(c, d_ctl, ctl.host.data(), ctl.host.size());
    c.stats.h2d_bytes += ctl.host.size();

    // ---- join, depth by depth: a left-driven pass over the segments of fmap, then (depth 1) a right-driven pass
    // over those of rmap; both publish where every query's viable list starts (gseg / glen, indexed by query)
    for (uint32_t d = 1; d < kmax; d++) {
        DepthHost& h = dh[d];
        const uint32_t nf = (uint32_t)h.fmap.size(), nr = (uint32_t)h.rmap.size();
        const uint32_t T = h.slot_off[nf];
        VLGB_CUDA(cudaMemsetAsync(h.gseg, 0, (size_t)(h.nseg + 1) * 4, c.stream));  // segments no pass publishes are empty
        VLGB_CUDA(cudaMemsetAsync(h.glen, 0, (size_t)(h.nseg + 1) * 4, c.stream));
        if (nf && T) {
            DepthDesc D;
            D.nseg = nf;
            D.total_slots = T;
            D.slot_off = dev_at<const uint32_t>(d_ctl, h.o_slot_off);
            D.left = dev_at<const uint32_t* const>(d_ctl, h.o_left);
            D.right_list = d == 1 ? dev_at<const uint32_t* const>(d_ctl, h.o_right_list) : nullptr;
            D.right_len = d == 1 ? dev_at<const uint32_t>(d_ctl, h.o_right_len) : nullptr;
            D.prev_vpos = d > 1 ? dh[d - 1].vpos : nullptr;
            D.prev_vlast = d > 1 ? dh[d - 1].vlast : nullptr;
            D.prev_vseg = d > 1 ? dh[d - 1].gseg : nullptr;  // at depth > 1 segment i IS query i (fmap = identity)
            D.prev_vlen = d > 1 ? dh[d - 1].glen : nullptr;
            D.lo = dev_at<const uint64_t>(d_ctl, h.o_lo);
            D.hi = dev_at<const uint64_t>(d_ctl, h.o_hi);
            D.stage_x = stage_x;
            D.stage_j = stage_j;
            D.stage_last = stage_last;
            D.seg_rank = seg_rank;
            D.tile_count = jtile_count;
            D.tile_prefix = jtile_prefix;
            D.tiles = jtile_rec;
            D.dbg = nullptr;
            static const bool join_debug = getenv("VLGB_JOIN_DEBUG") != nullptr;
            static thread_local DeviceBuffer dbg_buf;
            if (join_debug) {
                D.dbg = (unsigned long long*)dbg_buf.get(64);
                VLGB_CUDA(cudaMemsetAsync(D.dbg, 0, 48, c.stream));
            }
            D.total_out = h.ftotal;
            D.vpos = h.vpos;
            D.vnxt = h.vnxt;
            D.vlast = h.vlast;
            D.vseg = h.vseg_f;
            const uint32_t jtiles = (T + JF_TILE - 1) / JF_TILE;
            uint64_t right_words = 0;
            if (d == 1)
                for (uint32_t s = 0; s < nf; s++) right_words += h.right_len[s];
            c.begin(ST_JOIN_PROBE);
            join_tileseg_kernel<<<(jtiles + 1 + 127) / 128, 128, 0, c.stream>>>(D, jtiles);
            join_probe_kernel<<<jtiles, JF_THREADS, 0, c.stream>>>(D);
            c.end(ST_JOIN_PROBE, 2, (uint64_t)T * 4 + right_words * 4);  // + 12 B per viable slot staged, added below
            if (D.dbg) {
                unsigned long long h[6];
                VLGB_CUDA(cudaMemcpyAsync(h, D.dbg, 48, cudaMemcpyDeviceToHost, c.stream));
                VLGB_CUDA(cudaStreamSynchronize(c.stream));
                const char* names[3] = {"window", "dense", "multi-segment"};
                for (int i = 0; i < 3; i++)
                    fprintf(stderr, "[vlgb join] depth %u %-14s tiles %8llu  mean %8.1f us  sum %10.1f us\n", d, names[i], h[2 * i + 1],
                            h[2 * i + 1] ? h[2 * i] / 1e3 / h[2 * i + 1] : 0.0, h[2 * i] / 1e3);
            }
            c.begin(ST_JOIN_COMPACT);
            const uint32_t l = exclusive_scan(ValueU32{jtile_count}, jtiles, jtile_prefix, scratch, c.stream);
            join_vseg_kernel<<<(D.nseg + 1 + 255) / 256, 256, 0, c.stream>>>(D, jtiles);
            join_compact_kernel<<<jtiles, JC_THREADS, 0, c.stream>>>(D);
            join_publish_kernel<<<(nf + 1 + 255) / 256, 256, 0, c.stream>>>(h.vseg_f, dev_at<const uint32_t>(d_ctl, h.o_fmap), nf, nullptr,
                                                                          h.gseg, h.glen, h.ftotal);
            c.end(ST_JOIN_COMPACT, l + 3, (uint64_t)jtiles * 12);  // + 24 B per viable slot, added below
        } else {
            VLGB_CUDA(cudaMemsetAsync(h.ftotal, 0, 4, c.stream));
        }
        const uint32_t RT = nr ? h.r_slot_off[nr] : 0;
        if (nr && RT) {
            RevDesc RD;
            RD.nseg = nr;
            RD.total_slots = RT;
            RD.slot_off = dev_at<const uint32_t>(d_ctl, h.o_r_slot_off);
            RD.left = dev_at<const uint32_t* const>(d_ctl, h.o_r_left);
            RD.left_len = dev_at<const uint32_t>(d_ctl, h.o_r_left_len);
            RD.right = dev_at<const uint32_t* const>(d_ctl, h.o_r_right);
            RD.lo = dev_at<const uint64_t>(d_ctl, h.o_r_lo);
            RD.hi = dev_at<const uint64_t>(d_ctl, h.o_r_hi);
            RD.first = rev_first;
            RD.count = rev_count;
            RD.prefix = rev_prefix;
            RD.base = h.ftotal;
            RD.vpos = h.vpos;
            RD.vnxt = h.vnxt;
            RD.vlast = h.vlast;
            c.begin(ST_JOIN_PROBE);
            join_rev_count_kernel<<<(RT + 255) / 256, 256, 0, c.stream>>>(RD);
            c.end(ST_JOIN_PROBE, 1, (uint64_t)RT * 12);
            c.begin(ST_JOIN_COMPACT);
            const uint32_t l = exclusive_scan(ValueU32{rev_count}, RT, rev_prefix, scratch, c.stream);
            join_rev_vseg_kernel<<<(nr + 1 + 255) / 256, 256, 0, c.stream>>>(RD, h.vseg_r);
            join_rev_expand_kernel<<<(RT + 255) / 256, 256, 0, c.stream>>>(RD);
            join_publish_kernel<<<(nr + 1 + 255) / 256, 256, 0, c.stream>>>(h.vseg_r, dev_at<const uint32_t>(d_ctl, h.o_rmap), nr, h.ftotal,
                                                                          h.gseg, h.glen, d_viable + d);
            c.end(ST_JOIN_COMPACT, l + 3, (uint64_t)RT * 16);  // + 16 B per viable slot, added below
        } else {
            VLGB_CUDA(cudaMemcpyAsync(d_viable + d, h.ftotal, 4, cudaMemcpyDeviceToDevice, c.stream));
        }
    }

    // ---- selection
    SelDesc S;
    S.nq = nq;
    S.lay = nullptr;       // the layout fields are set below (built on the device once the V_0 sizes exist)
    S.sel_off = nullptr;
    S.tile_q = nullptr;
    S.kq = dev_at<const uint32_t>(d_ctl, o_kq);
    S.last_len = dev_at<const uint32_t>(d_ctl, o_last_len);
    S.entry_local = dev_at<const uint64_t>(d_ctl, o_entry);
    S.list0 = dev_at<const uint32_t* const>(d_ctl, o_list0);
    S.list0_len = dev_at<const uint32_t>(d_ctl, o_list0_len);
    S.depth_vpos = dev_at<const uint32_t* const>(d_ctl, o_tab);
    S.depth_vnxt = S.depth_vpos + MAX_K;
    S.depth_vlast = S.depth_vpos + 2 * MAX_K;
    S.depth_vseg = S.depth_vpos + 3 * MAX_K;
    S.depth_vlen = S.depth_vpos + 4 * MAX_K;
    S.last_list = dev_at<const uint32_t* const>(d_ctl, o_last_list);
    S.global_offset = v.global_offset;
    S.sym_shift = v.sym_shift;
    S.v0 = d_v0;
    S.vlast = d_vlast;
    S.v0_len = d_v0_len;
    S.t0 = d_t0;
    S.succ = d_succ;
    S.exitp = d_exitp;
    S.flagbits = d_flagbits;
    S.tile_count = d_tile_count;
    S.tile_prefix = d_tile_prefix;
    S.entry = d_entry;
    S.chunk_q = S.chunk_tile0 = S.q_chunk0 = nullptr;
    S.spec = d_spec;
    S.spec_exit = d_spec_exit;
    S.coalesce = d_coalesce;
    S.tuple_off = dev_at<const uint64_t>(d_ctl, o_tuple_off);
    S.starts = nullptr;
    S.tuples = nullptr;
    S.exit_cursor = dev_at<uint64_t>(d_ctl, o_exit);
    S.first_start = dev_at<uint64_t>(d_ctl, o_first);
    S.counts = d_counts;

    co.counts.assign(nq, 0);
    co.exit_cursor = exit_init;
    co.first_start = first_init;
    co.tuple_off.assign(nq, 0);
    co.matches = co.tuple_words = 0;
    co.h_starts = co.h_tuples = nullptr;
    std::vector<uint32_t> viable(kmax + 1, 0);

    // ---- V_0 of every query, then the selection layout -- tiles over the VIABLE starts only (on the configs of
    // BASELINE.json fewer than a sixth of the occurrences of P_1 are viable) -- built on the device: persistent
    // kernels read the tile / chunk counts from there, no size travels to the host between join and selection
    uint32_t lay[3] = {0, 0, 0};  // slots, tiles, chunks (read back with the match counts)
    if (sel_total_ub) {
        uint32_t* d_lay = reinterpret_cast<uint32_t*>(d_ctl2);
        size_t o2 = 256;
        S.lay = d_lay;
        S.sel_off = dev_at<uint32_t>(d_ctl2, o2);
        o2 += Arena::padded(nq + 1, 4);
        S.q_chunk0 = dev_at<uint32_t>(d_ctl2, o2);
        o2 += Arena::padded(nq, 4);
        S.chunk_q = dev_at<uint32_t>(d_ctl2, o2);
        o2 += Arena::padded(nchunks_ub, 4);
        S.chunk_tile0 = dev_at<uint32_t>(d_ctl2, o2);
        o2 += Arena::padded(nchunks_ub, 4);
        S.tile_q = dev_at<uint32_t>(d_ctl2, o2);
        // persistent kernels: CTAs that take tiles in turn.  Four waves' worth: tiles differ in cost (a dense V_0 gallops
        // beyond the staged window), finer turns balance better than exactly one resident wave (measured)
        const uint32_t pgrid = (uint32_t)c.sm_count * 32u;
        c.begin(ST_SELECT_SUCC);
        sel_resolve_kernel<<<(nq + 127) / 128, 128, 0, c.stream>>>(S);
        sel_layout_kernel<<<1, 1024, 0, c.stream>>>(S, d_lay);
        sel_fill_kernel<<<std::min<uint32_t>((ntiles_ub + 255) / 256, pgrid), 256, 0, c.stream>>>(S);
        c.end(ST_SELECT_SUCC, 3, (uint64_t)nq * 48);
        c.begin(ST_SELECT_TILE);
        sel_succtile_kernel<<<std::min<uint32_t>(ntiles_ub, pgrid), SELT_THREADS, 0, c.stream>>>(S);
        c.end(ST_SELECT_TILE, 1, 0);  // 16 B per slot (V_0 + tuple ends read, succ + exit written), added below
        c.begin(ST_SELECT_WALK);
        VLGB_CUDA(cudaMemsetAsync(d_entry, 0xFF, (size_t)(ntiles_ub + 1) * 4, c.stream));
        VLGB_CUDA(cudaMemsetAsync(d_spec, 0xFF, (size_t)(ntiles_ub + 1) * 4, c.stream));
        VLGB_CUDA(cudaMemsetAsync(d_coalesce, 0xFF, (size_t)(nchunks_ub + 1) * 4, c.stream));
        VLGB_CUDA(cudaMemsetAsync(d_tile_count, 0, (size_t)(ntiles_ub + 2) * 4, c.stream));  // tiles beyond the real count stay 0
        sel_spec_kernel<<<std::min<uint32_t>((nchunks_ub + 31) / 32, pgrid), 32, 0, c.stream>>>(S);
        sel_walk_kernel<<<(nq + 31) / 32, 32, 0, c.stream>>>(S);
        sel_entry_kernel<<<std::min<uint32_t>((ntiles_ub + 255) / 256, pgrid), 256, 0, c.stream>>>(S);
        sel_mark_kernel<<<std::min<uint32_t>((ntiles_ub + MARK_WARPS - 1) / MARK_WARPS, pgrid), MARK_WARPS * 32, 0, c.stream>>>(S);
        c.end(ST_SELECT_WALK, 4, 0);  // 5 B per slot, added below
        c.begin(ST_EMIT);
        uint32_t l = exclusive_scan(ValueU32{d_tile_count}, ntiles_ub, d_tile_prefix, scratch, c.stream);
        sel_counts_kernel<<<(nq + 127) / 128, 128, 0, c.stream>>>(S);
        c.end(ST_EMIT, l + 1, (uint64_t)ntiles_ub * 12);
        d2h(c, co.counts.data(), d_counts, (size_t)nq * 8);
        d2h(c, lay, d_lay, sizeof lay);
        if (kmax > 1)
            d2h(c, viable.data(), d_viable, (size_t)(kmax + 1) * 4);
        VLGB_CUDA(cudaGetLastError());
        trace().mark("chunk: join + select launched");
        sync_stream(c);  // sync: match counts
        trace().mark("chunk: sync counts");
        c.stats.d2h_bytes += (uint64_t)nq * 8 + (kmax + 1) * 4 + sizeof lay;
        const uint32_t sel_total = lay[0], ntiles = lay[1];
        c.stats.bytes[ST_SELECT_TILE] += (uint64_t)sel_total * 16;
        c.stats.bytes[ST_SELECT_WALK] += (uint64_t)sel_total * 5;
        for (uint32_t d = 1; d < kmax; d++)  // viable slots written; deeper levels read the previous viable list
        {
            c.stats.bytes[ST_JOIN_PROBE] += (uint64_t)viable[d] * 12;
            c.stats.bytes[ST_JOIN_COMPACT] += (uint64_t)viable[d] * 24 + (d > 1 ? (uint64_t)viable[d] * 4 : 0);
        }
        for (uint32_t d = 2; d < kmax; d++) c.stats.bytes[ST_JOIN_PROBE] += (uint64_t)viable[d - 1] * 4;

        for (uint32_t s = 0; s < nq; s++) {
            co.tuple_off[s] = co.tuple_words;
            co.matches += co.counts[s];
            co.tuple_words += co.counts[s] * qs[s].k;
        }
        const bool want_starts = (want & 3u) >= VLGB_WANT_STARTS;
        const bool want_tuples = (want & 3u) >= VLGB_WANT_TUPLES;
        uint64_t* d_starts = nullptr;
        uint64_t* d_tuples = nullptr;
        if (co.matches) {
            size_t ob = 0;
            if (want_starts) ob += Arena::padded(co.matches, 8);
            if (want_tuples) ob += Arena::padded(co.tuple_words, 8);
            uint8_t* o = (uint8_t*)device_alloc(ob + 256);
            if (want_starts) {
                d_starts = (uint64_t*)o;
                o += Arena::padded(co.matches, 8);
            }
            if (want_tuples) {
                d_tuples = (uint64_t*)o;
                h2d(c, d_ctl + o_tuple_off, co.tuple_off.data(), (size_t)nq * 8);
                c.stats.h2d_bytes += (uint64_t)nq * 8;
            }
            S.starts = d_starts;
            S.tuples = d_tuples;
            co.d_starts = d_starts;
            co.d_tuples = d_tuples;
            c.begin(ST_EMIT);
            sel_emit_kernel<<<(ntiles + EMIT_WARPS - 1) / EMIT_WARPS, EMIT_WARPS * 32, 0, c.stream>>>(S);
            c.end(ST_EMIT, 1, co.matches * (4 + (want_starts ? 8 : 0)) + (want_tuples ? co.tuple_words * 12 : 0));
            d2h(c, co.exit_cursor.data(), S.exit_cursor, (size_t)nq * 8);
            d2h(c, co.first_start.data(), S.first_start, (size_t)nq * 8);
            c.stats.d2h_bytes += (uint64_t)nq * 16;
            if (!keep_device && (want_starts || want_tuples)) {
                size_t hb = (want_starts ? co.matches * 8 : 0) + (want_tuples ? co.tuple_words * 8 : 0);
                uint8_t* h = (uint8_t*)pinned_alloc(hb);
                // VLGB_ASYNC_COPY: the lists travel on the copy stream, ordered behind the emit kernel by an event; the
                // caller's stream is free for the next search (the device block belongs to the result, see do_search)
                cudaStream_t cs = c.stream;
                if (want & VLGB_ASYNC_COPY) {
                    if (!c.copy_stream) {
                        VLGB_CUDA(cudaStreamCreateWithFlags(&c.copy_stream, cudaStreamNonBlocking));
                        VLGB_CUDA(cudaEventCreateWithFlags(&c.emit_done, cudaEventDisableTiming));
                    }
                    VLGB_CUDA(cudaEventRecord(c.emit_done, c.stream));
                    VLGB_CUDA(cudaStreamWaitEvent(c.copy_stream, c.emit_done, 0));
                    cs = c.copy_stream;
                    c.copy_issued = true;
                }
                if (want_starts) {
                    co.h_starts = (uint64_t*)h;
                    d2h_in_pieces(h, d_starts, co.matches * 8, cs, cs != c.stream);
                    h += co.matches * 8;
                    c.stats.d2h_bytes += co.matches * 8;
                }
                if (want_tuples) {
                    co.h_tuples = (uint64_t*)h;
                    d2h_in_pieces(h, d_tuples, co.tuple_words * 8, cs, cs != c.stream);
                    c.stats.d2h_bytes += co.tuple_words * 8;
                }
            }
            VLGB_CUDA(cudaGetLastError());
            sync_stream(c);
            trace().mark("chunk: emit + copies done");
        }
        Retained& T = retained();
        T.S = S;
        T.nq = nq;
        T.ntiles = ntiles;
        T.qs = qs;
        T.d_tile_count = d_tile_count;
        T.d_tile_prefix = d_tile_prefix;
        T.scratch = scratch;
        T.d_counts = d_counts;
        T.d_tuple_off = dev_at<uint64_t>(d_ctl, o_tuple_off);
        T.valid = true;
    }
}

// Re-entry of some queries of the retained chunk with new cursors (sel_refix_kernel).  `redo` holds
// chunk-local query indices in ascending order, `entry` their new whole-text cursors.  Outputs per redo query.
struct RefixOut {
    std::vector<uint64_t> counts, exit_cursor, first_start, tuple_off;
    uint64_t matches = 0, tuple_words = 0;
    uint64_t* h_starts = nullptr;
    uint64_t* h_tuples = nullptr;
    uint64_t* d_starts = nullptr;
    uint64_t* d_tuples = nullptr;
};

template <class PinnedAlloc, class DeviceAlloc>
inline void run_refix(Context& c, Retained& T, const std::vector<uint32_t>& redo, const std::vector<uint64_t>& entry,
                      uint32_t want, bool keep_device, PinnedAlloc&& pinned_alloc, DeviceAlloc&& device_alloc, RefixOut& ro)
{
    trace() = HostTrace();
    const uint32_t nr = (uint32_t)redo.size(), nq = T.nq;
    SelDesc S = T.S;
    std::vector<uint64_t> entry_local(nr);
    for (uint32_t i = 0; i < nr; i++) {
        const uint64_t eb = sat_shl64(entry[i], S.sym_shift);
        entry_local[i] = eb > S.global_offset ? eb - S.global_offset : 0;
    }
    // device scratch: redo list | cursors | flags | exit / first (nq each)
    const size_t o_q = 0, o_e = Arena::padded(nr, 4), o_flag = o_e + Arena::padded(nr, 8), o_exit = o_flag + Arena::padded(nq, 4),
                 o_first = o_exit + Arena::padded(nq, 8), total = o_first + Arena::padded(nq, 8);
    uint8_t* fb = (uint8_t*)T.fix_buf.get(total);
    h2d(c, fb + o_q, redo.data(), (size_t)nr * 4);
    h2d(c, fb + o_e, entry_local.data(), (size_t)nr * 8);
    VLGB_CUDA(cudaMemsetAsync(fb + o_flag, 0, (size_t)nq * 4, c.stream));
    VLGB_CUDA(cudaMemsetAsync(fb + o_first, 0xFF, (size_t)nq * 8, c.stream));
    c.stats.h2d_bytes += (uint64_t)nr * 12;
    RefixDesc rd;
    rd.n = nr;
    rd.q = (const uint32_t*)(fb + o_q);
    rd.entry_local = (const uint64_t*)(fb + o_e);
    rd.redo_flag = (uint32_t*)(fb + o_flag);
    rd.dirty = S.spec;  // [tiles]; the speculative entries of the search are not needed any more
    VLGB_CUDA(cudaMemsetAsync(S.spec, 0xFF, (size_t)T.ntiles * 4, c.stream));
    S.exit_cursor = (uint64_t*)(fb + o_exit);
    S.first_start = (uint64_t*)(fb + o_first);
    S.starts = nullptr;
    S.tuples = nullptr;
    c.begin(ST_SELECT_WALK);
    // (one re-entered query per warp: a long tile walk does not hold up 31 short ones)
    sel_refix_kernel<<<(nr + 3) / 4, 128, 0, c.stream>>>(S, rd);
    sel_remark_kernel<<<(T.ntiles + MARK_WARPS - 1) / MARK_WARPS, MARK_WARPS * 32, 0, c.stream>>>(S, rd, T.ntiles);
    c.end(ST_SELECT_WALK, 2, (uint64_t)T.ntiles * 136);
    c.begin(ST_EMIT);
    uint32_t l = exclusive_scan(ValueU32{T.d_tile_count}, T.ntiles, T.d_tile_prefix, T.scratch, c.stream);
    sel_counts_kernel<<<(nq + 127) / 128, 128, 0, c.stream>>>(S);
    c.end(ST_EMIT, l + 1, (uint64_t)T.ntiles * 12);
    std::vector<uint64_t> all_counts(nq, 0);
    d2h(c, all_counts.data(), T.d_counts, (size_t)nq * 8);
    VLGB_CUDA(cudaGetLastError());
    trace().mark("refix: walk + recount launched");
    sync_stream(c);
    trace().mark("refix: sync counts");
    c.stats.d2h_bytes += (uint64_t)nq * 8;

    ro.counts.assign(nr, 0);
    ro.exit_cursor.assign(nr, 0);
    ro.first_start.assign(nr, ~uint64_t(0));
    ro.tuple_off.assign(nr, 0);
    ro.matches = ro.tuple_words = 0;
    std::vector<uint64_t> tuple_off_all(nq, 0);
    for (uint32_t i = 0; i < nr; i++) {
        ro.counts[i] = all_counts[redo[i]];
        ro.tuple_off[i] = ro.tuple_words;
        tuple_off_all[redo[i]] = ro.tuple_words;
        ro.matches += ro.counts[i];
        ro.tuple_words += ro.counts[i] * T.qs[redo[i]].k;
    }
    if (!ro.matches) return;
    const bool want_starts = (want & 3u) >= VLGB_WANT_STARTS, want_tuples = (want & 3u) >= VLGB_WANT_TUPLES;
    size_t ob = 256;
    if (want_starts) ob += Arena::padded(ro.matches, 8);
    if (want_tuples) ob += Arena::padded(ro.tuple_words, 8);
    uint8_t* o = (uint8_t*)device_alloc(ob);
    uint64_t *d_starts = nullptr, *d_tuples = nullptr;
    if (want_starts) {
        d_starts = (uint64_t*)o;
        o += Arena::padded(ro.matches, 8);
    }
    if (want_tuples) {
        d_tuples = (uint64_t*)o;
        h2d(c, T.d_tuple_off, tuple_off_all.data(), (size_t)nq * 8);
        c.stats.h2d_bytes += (uint64_t)nq * 8;
    }
    S.starts = d_starts;
    S.tuples = d_tuples;
    ro.d_starts = d_starts;
    ro.d_tuples = d_tuples;
    c.begin(ST_EMIT);
    sel_emit_kernel<<<(T.ntiles + EMIT_WARPS - 1) / EMIT_WARPS, EMIT_WARPS * 32, 0, c.stream>>>(S);
    c.end(ST_EMIT, 1, ro.matches * (4 + (want_starts ? 8 : 0)) + (want_tuples ? ro.tuple_words * 12 : 0));
    std::vector<uint64_t> ex(nq), fs(nq);
    d2h(c, ex.data(), S.exit_cursor, (size_t)nq * 8);
    d2h(c, fs.data(), S.first_start, (size_t)nq * 8);
    c.stats.d2h_bytes += (uint64_t)nq * 16;
    if (!keep_device && (want_starts || want_tuples)) {
        size_t hb = (want_starts ? ro.matches * 8 : 0) + (want_tuples ? ro.tuple_words * 8 : 0);
        uint8_t* h = (uint8_t*)pinned_alloc(hb);
        cudaStream_t cs = c.stream;
        if ((want & VLGB_ASYNC_COPY) && c.copy_stream) {  // behind the result's first copy, which is still on its way
            VLGB_CUDA(cudaEventRecord(c.emit_done, c.stream));
            VLGB_CUDA(cudaStreamWaitEvent(c.copy_stream, c.emit_done, 0));
            cs = c.copy_stream;
            c.copy_issued = true;
        }
        if (want_starts) {
            ro.h_starts = (uint64_t*)h;
            d2h_in_pieces(h, d_starts, ro.matches * 8, cs, cs != c.stream);
            h += ro.matches * 8;
            c.stats.d2h_bytes += ro.matches * 8;
        }
        if (want_tuples) {
            ro.h_tuples = (uint64_t*)h;
            d2h_in_pieces(h, d_tuples, ro.tuple_words * 8, cs, cs != c.stream);
            c.stats.d2h_bytes += ro.tuple_words * 8;
        }
    }
    VLGB_CUDA(cudaGetLastError());
    trace().mark("refix: emit launched");
    sync_stream(c);
    trace().mark("refix: done");
    for (uint32_t i = 0; i < nr; i++) {
        if (!ro.counts[i]) continue;
        ro.exit_cursor[i] = ex[redo[i]];
        ro.first_start[i] = fs[redo[i]];
    }
}

}  // namespace vlgb
