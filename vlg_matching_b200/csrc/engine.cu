// engine.cu — libvlgb200.so: the C ABI of include/vlgb200.h on top of the sm_100a kernels
// (scan.cuh -> join.cuh -> select.cuh, prefix.cuh).  No CPU fallback: every compute entry point needs
// a CUDA device and fails with VLGB_ERR_CUDA otherwise.
#include <atomic>
#include <algorithm>
#include <cstdlib>
#include <memory>
#include <numeric>

#include "common.cuh"
#include "parse.hpp"
#include <mutex>
#include "search.cuh"
#include "phrases.cuh"
#include <queue>

using namespace vlgb;

// ---------------------------------------------------------------------------------------------
// handles
// ---------------------------------------------------------------------------------------------
struct vlgb_index {
    uint8_t* d_text = nullptr;  // n + PAD readable bytes
    bool owns = false;
    uint64_t n = 0, global_offset = 0, owned_len = 0, total_len = 0;  // all in BYTES
    uint32_t sym_shift = 0;  // log2(bytes per symbol): 0 = byte text, 1..3 = encoded integer alphabet
    WtSaOwned* rank = nullptr;  // suffix array + wavelet tree in rank blocks (vlgb_index_attach_rank), else nullptr
    uint32_t rank_flags = 0;
};
static constexpr uint64_t TEXT_PAD = 64;
static constexpr uint64_t MAX_SHARD = 0xFFFFF000ull;  // positions are shard-local u32 on the device

struct HostBlock {  // pinned result memory
    void* p = nullptr;
    size_t cap = 0;
};

struct DeviceBlock {  // device result memory owned by a result (VLGB_KEEP_ON_DEVICE)
    void* p = nullptr;
    size_t cap = 0;
};

struct vlgb_result {
    uint32_t nq = 0;
    uint32_t want = 0;
    std::vector<int32_t> status;
    std::vector<uint64_t> counts, exit_cursor, occ;
    std::vector<const uint64_t*> starts, tuples;          // host (pinned blocks)
    std::vector<const uint64_t*> dev_starts, dev_tuples;  // device (VLGB_KEEP_ON_DEVICE), nullptr otherwise
    std::vector<HostBlock> blocks;
    std::vector<DeviceBlock> dev_blocks;
    uint64_t total = 0;
    // VLGB_ASYNC_COPY: recorded on the copy stream behind the last host copy of this result; null once waited for
    mutable cudaEvent_t copy_done = nullptr;
};

namespace {

// pinned blocks are recycled across results (cudaHostAlloc is slow)
struct PinnedPool {
    static constexpr size_t PINNED_POOL_BLOCKS = 16;
    std::vector<HostBlock> free_list;
    HostBlock take(size_t bytes)
    {
        size_t best = SIZE_MAX;
        for (size_t i = 0; i < free_list.size(); i++)
            if (free_list[i].cap >= bytes && (best == SIZE_MAX || free_list[i].cap < free_list[best].cap)) best = i;
        // a much larger block is left for a larger request -- unless the pool is full: allocating would then end in a
        // cudaFreeHost when the block comes back, on every call (a small batch after a large one: 100 ms per search)
        if (best != SIZE_MAX && (free_list[best].cap <= 4 * bytes + (1u << 20) || free_list.size() >= PINNED_POOL_BLOCKS)) {
            HostBlock b = free_list[best];
            free_list.erase(free_list.begin() + best);
            return b;
        }
        HostBlock b;
        b.cap = bytes + bytes / 4 + 4096;
        if (cudaMallocHost(&b.p, b.cap) != cudaSuccess) throw CudaError{cudaErrorMemoryAllocation, __FILE__, __LINE__};
        return b;
    }
    void give(HostBlock b)
    {
        if (free_list.size() >= PINNED_POOL_BLOCKS) {  // several results in flight (VLGB_ASYNC_COPY) each hold a list block and re-entry blocks
            size_t small = 0;
            for (size_t i = 1; i < free_list.size(); i++)
                if (free_list[i].cap < free_list[small].cap) small = i;
            if (free_list[small].cap < b.cap) std::swap(free_list[small], b);
            cudaFreeHost(b.p);
            return;
        }
        free_list.push_back(b);
    }
};
PinnedPool& pinned_pool()
{
    static thread_local PinnedPool p;
    return p;
}

// device result blocks are recycled the same way (cudaMalloc / cudaFree synchronise the device)
struct DevicePool {
    static constexpr size_t DEVICE_POOL_BLOCKS = 8;
    std::vector<DeviceBlock> free_list;
    DeviceBlock take(size_t bytes)
    {
        size_t best = SIZE_MAX;
        for (size_t i = 0; i < free_list.size(); i++)
            if (free_list[i].cap >= bytes && (best == SIZE_MAX || free_list[i].cap < free_list[best].cap)) best = i;
        if (best != SIZE_MAX && (free_list[best].cap <= 4 * bytes + (1u << 20) || free_list.size() >= DEVICE_POOL_BLOCKS)) {  // as above
            DeviceBlock b = free_list[best];
            free_list.erase(free_list.begin() + best);
            return b;
        }
        DeviceBlock b;
        b.cap = bytes + bytes / 4 + 4096;
        cudaError_t e = cudaMalloc(&b.p, b.cap);
        if (e != cudaSuccess) {
            for (auto& f : free_list) cudaFree(f.p);  // give the cached blocks back and try once more
            free_list.clear();
            e = cudaMalloc(&b.p, b.cap);
        }
        if (e != cudaSuccess) throw CudaError{e, __FILE__, __LINE__};
        return b;
    }
    void give(DeviceBlock b)
    {
        if (free_list.size() >= DEVICE_POOL_BLOCKS) {
            size_t small = 0;
            for (size_t i = 1; i < free_list.size(); i++)
                if (free_list[i].cap < free_list[small].cap) small = i;
            if (free_list[small].cap < b.cap) std::swap(free_list[small], b);
            cudaFree(b.p);
            return;
        }
        free_list.push_back(b);
    }
};
DevicePool& device_pool()
{
    static thread_local DevicePool p;
    return p;
}

int cuda_fail(const CudaError& e)
{
    cudaGetLastError();
    int code = e.e == cudaErrorMemoryAllocation ? VLGB_ERR_NOMEM : VLGB_ERR_CUDA;
    return fail(code, "CUDA error %d (%s) at %s:%d", (int)e.e, cudaGetErrorString(e.e), e.file, e.line);
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// device / errors
// ---------------------------------------------------------------------------------------------
extern "C" int vlgb_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

extern "C" int vlgb_init(int device)
{
    try {
        Context& c = ctx();
        if (c.ready && c.device == device) return VLGB_OK;
        if (c.ready) return fail(VLGB_ERR_ARG, "vlgb_init: thread already bound to device %d", c.device);
        c.device = device;
        ensure_ready();
        return VLGB_OK;
    } catch (const CudaError& e) {
        ctx().device = -1;
        return cuda_fail(e);
    }
}

extern "C" const char* vlgb_last_error(void) { return last_error().c_str(); }
extern "C" const char* vlgb_version(void) { return "vlgb200 0.1 (sm_100a, abi 1)"; }
extern "C" const char* vlgb_stage_name(uint32_t stage) { return stage_name(stage); }

extern "C" int vlgb_last_stats(vlgb_stats* out)
{
    if (!out) return fail(VLGB_ERR_ARG, "vlgb_last_stats: null");
    Context& c = ctx();
    c.flush_timers();
    *out = c.stats;
    out->n_stages = ST_COUNT;
    return VLGB_OK;
}

extern "C" void* vlgb_stream_handle(void) { return ctx().ready ? (void*)ctx().stream : nullptr; }

extern "C" int vlgb_set_tuning(uint64_t stream_budget_bytes, uint64_t join_budget_bytes)
{
    Context& c = ctx();
    if (stream_budget_bytes) c.stream_budget = stream_budget_bytes;
    if (join_budget_bytes) c.join_budget = join_budget_bytes;
    return VLGB_OK;
}

// ---------------------------------------------------------------------------------------------
// parsing
// ---------------------------------------------------------------------------------------------
extern "C" int vlgb_parse_query(const char* str, size_t len, int syntax, vlgb_batch** out)
{
    if (!str || !out) return fail(VLGB_ERR_ARG, "vlgb_parse_query: null argument");
    if (syntax != VLGB_SYNTAX_LIBRARY && syntax != VLGB_SYNTAX_BENCHMARK)
        return fail(VLGB_ERR_ARG, "vlgb_parse_query: unknown syntax %d", syntax);
    *out = nullptr;
    ParsedQuery q;
    int rc = parse_query(str, len, syntax, q);
    if (rc == VLGB_ERR_LAZY) return fail(rc, "invalid gap description: expected '?' (lazy semantics)");
    if (rc == VLGB_ERR_MINMAX) return fail(rc, "invalid gap description: min-gap > max-gap");
    if (rc != VLGB_OK) return fail(rc, "invalid gap description: not a number");
    OwnedBatch* b = new (std::nothrow) OwnedBatch();
    if (!b) return fail(VLGB_ERR_NOMEM, "out of memory");
    b->add(q);
    b->seal();
    *out = &b->pub;
    return VLGB_OK;
}

extern "C" int vlgb_parse_batch(const char* const* strs, const size_t* lens, uint32_t n, int syntax, int32_t* status,
                                vlgb_batch** out)
{
    if ((!strs && n) || (!lens && n) || !out) return fail(VLGB_ERR_ARG, "vlgb_parse_batch: null argument");
    if (syntax != VLGB_SYNTAX_LIBRARY && syntax != VLGB_SYNTAX_BENCHMARK)
        return fail(VLGB_ERR_ARG, "vlgb_parse_batch: unknown syntax %d", syntax);
    *out = nullptr;
    OwnedBatch* b = new (std::nothrow) OwnedBatch();
    if (!b) return fail(VLGB_ERR_NOMEM, "out of memory");
    int first_err = VLGB_OK;
    for (uint32_t i = 0; i < n; i++) {
        ParsedQuery q;
        int rc = strs[i] ? parse_query(strs[i], lens[i], syntax, q) : VLGB_ERR_ARG;
        if (status) status[i] = rc;
        if (rc != VLGB_OK) {
            if (first_err == VLGB_OK) first_err = rc;
            q = ParsedQuery();  // keeps the numbering: an empty query (0 subpatterns) stands in
        }
        b->add(q);
    }
    b->seal();
    if (first_err != VLGB_OK && !status) {  // nobody to tell which query failed: fail the call like vlgb_parse_query
        delete b;
        if (first_err == VLGB_ERR_LAZY) return fail(first_err, "invalid gap description: expected '?' (lazy semantics)");
        if (first_err == VLGB_ERR_MINMAX) return fail(first_err, "invalid gap description: min-gap > max-gap");
        return fail(first_err, "invalid gap description: not a number");
    }
    *out = &b->pub;
    return VLGB_OK;
}

extern "C" int vlgb_parse_lines(const char* buf, size_t len, int syntax, vlgb_batch** out)
{
    if ((!buf && len) || !out) return fail(VLGB_ERR_ARG, "vlgb_parse_lines: null argument");
    std::vector<const char*> strs;
    std::vector<size_t> lens;
    size_t at = 0;
    while (at < len) {
        const char* nl = static_cast<const char*>(memchr(buf + at, '\n', len - at));
        const size_t end = nl ? (size_t)(nl - buf) : len;
        strs.push_back(buf + at);
        lens.push_back(end - at);
        at = end + 1;
    }
    return vlgb_parse_batch(strs.data(), lens.data(), (uint32_t)strs.size(), syntax, nullptr, out);
}

extern "C" void vlgb_batch_free(vlgb_batch* b)
{
    if (b) delete reinterpret_cast<OwnedBatch*>(b);  // pub is the first member
}

// ---------------------------------------------------------------------------------------------
// index: a text shard resident in HBM
// ---------------------------------------------------------------------------------------------
static int index_check(uint64_t n, uint64_t owned_len, const char* who)
{
    if (n > MAX_SHARD) return fail(VLGB_ERR_TOO_LARGE, "%s: shard of %llu bytes exceeds %llu (split the text into shards)", who,
                                   (unsigned long long)n, (unsigned long long)MAX_SHARD);
    if (owned_len > n) return fail(VLGB_ERR_ARG, "%s: owned_len > n", who);
    return VLGB_OK;
}

extern "C" int vlgb_index_build(const uint8_t* text, uint64_t n, uint64_t global_offset, uint64_t owned_len,
                                uint64_t total_len, vlgb_index** out)
{
    if (!out || (!text && n)) return fail(VLGB_ERR_ARG, "vlgb_index_build: null argument");
    *out = nullptr;
    int rc = index_check(n, owned_len, "vlgb_index_build");
    if (rc) return rc;
    if (n && memchr(text, 0, n))
        return fail(VLGB_ERR_ZERO_SYMBOL, "Error: File \"<memory>\" contains zero symbol.");  // construct.hpp:35-45
    try {
        ensure_ready();
        Context& c = ctx();
        std::unique_ptr<vlgb_index> idx(new vlgb_index());
        VLGB_CUDA(cudaMalloc(&idx->d_text, n + TEXT_PAD));
        idx->owns = true;
        VLGB_CUDA(cudaMemsetAsync(idx->d_text + n, 0, TEXT_PAD, c.stream));
        if (n) VLGB_CUDA(cudaMemcpyAsync(idx->d_text, text, n, cudaMemcpyHostToDevice, c.stream));
        VLGB_CUDA(cudaStreamSynchronize(c.stream));
        idx->n = n;
        idx->global_offset = global_offset;
        idx->owned_len = owned_len;
        idx->total_len = total_len;
        *out = idx.release();
        return VLGB_OK;
    } catch (const CudaError& e) {
        return cuda_fail(e);
    }
}

extern "C" int vlgb_index_build_device(const uint8_t* text_device, uint64_t n, uint64_t global_offset,
                                       uint64_t owned_len, uint64_t total_len, vlgb_index** out)
{
    if (!out || !text_device) return fail(VLGB_ERR_ARG, "vlgb_index_build_device: null argument");
    *out = nullptr;
    int rc = index_check(n, owned_len, "vlgb_index_build_device");
    if (rc) return rc;
    if (reinterpret_cast<uintptr_t>(text_device) & 15u)
        return fail(VLGB_ERR_ARG, "vlgb_index_build_device: the text must be 16-byte aligned");
    try {
        ensure_ready();
        vlgb_index* idx = new vlgb_index();
        idx->d_text = const_cast<uint8_t*>(text_device);
        idx->owns = false;
        idx->n = n;
        idx->global_offset = global_offset;
        idx->owned_len = owned_len;
        idx->total_len = total_len;
        *out = idx;
        return VLGB_OK;
    } catch (const CudaError& e) {
        return cuda_fail(e);
    }
}

static int read_file(const char* path, std::vector<uint8_t>& buf, uint64_t skip)
{
    FILE* f = fopen(path, "rb");
    if (!f) return fail(VLGB_ERR_IO, "cannot open %s", path);
    fseek(f, 0, SEEK_END);
    long long sz = ftell(f);
    if (sz < (long long)skip) {
        fclose(f);
        return fail(VLGB_ERR_IO, "%s: truncated", path);
    }
    fseek(f, (long)skip, SEEK_SET);
    buf.resize((size_t)(sz - skip));
    size_t got = buf.empty() ? 0 : fread(buf.data(), 1, buf.size(), f);
    fclose(f);
    if (got != buf.size()) return fail(VLGB_ERR_IO, "%s: short read", path);
    return VLGB_OK;
}

extern "C" int vlgb_index_build_file(const char* path, vlgb_index** out)
{
    if (!path || !out) return fail(VLGB_ERR_ARG, "vlgb_index_build_file: null argument");
    std::vector<uint8_t> buf;
    int rc = read_file(path, buf, 0);
    if (rc) return rc;
    rc = vlgb_index_build(buf.data(), buf.size(), 0, buf.size(), buf.size(), out);
    if (rc == VLGB_ERR_ZERO_SYMBOL) fail(rc, "Error: File \"%s\" contains zero symbol.", path);
    return rc;
}

// on-disk: magic, {n, global_offset, owned_len, total_len (bytes), sym_shift}, text bytes (vlg_index::serialize
// writes text then wt, vlg_index.hpp:183-191; there is no wavelet tree here)
static const char INDEX_MAGIC[8] = {'V', 'L', 'G', 'B', '2', '0', '0', '\2'};
static constexpr size_t INDEX_HDR = 8 + 5 * 8;

extern "C" int vlgb_index_save(const vlgb_index* idx, const char* path)
{
    if (!idx || !path) return fail(VLGB_ERR_ARG, "vlgb_index_save: null argument");
    try {
        ensure_ready();
        std::vector<uint8_t> buf(idx->n);
        if (idx->n) VLGB_CUDA(cudaMemcpy(buf.data(), idx->d_text, idx->n, cudaMemcpyDeviceToHost));
        FILE* f = fopen(path, "wb");
        if (!f) return fail(VLGB_ERR_IO, "cannot create %s", path);
        // the rank structures are rebuilt on the device at load time (seconds; cheaper than reading them back)
        uint64_t hdr[5] = {idx->n, idx->global_offset, idx->owned_len, idx->total_len,
                           idx->sym_shift | (idx->rank ? 0x100ull | ((uint64_t)idx->rank_flags << 16) : 0ull)};
        bool ok = fwrite(INDEX_MAGIC, 1, 8, f) == 8 && fwrite(hdr, 8, 5, f) == 5 &&
                  (idx->n == 0 || fwrite(buf.data(), 1, idx->n, f) == idx->n);
        ok = (fclose(f) == 0) && ok;
        return ok ? VLGB_OK : fail(VLGB_ERR_IO, "write to %s failed", path);
    } catch (const CudaError& e) {
        return cuda_fail(e);
    }
}

extern "C" int vlgb_index_load(const char* path, vlgb_index** out)
{
    if (!path || !out) return fail(VLGB_ERR_ARG, "vlgb_index_load: null argument");
    *out = nullptr;
    std::vector<uint8_t> buf;
    int rc = read_file(path, buf, 0);
    if (rc) return rc;
    if (buf.size() < INDEX_HDR || memcmp(buf.data(), INDEX_MAGIC, 8) != 0) return fail(VLGB_ERR_IO, "%s: not a vlgb200 index", path);
    uint64_t hdr[5];
    memcpy(hdr, buf.data() + 8, 40);
    if (buf.size() - INDEX_HDR != hdr[0] || (hdr[4] & 0xFF) > 3) return fail(VLGB_ERR_IO, "%s: size mismatch", path);
    rc = vlgb_index_build(buf.data() + INDEX_HDR, hdr[0], hdr[1], hdr[2], hdr[3], out);
    if (rc == VLGB_OK) (*out)->sym_shift = (uint32_t)(hdr[4] & 0xFF);
    if (rc == VLGB_OK && (hdr[4] & 0x100)) {
        rc = vlgb_index_attach_rank(*out, (uint32_t)(hdr[4] >> 16) & 0xFFu);
        if (rc != VLGB_OK) {
            vlgb_index_free(*out);
            *out = nullptr;
        }
    }
    return rc;
}

// ---- integer alphabets (vlg_index<int_alphabet_tag>, vlg_index.hpp:45,133): every symbol becomes W bytes
// (W = 1, 2, 4 or 8, the smallest that holds the largest symbol of the text), none of them 0, so the byte
// kernels apply unchanged: occurrences are kept only at multiples of W and positions are divided by W.
//   W = 1: the byte IS the symbol (1..255);  W > 1: big-endian base-255 digits, each stored + 1.
static const uint64_t SYM_LIMIT[4] = {256ull, 255ull * 255ull, 255ull * 255ull * 255ull * 255ull,
                                      17878103347812890625ull /* 255^8 */};

extern "C" int vlgb_encode_symbols(uint32_t width, const uint64_t* symbols, uint64_t n, uint8_t* out)
{
    if ((width != 1 && width != 2 && width != 4 && width != 8) || (n && (!symbols || !out)))
        return fail(VLGB_ERR_ARG, "vlgb_encode_symbols: bad argument");
    const uint32_t sh = width == 1 ? 0 : width == 2 ? 1 : width == 4 ? 2 : 3;
    for (uint64_t i = 0; i < n; i++) {
        uint64_t v = symbols[i];
        if (v == 0) return fail(VLGB_ERR_ZERO_SYMBOL, "Error: symbol sequence contains zero symbol.");
        if (v >= SYM_LIMIT[sh]) return fail(VLGB_ERR_TOO_LARGE, "symbol %llu does not fit %u encoded bytes", (unsigned long long)v, width);
        uint8_t* o = out + i * width;
        if (width == 1) { o[0] = (uint8_t)v; continue; }
        for (int d = (int)width - 1; d >= 0; d--) {
            o[d] = (uint8_t)(v % 255 + 1);
            v /= 255;
        }
    }
    return VLGB_OK;
}

extern "C" int vlgb_index_build_symbols(const uint64_t* symbols, uint64_t n, uint64_t global_offset, uint64_t owned_len,
                                        uint64_t total_len, vlgb_index** out)
{
    if (!out || (!symbols && n)) return fail(VLGB_ERR_ARG, "vlgb_index_build_symbols: null argument");
    *out = nullptr;
    uint64_t mx = 0;
    for (uint64_t i = 0; i < n; i++) {
        if (symbols[i] == 0) return fail(VLGB_ERR_ZERO_SYMBOL, "Error: File \"<memory>\" contains zero symbol.");
        mx = std::max(mx, symbols[i]);
    }
    uint32_t sh = 0;
    while (sh < 3 && mx >= SYM_LIMIT[sh]) sh++;
    if (mx >= SYM_LIMIT[sh]) return fail(VLGB_ERR_TOO_LARGE, "symbol %llu is not representable", (unsigned long long)mx);
    if (n > (MAX_SHARD >> sh)) return fail(VLGB_ERR_TOO_LARGE, "vlgb_index_build_symbols: shard too large (split the text into shards)");
    std::vector<uint8_t> enc((size_t)n << sh);
    int rc = vlgb_encode_symbols(1u << sh, symbols, n, enc.data());
    if (rc) return rc;
    rc = vlgb_index_build(enc.data(), n << sh, global_offset << sh, owned_len << sh, total_len << sh, out);
    if (rc == VLGB_OK) (*out)->sym_shift = sh;
    return rc;
}

extern "C" uint32_t vlgb_index_symbol_width(const vlgb_index* idx) { return idx ? 1u << idx->sym_shift : 0; }

extern "C" uint64_t vlgb_index_size(const vlgb_index* idx) { return idx ? idx->n >> idx->sym_shift : 0; }

extern "C" int vlgb_index_extract(const vlgb_index* idx, uint64_t pos, uint64_t len, uint8_t* out)
{
    if (!idx || (!out && len)) return fail(VLGB_ERR_ARG, "vlgb_index_extract: null argument");
    if (pos > idx->n || len > idx->n - pos) return fail(VLGB_ERR_ARG, "vlgb_index_extract: range outside the shard");
    try {
        ensure_ready();
        if (len) VLGB_CUDA(cudaMemcpy(out, idx->d_text + pos, len, cudaMemcpyDeviceToHost));
        return VLGB_OK;
    } catch (const CudaError& e) {
        return cuda_fail(e);
    }
}

// ---- rank path: suffix array + wavelet tree in flat rank blocks, built on the device (sa_build.cu)
extern "C" int vlgb_index_attach_rank(vlgb_index* idx, uint32_t flags)
{
    if (!idx) return fail(VLGB_ERR_ARG, "vlgb_index_attach_rank: null index");
    if (flags > VLGB_RANK_COMPACT) return fail(VLGB_ERR_ARG, "vlgb_index_attach_rank: unknown flags");
    if (idx->sym_shift) return fail(VLGB_ERR_ARG, "vlgb_index_attach_rank: byte texts only (symbol width %u)", 1u << idx->sym_shift);
    if (idx->n < 2) return fail(VLGB_ERR_ARG, "vlgb_index_attach_rank: text too short");
    if (idx->rank && idx->rank_flags == flags) return VLGB_OK;
    try {
        ensure_ready();
        Context& c = ctx();
        if (idx->rank) {
            wtsa_free(*idx->rank);
            delete idx->rank;
            idx->rank = nullptr;
        }
        std::unique_ptr<WtSaOwned> w(new WtSaOwned());
        try {
            wtsa_build(c.stream, idx->d_text, idx->n, flags != VLGB_RANK_COMPACT, *w);
        } catch (...) {
            wtsa_free(*w);
            throw;
        }
        idx->rank = w.release();
        idx->rank_flags = flags;
        return VLGB_OK;
    } catch (const CudaError& e) {
        return cuda_fail(e);
    }
}

extern "C" int vlgb_index_rank_info(const vlgb_index* idx, vlgb_rank_info* out)
{
    if (!idx || !out) return fail(VLGB_ERR_ARG, "vlgb_index_rank_info: null argument");
    memset(out, 0, sizeof *out);
    if (!idx->rank) return VLGB_OK;
    out->present = 1;
    out->levels = idx->rank->view.levels;
    out->compact = idx->rank->view.sa == nullptr;
    out->doubling_rounds = idx->rank->doubling_rounds;
    out->bytes = idx->rank->bytes();
    out->build_ms = idx->rank->build_ms;
    out->sa_ms = idx->rank->sa_ms;
    return VLGB_OK;
}

namespace {
__global__ void sa_extract_kernel(WtSa W, uint32_t first, uint32_t count, uint64_t* out)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) out[i] = wt_access(W, first + i);
}
}  // namespace

extern "C" int vlgb_index_sa_extract(const vlgb_index* idx, uint64_t first, uint64_t count, uint64_t* out)
{
    if (!idx || (!out && count)) return fail(VLGB_ERR_ARG, "vlgb_index_sa_extract: null argument");
    if (!idx->rank) return fail(VLGB_ERR_ARG, "vlgb_index_sa_extract: the index has no rank structures (vlgb_index_attach_rank)");
    if (first > idx->n || count > idx->n - first) return fail(VLGB_ERR_ARG, "vlgb_index_sa_extract: range outside the suffix array");
    if (!count) return VLGB_OK;
    try {
        ensure_ready();
        Context& c = ctx();
        DeviceBuffer buf;
        uint64_t* d = (uint64_t*)buf.get(count * 8);
        sa_extract_kernel<<<(unsigned)((count + 255) / 256), 256, 0, c.stream>>>(idx->rank->view, (uint32_t)first, (uint32_t)count, d);
        cudaError_t e = cudaMemcpyAsync(out, d, count * 8, cudaMemcpyDeviceToHost, c.stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(c.stream);
        buf.release();
        if (e != cudaSuccess) throw CudaError{e, __FILE__, __LINE__};
        return VLGB_OK;
    } catch (const CudaError& e) {
        return cuda_fail(e);
    }
}

// ---- vlgb_index_sa_extract_static: the same read, allocating nothing -----------------------------------------
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  vlgb_index_sa_extract
// stages the values in a device buffer it allocates; this variant stages them in a file-scope device array of
// VLGB_SA_EXTRACT_STATIC_MAX entries (1 MiB, held by every GPU that loads the library until the process exits).
namespace {
static __device__ __align__(256) uint64_t g_sa_static_scratch[VLGB_SA_EXTRACT_STATIC_MAX];
constexpr int SA_STATIC_MAX_DEVICES = 64;
uint64_t* g_sa_static_ptr[SA_STATIC_MAX_DEVICES] = {};  // device address of the array, per device (set-up, once)
std::mutex g_sa_static_mutex;

// In host code `g_sa_static_scratch` is a host-side stand-in, not a device address: the address a kernel can use
// comes from cudaGetSymbolAddress, asked once per device with THAT device current.  No allocation.
cudaError_t sa_static_setup(int device, uint64_t** out)
{
    std::lock_guard<std::mutex> lock(g_sa_static_mutex);
    if (!g_sa_static_ptr[device]) {
        int cur = -1;
        cudaError_t e = cudaGetDevice(&cur);
        if (e != cudaSuccess) return e;
        if (cur != device) return cudaErrorInvalidDevice;
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_sa_static_scratch);
        if (e != cudaSuccess) return e;
        g_sa_static_ptr[device] = static_cast<uint64_t*>(p);
    }
    *out = g_sa_static_ptr[device];
    return cudaSuccess;
}
}  // namespace

extern "C" int vlgb_index_sa_extract_static(const vlgb_index* idx, uint64_t first, uint64_t count, uint64_t* out)
{
    if (!idx || (!out && count)) return fail(VLGB_ERR_ARG, "vlgb_index_sa_extract_static: null argument");
    if (!idx->rank) return fail(VLGB_ERR_ARG, "vlgb_index_sa_extract_static: the index has no rank structures (vlgb_index_attach_rank)");
    if (count > (uint64_t)VLGB_SA_EXTRACT_STATIC_MAX)
        return fail(VLGB_ERR_ARG, "vlgb_index_sa_extract_static: count %llu exceeds the static scratch (VLGB_SA_EXTRACT_STATIC_MAX = %llu entries); "
                                  "split the range or use vlgb_index_sa_extract", (unsigned long long)count, (unsigned long long)VLGB_SA_EXTRACT_STATIC_MAX);
    if (first > idx->n || count > idx->n - first) return fail(VLGB_ERR_ARG, "vlgb_index_sa_extract_static: range outside the suffix array");
    Context& c = ctx();
    // the stream is created by vlgb_init (that call allocates); this one must not
    if (!c.ready) return fail(VLGB_ERR_ARG, "vlgb_index_sa_extract_static: call vlgb_init on this thread first");
    if (c.device < 0 || c.device >= SA_STATIC_MAX_DEVICES) return fail(VLGB_ERR_ARG, "vlgb_index_sa_extract_static: device %d not supported", c.device);
    if (!count) return VLGB_OK;
    try {
        int cur = -1;
        VLGB_CUDA(cudaGetDevice(&cur));
        if (cur != c.device) VLGB_CUDA(cudaSetDevice(c.device));
        uint64_t* scratch = nullptr;
        cudaError_t e = sa_static_setup(c.device, &scratch);
        if (e != cudaSuccess) throw CudaError{e, __FILE__, __LINE__};
        // every entry [0, count) of the scratch is written before it is read: nothing has to start at zero, and what the
        // previous call left behind beyond `count` is never looked at
        sa_extract_kernel<<<(unsigned)((count + 255) / 256), 256, 0, c.stream>>>(idx->rank->view, (uint32_t)first, (uint32_t)count, scratch);
        VLGB_CUDA(cudaGetLastError());
        VLGB_CUDA(cudaMemcpyAsync(out, scratch, count * 8, cudaMemcpyDefault, c.stream));  // out: host or device memory
        VLGB_CUDA(cudaStreamSynchronize(c.stream));
        return VLGB_OK;
    } catch (const CudaError& e) {
        return cuda_fail(e);
    }
}

// examples/generate-pattern.cpp:1-120 -- the `count` most frequent `length`-symbol phrases over a charset, in the order
// the reference prints them (its min-heap popped: ascending (occurrences, phrase)), padded by repetition (:107-114).
extern "C" int vlgb_index_frequent_phrases(const vlgb_index* idx, uint32_t length, uint32_t count, uint32_t charset, uint8_t* out,
                                           uint64_t* occ, uint32_t* n_distinct)
{
    if (!idx || (!out && count && length)) return fail(VLGB_ERR_ARG, "vlgb_index_frequent_phrases: null argument");
    if (!idx->rank) return fail(VLGB_ERR_ARG, "vlgb_index_frequent_phrases: the index has no rank structures (vlgb_index_attach_rank)");
    if (length == 0 || charset > 2) return fail(VLGB_ERR_ARG, "vlgb_index_frequent_phrases: bad length / charset");
    if (n_distinct) *n_distinct = 0;
    if (!count) return VLGB_OK;
    try {
        ensure_ready();
        PhraseRuns runs;
        frequent_runs(ctx(), idx->rank->view, idx->d_text, length, charset, count, runs);
        // the reference's heap rule, replayed in lexicographic order (generate-pattern.cpp:55-100): a phrase enters while
        // the heap is not full, or when it occurs MORE often than the heap's minimum (which is then dropped)
        typedef std::pair<uint32_t, std::string> entry;
        std::priority_queue<entry, std::vector<entry>, std::greater<entry>> heap;
        for (size_t r = 0; r < runs.occ.size(); r++) {
            if (heap.size() == count && runs.occ[r] <= heap.top().first) continue;
            heap.emplace(runs.occ[r], std::string(reinterpret_cast<const char*>(runs.grams.data()) + r * length, length));
            if (heap.size() > count) heap.pop();
        }
        std::vector<entry> res;
        while (!heap.empty() && res.size() < count) {
            res.push_back(heap.top());
            heap.pop();
        }
        if (n_distinct) *n_distinct = (uint32_t)res.size();
        if (res.empty()) res.emplace_back(0u, std::string(length, 'x'));
        const size_t si = res.size();
        while (res.size() < count) res.push_back(res[res.size() % si]);
        for (uint32_t i = 0; i < count; i++) {
            memcpy(out + (size_t)i * length, res[i].second.data(), length);
            if (occ) occ[i] = res[i].first;
        }
        return VLGB_OK;
    } catch (const CudaError& e) {
        return cuda_fail(e);
    } catch (const std::bad_alloc&) {
        return fail(VLGB_ERR_NOMEM, "out of host memory");
    }
}

extern "C" int vlgb_set_stage_timing(int on)
{
    Context& c = ctx();
    if (c.ready) c.flush_timers();
    c.stage_timing = on != 0;
    return VLGB_OK;
}

extern "C" int vlgb_set_path(int path)
{
    if (path != VLGB_PATH_AUTO && path != VLGB_PATH_SCAN && path != VLGB_PATH_RANK) return fail(VLGB_ERR_ARG, "vlgb_set_path: unknown path %d", path);
    ctx().list_path = path;
    return VLGB_OK;
}

extern "C" void vlgb_index_free(vlgb_index* idx)
{
    if (!idx) return;
    if (idx->owns && idx->d_text) cudaFree(idx->d_text);
    if (idx->rank) {
        wtsa_free(*idx->rank);
        delete idx->rank;
    }
    delete idx;
}

// ---------------------------------------------------------------------------------------------
// search
// ---------------------------------------------------------------------------------------------
struct vlgb_result_ext {  // what vlgb_result_fixup needs to re-enter the shard
    const vlgb_index* idx = nullptr;
    OwnedBatch batch;
    std::vector<uint64_t> entry, first_start;
    uint64_t serial = 0;  // Context::serial of the search that produced the result
};


namespace {

struct ResultImpl : vlgb_result {
    vlgb_result_ext ext;
};

const uint64_t EMPTY_LIST[1] = {0};

void copy_batch(const vlgb_batch* b, OwnedBatch& o)
{
    uint32_t nq = b->n_queries;
    uint32_t ns = b->sub_off[nq];
    o.sub_off.assign(b->sub_off, b->sub_off + nq + 1);
    o.pat_off.assign(b->pat_off, b->pat_off + ns + 1);
    o.bytes.assign(b->pat_bytes, b->pat_bytes + b->pat_off[ns]);
    o.gmin.assign(b->gap_min, b->gap_min + ns);
    o.gmax.assign(b->gap_max, b->gap_max + ns);
    o.seal();
}

int batch_check(const vlgb_batch* b)
{
    if (!b) return fail(VLGB_ERR_ARG, "vlgb_search_batch: null batch");
    if (b->n_queries && (!b->sub_off || !b->pat_off || !b->gap_min || !b->gap_max))
        return fail(VLGB_ERR_ARG, "vlgb_search_batch: null array inside the batch");
    if (b->n_queries && b->sub_off[0] != 0) return fail(VLGB_ERR_ARG, "vlgb_search_batch: sub_off[0] != 0");
    for (uint32_t q = 0; q < b->n_queries; q++)
        if (b->sub_off[q + 1] < b->sub_off[q]) return fail(VLGB_ERR_ARG, "vlgb_search_batch: sub_off not ascending");
    uint32_t ns = b->n_queries ? b->sub_off[b->n_queries] : 0;
    for (uint32_t s = 0; s < ns; s++)
        if (b->pat_off[s + 1] < b->pat_off[s]) return fail(VLGB_ERR_ARG, "vlgb_search_batch: pat_off not ascending");
    if (ns && b->pat_off[ns] && !b->pat_bytes) return fail(VLGB_ERR_ARG, "vlgb_search_batch: null pat_bytes");
    return VLGB_OK;
}

void do_search(Context& c, const vlgb_index* idx, const vlgb_batch* batch, const uint64_t* entry_cursor, uint32_t want,
               ResultImpl& R)
{
    trace() = HostTrace();
    const uint32_t nq = batch->n_queries;
    const uint32_t ns = nq ? batch->sub_off[nq] : 0;
    R.nq = nq;
    R.want = want;
    R.status.assign(nq, VLGB_OK);
    R.counts.assign(nq, 0);
    R.occ.assign(nq, 0);
    R.exit_cursor.assign(nq, 0);
    R.starts.assign(nq, nullptr);
    R.tuples.assign(nq, nullptr);
    R.dev_starts.assign(nq, nullptr);
    R.dev_tuples.assign(nq, nullptr);
    R.total = 0;
    R.ext.idx = idx;
    R.ext.entry.assign(nq, 0);
    R.ext.first_start.assign(nq, ~uint64_t(0));
    const bool want_starts = (want & 3u) >= VLGB_WANT_STARTS, want_tuples = (want & 3u) >= VLGB_WANT_TUPLES;
    const bool keep_device = (want & VLGB_KEEP_ON_DEVICE) != 0;
    const bool async_copy = (want & VLGB_ASYNC_COPY) != 0 && !keep_device;
    if (!async_copy) want &= ~VLGB_ASYNC_COPY;
    c.copy_issued = false;
    for (uint32_t q = 0; q < nq; q++) {
        R.ext.entry[q] = entry_cursor ? entry_cursor[q] : 0;
        R.exit_cursor[q] = R.ext.entry[q];
        if (want_starts && !keep_device) R.starts[q] = EMPTY_LIST;
        if (want_tuples && !keep_device) R.tuples[q] = EMPTY_LIST;
    }

    // ---- classify queries, collect the distinct subpatterns.  The outcome depends on the batch and on the shard's
    // geometry only, so it is kept (per thread) and reused when the very same batch comes again -- the next shard of
    // a sharded search, a repeated query file: the comparison is on the batch's bytes, not on a hash.
    const uint64_t halo = idx->n - idx->owned_len;
    const bool halo_truncated_by_shard = idx->global_offset + idx->n < idx->total_len;
    struct PlanCache {
        std::string key;
        PatternDict dict;
        std::vector<uint32_t> sub_pid;
        std::vector<PlanQuery> plan;  // entry cursors are filled per call
        std::vector<int32_t> status;
    };
    static thread_local PlanCache pc;
    static std::atomic<uint64_t> next_stamp{1};
    std::string key;
    {
        const uint64_t geo[6] = {nq, ns, idx->n, halo, (uint64_t)halo_truncated_by_shard, idx->sym_shift};
        const uint64_t nbytes = ns ? batch->pat_off[ns] : 0;
        key.reserve(sizeof geo + (nq + 1) * 4 + (ns + 1) * 8 + nbytes + (size_t)ns * 16);
        key.append(reinterpret_cast<const char*>(geo), sizeof geo);
        if (nq) {
            key.append(reinterpret_cast<const char*>(batch->sub_off), (size_t)(nq + 1) * sizeof batch->sub_off[0]);
            key.append(reinterpret_cast<const char*>(batch->pat_off), (size_t)(ns + 1) * sizeof batch->pat_off[0]);
            key.append(reinterpret_cast<const char*>(batch->pat_bytes), (size_t)nbytes);
            key.append(reinterpret_cast<const char*>(batch->gap_min), (size_t)ns * sizeof batch->gap_min[0]);
            key.append(reinterpret_cast<const char*>(batch->gap_max), (size_t)ns * sizeof batch->gap_max[0]);
        }
    }
    if (key != pc.key || !pc.dict.stamp) {
        pc.key.clear();
        pc.dict = PatternDict();
        pc.sub_pid.assign(ns, NONE32);
        pc.plan.clear();
        pc.status.assign(nq, VLGB_OK);
        for (uint32_t q = 0; q < nq; q++) {
            uint32_t s0 = batch->sub_off[q], s1 = batch->sub_off[q + 1], k = s1 - s0;
            if (k == 0) { pc.status[q] = VLGB_ERR_ARG; continue; }
            if (k > (uint32_t)MAX_K) { pc.status[q] = VLGB_ERR_TOO_LARGE; continue; }
            bool empty_sub = false, bad_width = false;
            uint64_t span = 0;
            for (uint32_t s = s0; s < s1; s++) {
                uint64_t len = batch->pat_off[s + 1] - batch->pat_off[s];
                if (len == 0) empty_sub = true;
                if (len & ((1u << idx->sym_shift) - 1u)) bad_width = true;
                span = sat_add64(span, len);
                if (s + 1 < s1) span = sat_add64(span, sat_shl64(batch->gap_max[s], idx->sym_shift));
            }
            if (empty_sub) { pc.status[q] = VLGB_ERR_EMPTY_SUB; continue; }  // SURVEY A.4 fence
            if (bad_width) { pc.status[q] = VLGB_ERR_ARG; continue; }        // not a whole number of encoded symbols
            // a match that starts in the owned part must end inside owned + halo
            if (halo_truncated_by_shard && span > halo + 1) { pc.status[q] = VLGB_ERR_TOO_LARGE; continue; }
            bool impossible = false;
            for (uint32_t s = s0; s < s1; s++) {
                uint64_t len = batch->pat_off[s + 1] - batch->pat_off[s];
                if (len > idx->n) { impossible = true; continue; }  // longer than the shard: cannot occur
                pc.sub_pid[s] = pc.dict.add(batch->pat_bytes + batch->pat_off[s], len);
            }
            if (impossible) continue;
            pc.plan.push_back({q, k, s0, 0});
        }
        pc.dict.stamp = next_stamp.fetch_add(1);
        pc.key = key;
    }
    R.status = pc.status;
    const PatternDict& dict = pc.dict;
    const std::vector<uint32_t>& sub_pid = pc.sub_pid;
    std::vector<PlanQuery> plan = pc.plan;
    for (PlanQuery& pq : plan) pq.entry = R.ext.entry[pq.q];

    // ---- scan: one sorted occurrence list per distinct subpattern
    ScanOut so;
    trace().mark("plan + dictionary");
    bool listed = false;
    if (idx->rank && c.list_path != VLGB_PATH_SCAN)
        listed = run_rank_lists(c, idx->d_text, idx->n, idx->owned_len, idx->rank->view, dict, so, c.list_path == VLGB_PATH_RANK);
    c.stats.list_path = listed ? VLGB_PATH_RANK : VLGB_PATH_SCAN;
    if (!listed) run_scan(c, idx->d_text, idx->n, idx->owned_len, idx->sym_shift, dict, so);
    BatchView view{batch, &sub_pid, &so, idx->global_offset, idx->sym_shift};

    trace().mark("scan done");
    // queries with an empty list have no match (vlg_index.hpp:316)
    std::vector<PlanQuery> live;
    for (const PlanQuery& pq : plan) {
        uint64_t occ = 0;
        bool any_empty = false;
        for (uint32_t i = 0; i < pq.k; i++) {
            uint64_t l = list_len(view, pq.s0 + i, false);
            occ += l;
            if (list_len(view, pq.s0 + i, i == 0) == 0) any_empty = true;
        }
        R.occ[pq.q] = occ;
        if (!any_empty) live.push_back(pq);
    }
    std::stable_sort(live.begin(), live.end(), [](const PlanQuery& a, const PlanQuery& b) { return a.k > b.k; });

    // ---- join + select, chunked by device memory
    const uint64_t budget = c.join_budget ? c.join_budget : (uint64_t)24 << 30;
    size_t i0 = 0;
    while (i0 < live.size()) {
        std::vector<uint64_t> depth_slots(live[i0].k + 1, 0);
        uint64_t bytes = 0, sel_slots = 0;
        size_t i1 = i0;
        while (i1 < live.size()) {
            const PlanQuery& pq = live[i1];
            uint64_t qb = query_state_bytes(view, pq);
            bool fits = (i1 == i0) || bytes + qb <= budget;
            uint64_t sel = (list_len(view, pq.s0, true) + SEL_TILE - 1) / SEL_TILE * SEL_TILE;
            if (sel_slots + sel >= 0xF0000000ull) fits = (i1 == i0);
            for (uint32_t d = 1; d < pq.k && fits; d++)
                if (depth_slots[d] + list_len(view, pq.s0 + pq.k - 1 - d, pq.k - 1 - d == 0) >= 0xF0000000ull) fits = (i1 == i0);
            if (!fits) break;
            for (uint32_t d = 1; d < pq.k; d++) depth_slots[d] += list_len(view, pq.s0 + pq.k - 1 - d, pq.k - 1 - d == 0);
            sel_slots += sel;
            bytes += qb;
            i1++;
        }
        std::vector<PlanQuery> chunk(live.begin() + i0, live.begin() + i1);
        ChunkOut co;
        run_chunk(c, view, chunk, want, keep_device,
                  [&](size_t nbytes) {
                      HostBlock hb = pinned_pool().take(nbytes ? nbytes : 8);
                      R.blocks.push_back(hb);
                      return hb.p;
                  },
                  [&](size_t nbytes) -> void* {
                      // results that outlive the call on the device (kept there, or still being copied) own their block
                      if (!keep_device && !async_copy) return c.out_buf.get(nbytes);
                      DeviceBlock db = device_pool().take(nbytes);
                      R.dev_blocks.push_back(db);
                      return db.p;
                  },
                  co);
        for (size_t s = 0; s < chunk.size(); s++) {
            uint32_t q = chunk[s].q;
            R.counts[q] = co.counts[s];
            R.exit_cursor[q] = co.exit_cursor[s];
            R.ext.first_start[q] = co.first_start[s];
        }
        uint64_t m = 0;
        for (size_t s = 0; s < chunk.size(); s++) {
            uint32_t q = chunk[s].q;
            if (co.counts[s]) {
                if (co.h_starts) R.starts[q] = co.h_starts + m;
                if (co.h_tuples) R.tuples[q] = co.h_tuples + co.tuple_off[s];
                if (keep_device && co.d_starts) R.dev_starts[q] = co.d_starts + m;
                if (keep_device && co.d_tuples) R.dev_tuples[q] = co.d_tuples + co.tuple_off[s];
            }
            m += co.counts[s];
        }
        R.total += co.matches;
        c.stats.query_chunks++;
        if (!(i0 == 0 && i1 == live.size())) retained().valid = false;  // only a single-chunk search stays re-enterable
        i0 = i1;
    }
    retained().serial = c.serial;
    R.ext.serial = c.serial;
    c.stats.segments = 1;
    if (c.copy_issued) {  // VLGB_ASYNC_COPY: the result's host lists are complete once the copy stream passes this event
        VLGB_CUDA(cudaEventCreateWithFlags(&R.copy_done, cudaEventDisableTiming));
        VLGB_CUDA(cudaEventRecord(R.copy_done, c.copy_stream));
        c.copy_issued = false;
    }
    trace().mark("search done");
}

}  // namespace

extern "C" int vlgb_search_batch(const vlgb_index* idx, const vlgb_batch* batch, const uint64_t* entry_cursor,
                                 uint32_t want, vlgb_result** out)
{
    if (!idx || !out) return fail(VLGB_ERR_ARG, "vlgb_search_batch: null argument");
    *out = nullptr;
    int rc = batch_check(batch);
    if (rc) return rc;
    if ((want & 3u) > VLGB_WANT_TUPLES) return fail(VLGB_ERR_ARG, "vlgb_search_batch: bad `want`");
    if (ctx().list_path == VLGB_PATH_RANK && !idx->rank)
        return fail(VLGB_ERR_ARG, "vlgb_search_batch: VLGB_PATH_RANK but the index has no rank structures (vlgb_index_attach_rank)");
    try {
        ensure_ready();
        Context& c = ctx();
        c.flush_timers();
        memset(&c.stats, 0, sizeof c.stats);
        c.serial++;
        retained().valid = false;
        if (c.stage_used) {  // the pinned staging block is reused: nothing of the previous call may still be in flight
            sync_stream(c);
            c.stage_used = 0;
        }
        std::unique_ptr<ResultImpl> R(new ResultImpl());
        try {
            do_search(c, idx, batch, entry_cursor, want, *R);
        } catch (...) {
            if (c.copy_stream) cudaStreamSynchronize(c.copy_stream);  // nothing may still be writing into the blocks
            if (R->copy_done) cudaEventDestroy(R->copy_done);
            for (auto& b : R->blocks) pinned_pool().give(b);
            for (auto& b : R->dev_blocks) device_pool().give(b);
            throw;
        }
        copy_batch(batch, R->ext.batch);
        *out = R.release();
        return VLGB_OK;
    } catch (const CudaError& e) {
        return cuda_fail(e);
    } catch (const std::bad_alloc&) {
        return fail(VLGB_ERR_NOMEM, "out of host memory");
    }
}

extern "C" uint32_t vlgb_result_queries(const vlgb_result* r) { return r ? r->nq : 0; }
extern "C" int32_t vlgb_result_status(const vlgb_result* r, uint32_t q) { return r && q < r->nq ? r->status[q] : VLGB_ERR_ARG; }
extern "C" uint64_t vlgb_result_count(const vlgb_result* r, uint32_t q) { return r && q < r->nq ? r->counts[q] : 0; }
extern "C" uint64_t vlgb_result_total(const vlgb_result* r) { return r ? r->total : 0; }
extern "C" const uint64_t* vlgb_result_counts(const vlgb_result* r) { return r ? r->counts.data() : nullptr; }
// VLGB_ASYNC_COPY: the host lists may be read once the copy stream has passed the result's event
static int wait_copies(const vlgb_result* r)
{
    if (!r || !r->copy_done) return VLGB_OK;
    const cudaError_t e = cudaEventSynchronize(r->copy_done);
    cudaEventDestroy(r->copy_done);
    r->copy_done = nullptr;
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(VLGB_ERR_CUDA, "CUDA error %d (%s) while waiting for the result copy", (int)e, cudaGetErrorString(e));
    }
    return VLGB_OK;
}
extern "C" int vlgb_result_wait(const vlgb_result* r) { return wait_copies(r); }
extern "C" const uint64_t* vlgb_result_starts(const vlgb_result* r, uint32_t q)
{
    if (!r || q >= r->nq) return nullptr;
    if (wait_copies(r) != VLGB_OK) return nullptr;
    return r->starts[q];
}
extern "C" const uint64_t* vlgb_result_tuples(const vlgb_result* r, uint32_t q)
{
    if (!r || q >= r->nq) return nullptr;
    if (wait_copies(r) != VLGB_OK) return nullptr;
    return r->tuples[q];
}
extern "C" const uint64_t* vlgb_result_exit_cursors(const vlgb_result* r) { return r ? r->exit_cursor.data() : nullptr; }
extern "C" const uint64_t* vlgb_result_first_starts(const vlgb_result* r)
{
    return r ? static_cast<const ResultImpl*>(r)->ext.first_start.data() : nullptr;
}
extern "C" uint64_t vlgb_result_occurrences(const vlgb_result* r, uint32_t q) { return r && q < r->nq ? r->occ[q] : 0; }

extern "C" const uint64_t* vlgb_result_device_starts(const vlgb_result* r, uint32_t q) { return r && q < r->nq ? r->dev_starts[q] : nullptr; }
extern "C" const uint64_t* vlgb_result_device_tuples(const vlgb_result* r, uint32_t q) { return r && q < r->nq ? r->dev_tuples[q] : nullptr; }

extern "C" void vlgb_result_free(vlgb_result* r)
{
    if (!r) return;
    wait_copies(r);  // the blocks go back to the pools: nothing may still be writing into them
    for (auto& b : r->blocks) pinned_pool().give(b);
    for (auto& b : r->dev_blocks) device_pool().give(b);
    delete static_cast<ResultImpl*>(r);
}

// Re-enter the shard with the true cursors.  A query is re-searched only when its reported set can
// change: the orbit of succ() depends on the entry cursor only through its first element
// (SURVEY A.3/A.5), so an entry that still lies at or before the first reported start (and not before
// the entry the result was computed with) leaves every match in place.
extern "C" int vlgb_result_fixup(vlgb_result* r, const uint64_t* entry_cursor, uint32_t* n_changed)
{
    if (!r || !entry_cursor) return fail(VLGB_ERR_ARG, "vlgb_result_fixup: null argument");
    ResultImpl& R = *static_cast<ResultImpl*>(r);
    if (n_changed) *n_changed = 0;
    std::vector<uint32_t> redo;
    for (uint32_t q = 0; q < R.nq; q++) {
        if (R.status[q] != VLGB_OK) continue;
        const uint64_t e_old = R.ext.entry[q], e_new = entry_cursor[q];
        if (e_new == e_old) continue;
        if (e_new > e_old && (R.counts[q] == 0 || e_new <= R.ext.first_start[q])) {
            if (R.counts[q] == 0) R.exit_cursor[q] = e_new;
            R.ext.entry[q] = e_new;
            continue;
        }
        redo.push_back(q);
    }
    if (redo.empty()) return VLGB_OK;
    try {
        Context& c = ctx();
        Retained& T = retained();
        if (c.ready && T.valid && T.serial == R.ext.serial && c.serial == R.ext.serial) {
            // fast path: the selection state of the search is still on the device (sel_refix_kernel)
            std::vector<int32_t> chunk_of(R.nq, -1);
            for (size_t s = 0; s < T.qs.size(); s++) chunk_of[T.qs[s].q] = (int32_t)s;
            std::vector<std::pair<uint32_t, uint32_t>> order;  // (chunk index, batch index)
            for (uint32_t q : redo) {
                if (chunk_of[q] < 0) {  // a query without a viable start: nothing to report under any cursor
                    R.exit_cursor[q] = entry_cursor[q];
                    R.ext.entry[q] = entry_cursor[q];
                } else {
                    order.emplace_back((uint32_t)chunk_of[q], q);
                }
            }
            std::sort(order.begin(), order.end());
            if (!order.empty()) {
                std::vector<uint32_t> rs(order.size());
                std::vector<uint64_t> re(order.size());
                for (size_t i = 0; i < order.size(); i++) {
                    rs[i] = order[i].first;
                    re[i] = entry_cursor[order[i].second];
                }
                vlgb_stats keep = c.stats;
                RefixOut ro;
                const bool keep_dev = (R.want & VLGB_KEEP_ON_DEVICE) != 0;
                // the re-emitted lists follow the result's first copy on the copy stream while that one is pending
                const bool async_copy = !keep_dev && (R.want & VLGB_ASYNC_COPY) && R.copy_done;
                static thread_local DeviceBuffer fix_out;
                c.copy_issued = false;
                run_refix(c, T, rs, re, (R.want & 3u) | (async_copy ? VLGB_ASYNC_COPY : 0u), keep_dev,
                          [&](size_t nbytes) {
                              HostBlock hb = pinned_pool().take(nbytes ? nbytes : 8);
                              R.blocks.push_back(hb);
                              return hb.p;
                          },
                          [&](size_t nbytes) -> void* {
                              if (!keep_dev && !async_copy) return fix_out.get(nbytes);
                              DeviceBlock db = device_pool().take(nbytes);
                              R.dev_blocks.push_back(db);
                              return db.p;
                          },
                          ro);
                if (c.copy_issued) {  // one event behind everything this result has on the copy stream
                    cudaEventDestroy(R.copy_done);
                    R.copy_done = nullptr;
                    VLGB_CUDA(cudaEventCreateWithFlags(&R.copy_done, cudaEventDisableTiming));
                    VLGB_CUDA(cudaEventRecord(R.copy_done, c.copy_stream));
                    c.copy_issued = false;
                }
                c.flush_timers();
                c.stats = keep;
                uint64_t m = 0;
                const bool host_lists = !(R.want & VLGB_KEEP_ON_DEVICE);
                for (size_t i = 0; i < order.size(); i++) {
                    const uint32_t q = order[i].second;
                    R.total = R.total - R.counts[q] + ro.counts[i];
                    R.counts[q] = ro.counts[i];
                    R.exit_cursor[q] = ro.counts[i] ? ro.exit_cursor[i] : entry_cursor[q];
                    R.ext.first_start[q] = ro.counts[i] ? ro.first_start[i] : ~uint64_t(0);
                    R.ext.entry[q] = entry_cursor[q];
                    if (host_lists && (R.want & 3u) >= VLGB_WANT_STARTS) R.starts[q] = ro.counts[i] && ro.h_starts ? ro.h_starts + m : EMPTY_LIST;
                    if (host_lists && (R.want & 3u) >= VLGB_WANT_TUPLES)
                        R.tuples[q] = ro.counts[i] && ro.h_tuples ? ro.h_tuples + ro.tuple_off[i] : EMPTY_LIST;
                    if (keep_dev) {
                        R.dev_starts[q] = ro.counts[i] && ro.d_starts ? ro.d_starts + m : nullptr;
                        R.dev_tuples[q] = ro.counts[i] && ro.d_tuples ? ro.d_tuples + ro.tuple_off[i] : nullptr;
                    }
                    m += ro.counts[i];
                }
            }
            if (n_changed) *n_changed = (uint32_t)redo.size();
            return VLGB_OK;
        }
    } catch (const CudaError& e) {
        return cuda_fail(e);
    } catch (const std::bad_alloc&) {
        return fail(VLGB_ERR_NOMEM, "out of host memory");
    }
    OwnedBatch sub;
    std::vector<uint64_t> sub_entry;
    const vlgb_batch& B = R.ext.batch.pub;
    sub.sub_off.push_back(0);
    sub.pat_off.push_back(0);
    for (uint32_t q : redo) {
        for (uint32_t s = B.sub_off[q]; s < B.sub_off[q + 1]; s++) {
            sub.bytes.insert(sub.bytes.end(), B.pat_bytes + B.pat_off[s], B.pat_bytes + B.pat_off[s + 1]);
            sub.pat_off.push_back(sub.bytes.size());
            sub.gmin.push_back(B.gap_min[s]);
            sub.gmax.push_back(B.gap_max[s]);
        }
        sub.sub_off.push_back((uint32_t)(sub.pat_off.size() - 1));
        sub_entry.push_back(entry_cursor[q]);
    }
    sub.seal();
    vlgb_result* nr = nullptr;
    vlgb_stats keep = ctx().stats;
    int rc = vlgb_search_batch(R.ext.idx, &sub.pub, sub_entry.data(), R.want, &nr);
    ctx().flush_timers();
    ctx().stats = keep;
    if (rc) return rc;
    ResultImpl& N = *static_cast<ResultImpl*>(nr);
    for (size_t i = 0; i < redo.size(); i++) {
        uint32_t q = redo[i];
        R.total = R.total - R.counts[q] + N.counts[i];
        R.counts[q] = N.counts[i];
        R.exit_cursor[q] = N.exit_cursor[i];
        R.starts[q] = N.starts[i];
        R.tuples[q] = N.tuples[i];
        R.dev_starts[q] = N.dev_starts[i];
        R.dev_tuples[q] = N.dev_tuples[i];
        R.status[q] = N.status[i];
        R.ext.entry[q] = entry_cursor[q];
        R.ext.first_start[q] = N.ext.first_start[i];
    }
    for (auto& b : N.blocks) R.blocks.push_back(b);  // the spliced lists live in the new result's blocks
    N.blocks.clear();
    for (auto& b : N.dev_blocks) R.dev_blocks.push_back(b);
    N.dev_blocks.clear();
    vlgb_result_free(nr);
    if (n_changed) *n_changed = (uint32_t)redo.size();
    return VLGB_OK;
}

// ---------------------------------------------------------------------------------------------
// multi-GPU: overlap-sharded text, one rank per GPU, ONE small collective per round
// ---------------------------------------------------------------------------------------------
#include "comm.hpp"

extern "C" int vlgb_chain_plan(const uint64_t* rows, const uint64_t* used, uint32_t n_shards, uint32_t nq, uint64_t* entries,
                               uint8_t* redo)
{
    if ((!rows || !used || !entries) && n_shards && nq) return fail(VLGB_ERR_ARG, "vlgb_chain_plan: null argument");
    if (!redo && n_shards) return fail(VLGB_ERR_ARG, "vlgb_chain_plan: null argument");
    return (int)chain_plan(rows, used, n_shards, nq, entries, redo);
}

extern "C" int vlgb_comm_unique_id(uint8_t* id)
{
    if (!id) return fail(VLGB_ERR_ARG, "vlgb_comm_unique_id: null");
    NcclApi& api = nccl_api();
    if (!api.ok) return fail(VLGB_ERR_CUDA, "NCCL is not available (libnccl.so.2 could not be loaded)");
    static_assert(sizeof(ncclUniqueId) == VLGB_COMM_ID_BYTES, "ncclUniqueId size");
    ncclUniqueId u;
    ncclResult_t rc = api.GetUniqueId(&u);
    if (rc != ncclSuccess) return fail(VLGB_ERR_CUDA, "ncclGetUniqueId: %s", api.GetErrorString(rc));
    memcpy(id, &u, sizeof u);
    return VLGB_OK;
}

extern "C" int vlgb_comm_init(const uint8_t* id, int n_ranks, int rank)
{
    if (n_ranks < 1 || rank < 0 || rank >= n_ranks) return fail(VLGB_ERR_ARG, "vlgb_comm_init: rank %d of %d", rank, n_ranks);
    Comm& cm = comm();
    if (cm.comm) return fail(VLGB_ERR_ARG, "vlgb_comm_init: this thread already has a communicator");
    try {
        ensure_ready();
        if (n_ranks > 1) {
            if (!id) return fail(VLGB_ERR_ARG, "vlgb_comm_init: null id");
            NcclApi& api = nccl_api();
            if (!api.ok) return fail(VLGB_ERR_CUDA, "NCCL is not available (libnccl.so.2 could not be loaded)");
            ncclUniqueId u;
            memcpy(&u, id, sizeof u);
            ncclResult_t rc = api.CommInitRank(&cm.comm, n_ranks, u, rank);
            if (rc != ncclSuccess) {
                cm.comm = nullptr;
                return fail(VLGB_ERR_CUDA, "ncclCommInitRank: %s", api.GetErrorString(rc));
            }
        }
        cm.n_ranks = n_ranks;
        cm.rank = rank;
        return VLGB_OK;
    } catch (const CudaError& e) {
        return cuda_fail(e);
    }
}

extern "C" int vlgb_comm_destroy(void)
{
    Comm& cm = comm();
    if (cm.comm) nccl_api().CommDestroy(cm.comm);
    cm.comm = nullptr;
    cm.n_ranks = 1;
    cm.rank = 0;
    return VLGB_OK;
}

namespace {

// row of this rank: its local shards are consecutive in the text and already chained exactly
void local_row(vlgb_result* const* res, uint32_t n_local, uint32_t nq, uint64_t* row)
{
    for (uint32_t q = 0; q < nq; q++) {
        uint64_t cnt = 0, first = ~uint64_t(0);
        for (uint32_t j = 0; j < n_local; j++) {
            const ResultImpl& R = *static_cast<const ResultImpl*>(res[j]);
            if (R.counts[q] && !cnt) first = R.ext.first_start[q];
            cnt += R.counts[q];
        }
        row[3 * q] = cnt;
        row[3 * q + 1] = static_cast<const ResultImpl*>(res[n_local - 1])->exit_cursor[q];
        row[3 * q + 2] = first;
    }
}

}  // namespace

extern "C" int vlgb_search_sharded(const vlgb_index* const* shards, uint32_t n_local, const vlgb_batch* batch, uint32_t want,
                                   vlgb_result** results, uint64_t* global_counts, vlgb_shard_stats* st)
{
    if (!shards || !n_local || !results) return fail(VLGB_ERR_ARG, "vlgb_search_sharded: null argument");
    int rc = batch_check(batch);
    if (rc) return rc;
    const uint32_t nq = batch->n_queries;
    for (uint32_t j = 0; j < n_local; j++) results[j] = nullptr;
    auto cleanup = [&](int code) {
        for (uint32_t j = 0; j < n_local; j++) {
            vlgb_result_free(results[j]);
            results[j] = nullptr;
        }
        return code;
    };
    vlgb_shard_stats S;
    memset(&S, 0, sizeof S);
    Comm& cm = comm();
    S.n_ranks = (uint32_t)cm.n_ranks;
    // ---- the local shards, left to right; shard j enters with the cursors shard j-1 left (exact inside the rank)
    vlgb_stats acc;
    memset(&acc, 0, sizeof acc);
    std::vector<uint64_t> entry(nq, 0);
    for (uint32_t j = 0; j < n_local; j++) {
        rc = vlgb_search_batch(shards[j], batch, j ? entry.data() : nullptr, want, &results[j]);
        if (rc) return cleanup(rc);
        const ResultImpl& R = *static_cast<const ResultImpl*>(results[j]);
        entry = R.exit_cursor;
        if (n_local > 1) {  // stage times of all local shards add up in vlgb_last_stats
            Context& c = ctx();
            c.flush_timers();
            for (uint32_t s = 0; s < ST_COUNT; s++) {
                acc.ms[s] += c.stats.ms[s];
                acc.bytes[s] += c.stats.bytes[s];
                acc.launches[s] += c.stats.launches[s];
            }
            acc.h2d_bytes += c.stats.h2d_bytes;
            acc.d2h_bytes += c.stats.d2h_bytes;
            acc.distinct_patterns = c.stats.distinct_patterns;
            acc.scan_passes += c.stats.scan_passes;
            acc.segments += c.stats.segments;
            acc.query_chunks += c.stats.query_chunks;
            acc.total_occurrences += c.stats.total_occurrences;
            acc.join_slots += c.stats.join_slots;
            acc.list_path = c.stats.list_path;
            acc.host_syncs += c.stats.host_syncs;
        }
    }
    if (n_local > 1) ctx().stats = acc;
    const uint32_t G = (uint32_t)cm.n_ranks;
    std::vector<uint64_t> table((size_t)G * nq * 3);
    try {
        if (G > 1 && nq) {
            Context& c = ctx();
            NcclApi& api = nccl_api();
            const double t0 = HostTrace::now();
            const size_t row_words = (size_t)nq * 3;
            uint64_t* d_buf = (uint64_t*)cm.dev.get((G + 1) * row_words * 8);
            uint64_t* h_buf = (uint64_t*)cm.host.get((G + 1) * row_words * 8);
            std::vector<uint64_t> used((size_t)G * nq, 0), entries((size_t)G * nq, 0);
            std::vector<uint8_t> redo(G, 0);
            for (uint32_t round = 0; round <= G; round++) {
                local_row(results, n_local, nq, h_buf);
                VLGB_CUDA(cudaMemcpyAsync(d_buf, h_buf, row_words * 8, cudaMemcpyHostToDevice, c.stream));
                ncclResult_t nrc = api.AllGather(d_buf, d_buf + row_words, row_words, ncclUint64, cm.comm, c.stream);
                if (nrc != ncclSuccess) return cleanup(fail(VLGB_ERR_CUDA, "ncclAllGather: %s", api.GetErrorString(nrc)));
                // written into the pinned buffer by a kernel: the copy engine may be busy with a result copy (VLGB_ASYNC_COPY)
                small_d2h_kernel<<<8, 256, 0, c.stream>>>(reinterpret_cast<uint32_t*>(h_buf + row_words),
                                                          reinterpret_cast<const uint32_t*>(d_buf + row_words), (uint32_t)(G * row_words * 2));
                VLGB_CUDA(cudaGetLastError());
                VLGB_CUDA(cudaStreamSynchronize(c.stream));
                trace().mark("exchange: all-gather done (includes the wait for the slowest rank)");
                memcpy(table.data(), h_buf + row_words, G * row_words * 8);
                S.rounds++;
                const uint32_t n_redo = chain_plan(table.data(), used.data(), G, nq, entries.data(), redo.data());
                // record the cursors on this rank (harmless moves change no match but do change the exit cursor of
                // queries without one); search again only what can change -- vlgb_result_fixup decides per query
                const uint64_t* mine = entries.data() + (size_t)cm.rank * nq;
                if (memcmp(mine, used.data() + (size_t)cm.rank * nq, (size_t)nq * 8) != 0) {
                    std::vector<uint64_t> e(mine, mine + nq);
                    for (uint32_t j = 0; j < n_local; j++) {
                        uint32_t changed = 0;
                        vlgb_stats keep = c.stats;
                        rc = vlgb_result_fixup(results[j], e.data(), &changed);
                        c.stats = keep;
                        if (rc) return cleanup(rc);
                        S.reentered_queries += changed;
                        e = static_cast<const ResultImpl*>(results[j])->exit_cursor;
                    }
                }
                used = entries;
                trace().mark("exchange: cursors chained, re-entry done");
                if (!n_redo) break;
                S.reentered_shards += n_redo;
            }
            // rows of shards without a match carry the cursor through
            S.exchange_ms = (float)(HostTrace::now() - t0);
        } else {
            local_row(results, n_local, nq, table.data());
        }
    } catch (const CudaError& e) {
        return cleanup(cuda_fail(e));
    } catch (const std::bad_alloc&) {
        return cleanup(fail(VLGB_ERR_NOMEM, "out of host memory"));
    }
    if (global_counts)
        for (uint32_t q = 0; q < nq; q++) {
            uint64_t s = 0;
            for (uint32_t g = 0; g < G; g++) s += table[((size_t)g * nq + q) * 3];
            global_counts[q] = s;
        }
    for (uint32_t q = 0; q < nq; q++)
        for (uint32_t g = 0; g < G; g++) S.global_matches += table[((size_t)g * nq + q) * 3];
    if (st) *st = S;
    return VLGB_OK;
}
