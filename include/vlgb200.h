/* vlgb200.h — C ABI of libvlgb200.so, the B200 (sm_100a) variable-length-gap matcher.
 *
 * The reference (olydis/vlg_matching, an sdsl-lite fork) has no FFI boundary: its callers bind at
 * compile time to C++ templates.  These entry points are what a binding for that path would call;
 * each one names the reference interface it replaces (paths relative to /root/reference).
 * The C++ header vlg_matching_b200/csrc/host/vlg_b200.hpp wraps this ABI behind the reference's own
 * names (vlg_index / construct / locate / count and the benchmark's 7-method index interface);
 * INTEGRATION.md shows the binding a maintainer would add.
 *
 * Conventions
 *  - every call returns VLGB_OK (0) or a negative code and never throws across the ABI;
 *    vlgb_last_error() gives the text for the calling thread;
 *  - plain pointers and sizes only; inputs are HOST memory owned by the caller unless a name says
 *    "_device"; outputs are owned by the library until the matching *_free call;
 *  - all searching runs in CUDA kernels.  There is no CPU fallback: without a usable GPU every
 *    compute entry point fails with VLGB_ERR_CUDA.
 *  - positions are 0-based byte offsets into the WHOLE text (uint64), ascending per query.
 */
#ifndef VLGB200_H
#define VLGB200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VLGB_ABI_VERSION 2

enum vlgb_status {
    VLGB_OK = 0,
    VLGB_ERR_CUDA = -1,       /* no device / CUDA runtime failure */
    VLGB_ERR_ARG = -2,        /* null pointer, bad size, bad handle */
    VLGB_ERR_LAZY = -3,       /* std::runtime_error "invalid gap description: expected '?' (lazy semantics)"  include/sdsl/vlg_index.hpp:97-99 */
    VLGB_ERR_MINMAX = -4,     /* std::runtime_error "invalid gap description: min-gap > max-gap"              include/sdsl/vlg_index.hpp:92-94 */
    VLGB_ERR_NUMBER = -5,     /* std::stoull failure inside the gap                                            include/sdsl/vlg_index.hpp:90-91 */
    VLGB_ERR_EMPTY_SUB = -6,  /* FENCED: empty subpattern (reference matches every suffix)          */
    VLGB_ERR_ZERO_SYMBOL = -7,/* std::logic_error "... contains zero symbol."                                  include/sdsl/construct.hpp:35-45 */
    VLGB_ERR_NOMEM = -8,
    VLGB_ERR_IO = -9,
    VLGB_ERR_TOO_LARGE = -10  /* shard > 4 GiB - 4 KiB, query with > 64 subpatterns, or a query span that
                                 does not fit the shard's halo */
};

typedef struct vlgb_index vlgb_index;   /* a text shard resident in HBM  — replaces sdsl::vlg_index<> (vlg_index.hpp:116-199) */
typedef struct vlgb_result vlgb_result; /* the matches of one batch       — replaces container<vlg_iterator> (vlg_index.hpp:395-401) */

/* ---- device ------------------------------------------------------------------------------- */

/* Number of CUDA devices visible to this process (0 without a GPU; never an error). */
int vlgb_device_count(void);
/* Bind the calling process to one device (one process per GPU).  Idempotent. */
int vlgb_init(int device);
const char* vlgb_last_error(void);
const char* vlgb_version(void);

/* ---- query parsing  (replaces gapped_pattern_query ctor, vlg_index.hpp:54-105, and the benchmark's
 *      gapped_pattern ctor, benchmark/gapped-matching/include/utils.hpp:19-71) ---------------------- */

#define VLGB_SYNTAX_LIBRARY   0  /* P1.{a,b}?P2 ... : lazy '?' mandatory */
#define VLGB_SYNTAX_BENCHMARK 1  /* P1.{a,b}P2  ... : no '?'             */

/* A batch of parsed queries in CSR form.
 *   query q owns subpatterns s in [sub_off[q], sub_off[q+1])
 *   subpattern s owns bytes  pat_bytes[pat_off[s] .. pat_off[s+1])   (literal bytes, no escapes)
 *   gap_min[s], gap_max[s]   RAW gap {a,b} FOLLOWING subpattern s (ignored for a query's last one);
 *                            the library forms the start-to-start bounds a+|P_s|, b+|P_s| exactly
 *                            as vlg_index.hpp:95 does. */
typedef struct {
    uint32_t n_queries;
    const uint32_t* sub_off;  /* n_queries+1 */
    const uint64_t* pat_off;  /* n_subs+1    */
    const uint8_t* pat_bytes;
    const uint64_t* gap_min;  /* n_subs */
    const uint64_t* gap_max;  /* n_subs */
} vlgb_batch;

/* Parse ONE query string into a one-query batch owned by the library (free with vlgb_batch_free).
 * Returns VLGB_OK or VLGB_ERR_LAZY / _MINMAX / _NUMBER (same conditions, same order of checks as the
 * reference parser).  Empty subpatterns parse fine and are fenced at search time. */
int vlgb_parse_query(const char* str, size_t len, int syntax, vlgb_batch** out);
/* Parse n query strings into ONE batch (query i of the batch = strs[i]) with a single call: what the loop over
 * the pattern file in gm_search.cpp:84-104 does one string at a time.  status (nullable, n entries) receives the
 * per-query parse code; a query that fails to parse becomes an empty query (0 subpatterns, no matches).  Without
 * a status array the first parse error fails the whole call. */
int vlgb_parse_batch(const char* const* strs, const size_t* lens, uint32_t n, int syntax, int32_t* status,
                     vlgb_batch** out);
/* Same for a pattern FILE image: one query per line ('\n'-separated, a trailing newline is optional), exactly what
 * gm_search.cpp:84-104 reads with std::getline.  The first parse error fails the call. */
int vlgb_parse_lines(const char* buf, size_t len, int syntax, vlgb_batch** out);
void vlgb_batch_free(vlgb_batch* b);

/* ---- index  (replaces construct(vlg_index&, file, cfg, num_bytes), vlg_index.hpp:375-392; the
 *      benchmark's index_wcsearch3(collection&), index_wcsearch3.hpp:365-369; serialize/load,
 *      vlg_index.hpp:183-198) ---------------------------------------------------------------------- */

/* Copy a text shard to the device.
 *   text[0 .. n)      host bytes of the shard INCLUDING its right halo
 *   global_offset     position of text[0] in the whole text
 *   owned_len         text[0 .. owned_len) is owned by this shard (matches whose FIRST subpattern
 *                     starts there are reported by it); text[owned_len .. n) is halo, read only to
 *                     complete owned starts.  Whole text on one GPU: global_offset=0, owned_len=n.
 *   total_len         length of the whole text (tells the shard whether its halo is truncated)
 * Fails with VLGB_ERR_ZERO_SYMBOL if a 0 byte is present (construct.hpp:35-45). */
int vlgb_index_build(const uint8_t* text, uint64_t n, uint64_t global_offset, uint64_t owned_len,
                     uint64_t total_len, vlgb_index** out);
/* Same, from a raw byte file (the `construct(idx, file, 1)` call of examples/vlg_matching.cpp:34). */
int vlgb_index_build_file(const char* path, vlgb_index** out);
/* Same, the text is already in device memory (device pointer stays owned by the caller and must
 * outlive the index; it must be 16-byte aligned and >= n+64 bytes must be readable). */
int vlgb_index_build_device(const uint8_t* text_device, uint64_t n, uint64_t global_offset, uint64_t owned_len,
                            uint64_t total_len, vlgb_index** out);
/* Integer alphabet (vlg_index<int_alphabet_tag>, vlg_index.hpp:45,133; examples/vlg_matching.cpp:45-61).
 * symbols[0 .. n) are non-zero integers; global_offset / owned_len / total_len count SYMBOLS.  The
 * library stores every symbol as W = vlgb_index_symbol_width() bytes (1, 2, 4 or 8: the smallest width
 * that holds the largest symbol), none of them zero, and runs the same kernels: occurrences are kept
 * at multiples of W only and every reported position / cursor is a SYMBOL index.  Subpatterns of a
 * batch searched on such an index must be encoded with vlgb_encode_symbols(W, ...) (pat_bytes then
 * holds W bytes per symbol); gaps stay in symbols. */
int vlgb_index_build_symbols(const uint64_t* symbols, uint64_t n, uint64_t global_offset, uint64_t owned_len,
                             uint64_t total_len, vlgb_index** out);
uint32_t vlgb_index_symbol_width(const vlgb_index* idx);   /* bytes per symbol: 1 for byte texts */
/* out[0 .. n*width).  VLGB_ERR_TOO_LARGE if a symbol does not fit `width` bytes (such a subpattern
 * cannot occur in the text), VLGB_ERR_ZERO_SYMBOL for a 0. */
int vlgb_encode_symbols(uint32_t width, const uint64_t* symbols, uint64_t n, uint8_t* out);
/* ---- rank path  (replaces the structure the reference index actually stores and queries: the wavelet tree over the
 *      suffix array, vlg_index.hpp:116-138 wt_int<bit_vector_il<>>; its construction, vlg_index.hpp:375-392 ->
 *      construct_sa.hpp:145-171 + wt_int.hpp:180-269 + bit_vector_il.hpp:114-154; and its three query primitives:
 *      forward_search suffix_array_algorithm.hpp:48-115, wt_int::operator[] wt_int.hpp:339-361, the in-order range
 *      walk wt_helper.hpp:726-785 over rank_support_il::rank1 bit_vector_il.hpp:264-278) ------------------------
 * Builds, ON THE DEVICE, the suffix array of the shard (prefix doubling) and the wavelet tree over it in flat
 * 64-byte rank blocks.  With it, a batch whose subpatterns are RARE gets its occurrence lists from SA ranges
 * (cost proportional to the occurrences, like the reference) instead of a pass over the text; frequent
 * subpatterns keep using the scan.  vlgb_set_path() overrides the choice.  Byte texts only (symbol width 1).
 * One-time cost, reported by vlgb_index_rank_info (north_star: construction is reported separately). */
#define VLGB_RANK_KEEP_SA 0u  /* keep the plain suffix array next to the tree (4 B/char more, SA probes cost one read) */
#define VLGB_RANK_COMPACT 1u  /* tree only (what the reference stores): SA probes walk the tree (wt_int::operator[])   */
int vlgb_index_attach_rank(vlgb_index* idx, uint32_t flags);
typedef struct {
    uint32_t present, levels, compact, doubling_rounds;
    uint64_t bytes;              /* device bytes of suffix array + rank blocks */
    double build_ms, sa_ms;      /* whole construction / suffix sort alone, host wall clock around the device work */
} vlgb_rank_info;
int vlgb_index_rank_info(const vlgb_index* idx, vlgb_rank_info* out);
/* Copy SA[first .. first+count) to the host, read THROUGH the wavelet tree on the device (wt_int::operator[],
 * wt_int.hpp:339-361) -- what a caller of vlg_index::wt[i] gets.  Needs vlgb_index_attach_rank. */
int vlgb_index_sa_extract(const vlgb_index* idx, uint64_t first, uint64_t count, uint64_t* out);
/* The same read WITHOUT any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  vlgb_index_sa_extract stages the values in a device buffer it allocates; this variant stages
 * them in a file-scope device array of VLGB_SA_EXTRACT_STATIC_MAX u64 entries (1 MiB; every GPU that loads the
 * library holds it until the process exits) and calls no cudaMalloc / cudaFree / allocating library routine.
 *   - count > VLGB_SA_EXTRACT_STATIC_MAX is rejected with VLGB_ERR_ARG (split the range, or use the call above);
 *   - vlgb_init must have run on the calling thread (it creates the stream; that call allocates, this one does not);
 *     the first call on a device looks the array's device address up (cudaGetSymbolAddress, once per device);
 *   - `out` (count entries) may be host or device memory; the call returns after the copy has completed;
 *   - every call on a GPU shares the ONE array, which keeps the last call's contents: entries [0, count) are written
 *     before they are read, so nothing needs clearing, but calls that overlap on the same GPU (two host threads, each
 *     on its own stream) race on it -- serialise them.
 * Needs vlgb_index_attach_rank. */
#define VLGB_SA_EXTRACT_STATIC_MAX (1ull << 17)
int vlgb_index_sa_extract_static(const vlgb_index* idx, uint64_t first, uint64_t count, uint64_t* out);
/* Pattern generation for the reference's experiment grid (examples/generate-pattern.cpp:1-120, driven by
 * benchmark/gapped-matching/Makefile:90-106): the `count` most frequent substrings of `length` symbols whose symbols
 * all belong to `charset` (0 = alpha-numerical + '_', 1 = 0 plus ' ', 2 = any: the tool's human_readable = 0), in the
 * order the reference prints them, padded by repetition when the text has fewer.  The reference walks a compressed
 * suffix tree; here the runs of equal `length`-prefixes in the suffix array are counted on the device.
 * out: count * length bytes; occ (nullable): count occurrence numbers; n_distinct (nullable): phrases before padding.
 * Needs vlgb_index_attach_rank. */
int vlgb_index_frequent_phrases(const vlgb_index* idx, uint32_t length, uint32_t count, uint32_t charset, uint8_t* out,
                                uint64_t* occ, uint32_t* n_distinct);
/* Which path produces the occurrence lists of the calling thread's searches. */
#define VLGB_PATH_AUTO 0   /* cost model: rank path when the index has one and the batch is small and rare */
#define VLGB_PATH_SCAN 1
#define VLGB_PATH_RANK 2   /* fails with VLGB_ERR_ARG at search time if the index has no rank structures */
int vlgb_set_path(int path);

int vlgb_index_save(const vlgb_index* idx, const char* path);
int vlgb_index_load(const char* path, vlgb_index** out);
uint64_t vlgb_index_size(const vlgb_index* idx);          /* n (shard symbols incl. halo) */
/* Copy text[pos .. pos+len) of the shard back to the host (vlg_index::text, vlg_index.hpp:141).  For a
 * symbol index pos/len count ENCODED bytes (symbol i occupies bytes [i*W, (i+1)*W)). */
int vlgb_index_extract(const vlgb_index* idx, uint64_t pos, uint64_t len, uint8_t* out);
void vlgb_index_free(vlgb_index* idx);

/* ---- search  (replaces vlg_iterator ctor/next/operator++, vlg_index.hpp:227-356; locate/count,
 *      vlg_index.hpp:395-411; index_wcsearch3::search/info, index_wcsearch3.hpp:388-423) ---------- */

#define VLGB_WANT_COUNT  0u  /* counts only            (count(),           vlg_index.hpp:404-411)        */
#define VLGB_WANT_STARTS 1u  /* + match starts         (search().positions, index_wcsearch3.hpp:418-420) */
#define VLGB_WANT_TUPLES 2u  /* + all k positions/match (it[i],             vlg_index.hpp:337-340)        */
#define VLGB_KEEP_ON_DEVICE 16u /* do not copy starts/tuples to the host (counts still are)            */
#define VLGB_ASYNC_COPY 32u     /* return as soon as counts / cursors are known: the starts / tuples travel to the host on a
                                 * second stream while the caller goes on (e.g. issues the next batch: the copy of batch i
                                 * overlaps the search of batch i + 1).  vlgb_result_starts / _tuples / _wait / _free wait for
                                 * the copy; everything else (counts, cursors, vlgb_result_fixup) does not.  Ignored with
                                 * VLGB_KEEP_ON_DEVICE. */

/* Run a whole batch.  entry_cursor (nullable, n_queries entries, whole-text positions): the
 * non-overlap cursor each query enters this shard with (pull_forward state, vlg_index.hpp:253-264);
 * NULL means 0 = "search from the beginning".  Queries that hit a fence get a per-query status and
 * zero matches; the call itself still returns VLGB_OK. */
int vlgb_search_batch(const vlgb_index* idx, const vlgb_batch* batch, const uint64_t* entry_cursor,
                      uint32_t want, vlgb_result** out);

uint32_t vlgb_result_queries(const vlgb_result* r);
int32_t vlgb_result_status(const vlgb_result* r, uint32_t q);      /* VLGB_OK or a fence code        */
uint64_t vlgb_result_count(const vlgb_result* r, uint32_t q);      /* number of matches of query q    */
uint64_t vlgb_result_total(const vlgb_result* r);                  /* sum over the batch              */
const uint64_t* vlgb_result_counts(const vlgb_result* r);          /* n_queries counts                */
const uint64_t* vlgb_result_starts(const vlgb_result* r, uint32_t q);  /* host, ascending; NULL if not wanted */
const uint64_t* vlgb_result_tuples(const vlgb_result* r, uint32_t q);  /* host, k(q) per match            */
/* VLGB_KEEP_ON_DEVICE: the same lists in DEVICE memory (whole-text u64 positions; k(q) per match for tuples), NULL
 * for a query without matches or when not kept.  The memory belongs to the result and stays valid until
 * vlgb_result_free; it was written on vlgb_stream_handle() and the search call returns after that stream has been
 * synchronised, so any stream may read it. */
/* Blocks until the host copies of an VLGB_ASYNC_COPY result have landed (no-op otherwise). */
int vlgb_result_wait(const vlgb_result* r);
const uint64_t* vlgb_result_device_starts(const vlgb_result* r, uint32_t q);
const uint64_t* vlgb_result_device_tuples(const vlgb_result* r, uint32_t q);
/* Non-overlap cursor each query leaves this shard with (= entry cursor of the next shard). */
const uint64_t* vlgb_result_exit_cursors(const vlgb_result* r);
/* Whole-text position of the first reported match of every query (~0 if none). */
const uint64_t* vlgb_result_first_starts(const vlgb_result* r);
/* Sum over subpatterns of the occurrence-list sizes (what index_wcsearch3::info prints,
 * index_wcsearch3.hpp:388-397: the total SA-range size). */
uint64_t vlgb_result_occurrences(const vlgb_result* r, uint32_t q);
/* Multi-shard exactness: re-enter the shard with the true cursors gathered from the left
 * neighbour.  Only queries whose cursor lies beyond their first reported start can change; those are
 * searched again on the device.  Updates counts / starts / tuples / exit cursors in place;
 * *n_changed = number of queries searched again.  The index must still be alive. */
int vlgb_result_fixup(vlgb_result* r, const uint64_t* entry_cursor, uint32_t* n_changed);
void vlgb_result_free(vlgb_result* r);

/* ---- multi-GPU  (the text is cut into contiguous shards that overlap by the largest query span; a
 *      match belongs to the shard that contains its first subpattern's start.  The only coupling between shards is
 *      the non-overlap cursor pull_forward leaves behind, vlg_index.hpp:253-264.) -------------------------------------
 * One host thread drives one GPU = one rank (one process per GPU under torchrun, or one thread per GPU inside one
 * process: gm_search --gpus N).  The communicator is NCCL, bound at run time (libnccl.so.2); it carries ONE
 * all-gather of 24 bytes per query and rank per round.  n_ranks = 1 needs no NCCL and no id. */
#define VLGB_COMM_ID_BYTES 128
int vlgb_comm_unique_id(uint8_t* id /* VLGB_COMM_ID_BYTES, out */);  /* ncclGetUniqueId: call on ONE rank, hand the bytes to all */
int vlgb_comm_init(const uint8_t* id, int n_ranks, int rank);        /* after vlgb_init(device) on the calling thread */
int vlgb_comm_destroy(void);
typedef struct {
    uint32_t n_ranks, rounds;              /* all-gather rounds (1 = nobody had to re-enter) */
    uint32_t reentered_shards;             /* over all ranks and rounds */
    uint32_t reentered_queries;            /* on this rank */
    float exchange_ms;                     /* host wall clock of the exchange on this rank: collectives + re-entry + waiting */
    uint64_t global_matches;               /* over all ranks */
} vlgb_shard_stats;
/* Search the whole batch on this rank's shards.  The shards of the WHOLE text are ordered rank-major: all shards of
 * rank 0 in the order given, then rank 1's, ...; a rank's shards must be consecutive pieces of the text (each built
 * with its global_offset / owned_len / halo).  Inside the rank the shards are searched left to right and shard j
 * enters with the cursors shard j-1 left (exact, no re-entry needed: this is how a text larger than one shard, e.g.
 * 8 GiB on one GPU, is searched); across ranks every rank first searches speculatively from cursor 0, then the
 * ranks all-gather (count, exit cursor, first start) per query, derive every rank's true entry cursor, and the ranks
 * whose reported set can change re-enter (vlgb_result_fixup) -- repeated until no cursor moves (<= n_ranks
 * rounds, typically 1-2).  Collective: every rank must call it with the same batch.
 * results[j] receives shard j's result (its own matches, whole-text positions); global_counts (nullable, n_queries)
 * the per-query totals over ALL ranks. */
int vlgb_search_sharded(const vlgb_index* const* shards, uint32_t n_local, const vlgb_batch* batch, uint32_t want,
                        vlgb_result** results, uint64_t* global_counts, vlgb_shard_stats* st);
/* The host logic of the exchange, exported so that it is testable without a GPU: rows[n_shards][nq][3] = (count,
 * exit cursor, first start) as reported, used[n_shards][nq] = the entry cursor each row was computed with ->
 * entries[n_shards][nq] = true entry cursors implied by the rows, redo[n_shards] = 1 where a shard's reported set
 * can change.  Returns the number of shards to redo (or a negative status). */
int vlgb_chain_plan(const uint64_t* rows, const uint64_t* used, uint32_t n_shards, uint32_t nq, uint64_t* entries,
                    uint8_t* redo);

/* ---- measurement hooks (bench.py / profiles; no effect on results) ------------------------------ */

/* Per-stage device time (CUDA events on the library's stream) and algorithmic byte counts of the
 * last vlgb_search_batch on this thread.  stage names: see vlgb_stage_name(). */
#define VLGB_MAX_STAGES 16
typedef struct {
    uint32_t n_stages;
    float ms[VLGB_MAX_STAGES];          /* device time per stage, summed over launches */
    uint64_t bytes[VLGB_MAX_STAGES];    /* algorithmic bytes per stage   */
    uint32_t launches[VLGB_MAX_STAGES]; /* kernel launches per stage                    */
    uint64_t h2d_bytes, d2h_bytes;
    uint32_t distinct_patterns, scan_passes, segments, query_chunks;
    uint64_t total_occurrences;         /* sum of list sizes of the distinct patterns   */
    uint64_t join_slots;                /* sum over (query, level) of |L_i|             */
    uint32_t list_path;                 /* VLGB_PATH_SCAN or VLGB_PATH_RANK: where the occurrence lists came from */
    uint32_t host_syncs;                /* host<->device synchronisations inside the call */
} vlgb_stats;
int vlgb_last_stats(vlgb_stats* out);
/* The per-stage device times cost two CUDA event records per stage and one stream synchronisation + ~30 event queries
 * per search (~0.05-0.1 ms of host time).  on = 0 switches them off for the calling thread (ms[] then stays 0; bytes,
 * launches and the counters are still filled); default: on. */
int vlgb_set_stage_timing(int on);
const char* vlgb_stage_name(uint32_t stage);
/* The CUDA stream (cudaStream_t) every kernel of the calling thread is launched on, for callers that
 * bracket calls with their own CUDA events.  NULL before the first device call. */
void* vlgb_stream_handle(void);
/* Tunables (0 = keep default): device memory the scan may use for its hit streams (pass 1 records the
 * occurrences so that pass 2 need not read the text again; 1 = never, pass 2 always re-scans), and the device
 * memory budget for join state. */
int vlgb_set_tuning(uint64_t stream_budget_bytes, uint64_t join_budget_bytes);
/* Environment hooks read once per process, measurement only (stderr), no effect on results:
 *   VLGB_TRACE=1        host wall-clock trace of every search call
 *   VLGB_JOIN_DEBUG=1   per-tile-kind device time of the join probe kernel
 *   VLGB_REV_FACTOR=n   the join is driven from the right list when it is n times shorter than the left (default 16, 0 = never) */

#ifdef __cplusplus
}
#endif
#endif /* VLGB200_H */
