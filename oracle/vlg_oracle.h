/* TEST INFRASTRUCTURE ONLY — CPU oracle for the VLG matching hot path.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load
 * this library, and only as the checker or the timed CPU baseline.  Nothing under
 * vlg_matching_b200/ links, imports or calls it; the product path fails loudly when the CUDA
 * library is missing.
 *
 * What it restates: the RESULT SET of the reference matcher
 *   /root/reference/include/sdsl/vlg_index.hpp:227-322   (vlg_iterator: relax/next/pull_forward)
 *   /root/reference/include/sdsl/vlg_index.hpp:54-105    (gapped_pattern_query parser)
 *   /root/reference/benchmark/gapped-matching/include/utils.hpp:19-71 (benchmark parser)
 * as the list-level specification of SURVEY.md Appendix A (occurrence lists -> right-to-left
 * viability -> greedy non-overlapping selection).  The reference walks a wavelet tree over the
 * suffix array; this oracle scans the text.  Same tuples, different mechanics.
 *
 * Parity pin: PINNED.  tests/test_oracle_vs_reference.py diffs this oracle against the unmodified
 * reference (oracle/_ref/ref_vlg, built by oracle/Makefile from /root/reference) on the SURVEY
 * Appendix B vectors and on randomized texts/queries; tests/golden/ carries reference-generated
 * vectors (generator: tests/golden/make_golden.py) for boxes without /root/reference.
 */
#ifndef VLG_ORACLE_H
#define VLG_ORACLE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum {
    VLGO_OK = 0,
    VLGO_ERR_LAZY = -1,       /* "invalid gap description: expected '?' (lazy semantics)"  vlg_index.hpp:97-99 */
    VLGO_ERR_MINMAX = -2,     /* "invalid gap description: min-gap > max-gap"              vlg_index.hpp:92-94 */
    VLGO_ERR_NUMBER = -3,     /* std::stoull threw (vlg_index.hpp:90-91)                                    */
    VLGO_ERR_EMPTY_SUB = -4,  /* fenced: empty subpattern                                      */
    VLGO_ERR_NOMEM = -5,
    VLGO_ERR_ARG = -6
};

/* A batch of parsed queries in CSR form (the same layout the product's C ABI uses).
 *   query q owns subpatterns  s in [sub_off[q], sub_off[q+1])
 *   subpattern s owns bytes   pat_bytes[pat_off[s] .. pat_off[s+1])
 *   gap_min[s], gap_max[s]    RAW gap (a,b) that FOLLOWS subpattern s (unused for a query's last s)
 */
typedef struct {
    uint32_t n_queries;
    const uint32_t* sub_off;   /* n_queries+1 */
    const uint64_t* pat_off;   /* n_subs+1    */
    const uint8_t* pat_bytes;
    const uint64_t* gap_min;   /* n_subs */
    const uint64_t* gap_max;   /* n_subs */
} vlgo_batch;

/* Result of a batch: match m of query q has its k(q) positions at
 * tuples[tuple_off[q] + m*k(q) .. +k(q)); starts[match_off[q]+m] is its first component. */
typedef struct {
    uint64_t* match_off;  /* n_queries+1, prefix sums of match counts */
    uint64_t* tuple_off;  /* n_queries+1 */
    uint64_t* starts;
    uint64_t* tuples;     /* NULL unless want_tuples */
    int32_t* status;      /* per query: VLGO_OK or a fence code */
} vlgo_result;

/* Parse one query string.  syntax 0 = library (".{a,b}?", vlg_index.hpp:54-105),
 * 1 = benchmark (".{a,b}", utils.hpp:40-66).  On success fills malloc'ed arrays (caller frees
 * with vlgo_parsed_free).  Returns VLGO_OK or an error code. */
typedef struct {
    uint32_t k;
    uint64_t* pat_off;   /* k+1 */
    uint8_t* pat_bytes;
    uint64_t* gap_min;   /* k (last unused = 0) */
    uint64_t* gap_max;
} vlgo_parsed;
int vlgo_parse(const char* s, size_t len, int syntax, vlgo_parsed* out);
void vlgo_parsed_free(vlgo_parsed* p);

/* All occurrences (ascending, overlapping included) of pat in text.  Returns count, *out malloc'ed. */
uint64_t vlgo_occurrences(const uint8_t* text, uint64_t n, const uint8_t* pat, uint64_t m, uint64_t** out);

/* Search a batch over the whole text with `threads` worker threads (queries are independent).
 * Library semantics (vlg_index.hpp: per-gap bounds a_i+|P_i| :95, cursor uses |P_k| :262,306,
 * k=1 sibling-skip quirk wt_helper.hpp:779-780).  index_wcsearch3 (benchmark) gives the same set for
 * k=2 or uniform gaps+lengths; its k>=3 single-gap shortcut is NOT restated. */
int vlgo_search_batch(const uint8_t* text, uint64_t n, const vlgo_batch* b,
                      int want_tuples, int threads, vlgo_result* out);
void vlgo_result_free(vlgo_result* r);

#ifdef __cplusplus
}
#endif
#endif
