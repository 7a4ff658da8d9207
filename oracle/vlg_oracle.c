/* TEST INFRASTRUCTURE ONLY — see vlg_oracle.h for the scope and the parity pin.
 *
 * CPU restatement of the reference's VLG result set:
 *   A.1  occurrence lists            <- forward_search SA ranges, vlg_index.hpp:310-317
 *   A.2  right-to-left viability     <- relax(), vlg_index.hpp:227-248
 *   A.3  greedy non-overlap select   <- next()/pull_forward(), vlg_index.hpp:253-291
 *   A.4  k=1 sibling skip            <- wt_range_walker::next_leaf(), wt_helper.hpp:777-784
 * Written from the specification; no reference source is copied.
 */
#define _GNU_SOURCE
#include "vlg_oracle.h"

#include <ctype.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------ parsing */

/* std::stoull(str) as the reference calls it (vlg_index.hpp:90-91): leading whitespace, optional
 * sign, decimal digits, trailing garbage ignored; no digits / overflow are exceptions. */
static int stoull_like(const char* s, size_t len, uint64_t* out)
{
    size_t i = 0;
    while (i < len && isspace((unsigned char)s[i])) i++;
    int neg = 0;
    if (i < len && (s[i] == '+' || s[i] == '-')) { neg = s[i] == '-'; i++; }
    if (i >= len || !isdigit((unsigned char)s[i])) return VLGO_ERR_NUMBER;
    uint64_t v = 0;
    for (; i < len && isdigit((unsigned char)s[i]); i++) {
        uint64_t d = (uint64_t)(s[i] - '0');
        if (v > (UINT64_MAX - d) / 10) return VLGO_ERR_NUMBER;
        v = v * 10 + d;
    }
    *out = neg ? (uint64_t)(0 - v) : v;
    return VLGO_OK;
}

static size_t find2(const char* s, size_t len, size_t from, char a, char b)
{
    for (size_t i = from; i + 1 < len; i++)
        if (s[i] == a && s[i + 1] == b) return i;
    return (size_t)-1;
}

static size_t find1(const char* s, size_t len, size_t from, char a)
{
    for (size_t i = from; i < len; i++)
        if (s[i] == a) return i;
    return (size_t)-1;
}

void vlgo_parsed_free(vlgo_parsed* p)
{
    if (!p) return;
    free(p->pat_off); free(p->pat_bytes); free(p->gap_min); free(p->gap_max);
    memset(p, 0, sizeof *p);
}

int vlgo_parse(const char* s, size_t len, int syntax, vlgo_parsed* out)
{
    memset(out, 0, sizeof *out);
    size_t cap = 2;
    for (size_t i = 0; i + 1 < len; i++) if (s[i] == '.' && s[i + 1] == '{') cap++;
    out->pat_off = calloc(cap + 1, sizeof(uint64_t));
    out->pat_bytes = malloc(len + 1);
    out->gap_min = calloc(cap, sizeof(uint64_t));
    out->gap_max = calloc(cap, sizeof(uint64_t));
    if (!out->pat_off || !out->pat_bytes || !out->gap_min || !out->gap_max) {
        vlgo_parsed_free(out);
        return VLGO_ERR_NOMEM;
    }
    uint32_t k = 0;
    uint64_t nbytes = 0;
    size_t start = 0; /* last_gap_end + 1 */
    int rc = VLGO_OK;
    for (;;) {
        size_t gap_pos = find2(s, len, start, '.', '{');
        if (gap_pos == (size_t)-1) break;
        /* subpattern = literal bytes between the previous gap and this one (:78-81) */
        size_t sublen = gap_pos - start;
        memcpy(out->pat_bytes + nbytes, s + start, sublen);
        nbytes += sublen;
        out->pat_off[k + 1] = nbytes;
        /* gap "{a,b}" (:84-91) */
        size_t close = find1(s, len, gap_pos, '}');
        if (close == (size_t)-1) { rc = VLGO_ERR_NUMBER; break; } /* fenced: unterminated gap */
        size_t comma = find1(s, close, gap_pos, ',');
        uint64_t a, b;
        if (comma == (size_t)-1) {
            /* reference: first = everything after ".{", second = whole gap string minus '}' which
             * starts with '.', so stoull throws */
            rc = VLGO_ERR_NUMBER; break;
        }
        if ((rc = stoull_like(s + gap_pos + 2, comma - (gap_pos + 2), &a)) != VLGO_OK) break;
        if ((rc = stoull_like(s + comma + 1, close - (comma + 1), &b)) != VLGO_OK) break;
        if (a > b) { rc = VLGO_ERR_MINMAX; break; }
        out->gap_min[k] = a;
        out->gap_max[k] = b;
        k++;
        if (syntax == 0) {
            /* lazy marker required right after '}' (:97-99); next subpattern starts after it */
            if (close + 1 >= len || s[close + 1] != '?') { rc = VLGO_ERR_LAZY; break; }
            start = close + 2;
        } else {
            start = close + 1; /* utils.hpp:64 */
        }
    }
    if (rc != VLGO_OK) { vlgo_parsed_free(out); return rc; }
    size_t sublen = len - start;
    memcpy(out->pat_bytes + nbytes, s + start, sublen);
    nbytes += sublen;
    out->pat_off[k + 1] = nbytes;
    k++;
    out->k = k;
    return VLGO_OK;
}

/* --------------------------------------------------------------- occurrences */

uint64_t vlgo_occurrences(const uint8_t* text, uint64_t n, const uint8_t* pat, uint64_t m, uint64_t** out)
{
    *out = NULL;
    if (m == 0 || m > n) return 0;
    uint64_t cap = 1024, cnt = 0;
    uint64_t* v = malloc(cap * sizeof *v);
    if (!v) return 0;
    const uint8_t* p = text;
    const uint8_t* last = text + (n - m); /* last admissible start */
    while (p <= last) {
        const uint8_t* c = memchr(p, pat[0], (size_t)(last - p) + 1);
        if (!c) break;
        if (m == 1 || memcmp(c + 1, pat + 1, m - 1) == 0) {
            if (cnt == cap) {
                cap *= 2;
                uint64_t* nv = realloc(v, cap * sizeof *v);
                if (!nv) { free(v); return 0; }
                v = nv;
            }
            v[cnt++] = (uint64_t)(c - text);
        }
        p = c + 1;
    }
    *out = v;
    return cnt;
}

/* --------------------------------------------------------------- one query */

static uint64_t sat_add(uint64_t a, uint64_t b) { return a > UINT64_MAX - b ? UINT64_MAX : a + b; }

typedef struct {
    uint64_t count;
    uint64_t* starts;
    uint64_t* tuples; /* count*k or NULL */
    int32_t status;
} one_result;

static void search_one(const uint8_t* text, uint64_t n, const vlgo_batch* b, uint32_t q, int want_tuples,
                       one_result* r)
{
    memset(r, 0, sizeof *r);
    uint32_t s0 = b->sub_off[q], s1 = b->sub_off[q + 1];
    uint32_t k = s1 - s0;
    if (k == 0) { r->status = VLGO_ERR_ARG; return; }
    for (uint32_t s = s0; s < s1; s++)
        if (b->pat_off[s + 1] == b->pat_off[s]) { r->status = VLGO_ERR_EMPTY_SUB; return; }

    uint64_t** L = calloc(k, sizeof *L);     /* occurrence lists, later compacted to the viable ones */
    uint64_t* len = calloc(k, sizeof *len);
    uint64_t** nxt = calloc(k, sizeof *nxt); /* nxt[i][t] = index into viable list i+1 */
    if (!L || !len || !nxt) { r->status = VLGO_ERR_NOMEM; goto done; }

    /* A.1: lists; any empty list -> no match (vlg_index.hpp:316) */
    for (uint32_t i = 0; i < k; i++) {
        uint32_t s = s0 + i;
        len[i] = vlgo_occurrences(text, n, b->pat_bytes + b->pat_off[s], b->pat_off[s + 1] - b->pat_off[s], &L[i]);
        if (len[i] == 0) goto done;
    }

    /* A.2: right-to-left viability; start-to-start bounds lo = a+|P_i|, hi = b+|P_i| (:95) */
    for (int64_t i = (int64_t)k - 2; i >= 0; i--) {
        uint32_t s = s0 + (uint32_t)i;
        uint64_t plen = b->pat_off[s + 1] - b->pat_off[s];
        uint64_t lo = sat_add(b->gap_min[s], plen), hi = sat_add(b->gap_max[s], plen);
        nxt[i] = malloc((len[i] ? len[i] : 1) * sizeof(uint64_t));
        if (!nxt[i]) { r->status = VLGO_ERR_NOMEM; goto done; }
        uint64_t j = 0, w = 0;
        const uint64_t* V = L[i + 1];
        uint64_t m = len[i + 1];
        for (uint64_t t = 0; t < len[i]; t++) {
            uint64_t x = L[i][t];
            uint64_t need = sat_add(x, lo);
            while (j < m && V[j] < need) j++;
            if (j == m) break; /* no later x can be viable either (need is monotone) */
            if (V[j] <= sat_add(x, hi)) {
                L[i][w] = x;
                nxt[i][w] = j;
                w++;
            }
        }
        len[i] = w;
        if (w == 0) goto done;
    }

    /* A.3: greedy left-to-right selection */
    {
        uint64_t last_len = b->pat_off[s1] - b->pat_off[s1 - 1];
        uint64_t cap = 256, cnt = 0;
        uint64_t* starts = malloc(cap * sizeof *starts);
        uint64_t* tuples = want_tuples ? malloc(cap * k * sizeof *tuples) : NULL;
        if (!starts || (want_tuples && !tuples)) { free(starts); free(tuples); r->status = VLGO_ERR_NOMEM; goto done; }
        uint64_t cursor = 0;
        for (uint64_t t = 0; t < len[0]; t++) {
            uint64_t x = L[0][t];
            if (x < cursor) continue;
            if (cnt == cap) {
                cap *= 2;
                uint64_t* ns = realloc(starts, cap * sizeof *starts);
                uint64_t* nt = want_tuples ? realloc(tuples, cap * k * sizeof *tuples) : NULL;
                if (!ns || (want_tuples && !nt)) { r->status = VLGO_ERR_NOMEM; free(ns ? ns : starts); free(nt ? nt : tuples); goto done; }
                starts = ns; tuples = nt;
            }
            uint64_t pos = x, idx = t;
            starts[cnt] = x;
            if (want_tuples) tuples[cnt * k] = x;
            for (uint32_t i = 0; i + 1 < k; i++) {
                idx = nxt[i][idx];
                pos = L[i + 1][idx];
                if (want_tuples) tuples[cnt * k + i + 1] = pos;
            }
            cnt++;
            cursor = pos + last_len; /* pull_forward, :253-264 */
            /* A.4: k=1, |P|=1: the sibling leaf of an even position is popped unconditionally */
            if (k == 1 && last_len == 1 && (x & 1) == 0) cursor = x + 2;
        }
        r->count = cnt;
        r->starts = starts;
        r->tuples = tuples;
    }
done:
    if (L) for (uint32_t i = 0; i < k; i++) free(L[i]);
    if (nxt) for (uint32_t i = 0; i < k; i++) free(nxt[i]);
    free(L); free(len); free(nxt);
}

/* --------------------------------------------------------------- batch */

typedef struct {
    const uint8_t* text;
    uint64_t n;
    const vlgo_batch* b;
    int want_tuples;
    one_result* res;
    volatile uint32_t* next;
} worker_arg;

static void* worker(void* p)
{
    worker_arg* a = p;
    for (;;) {
        uint32_t q = __atomic_fetch_add(a->next, 1, __ATOMIC_RELAXED);
        if (q >= a->b->n_queries) break;
        search_one(a->text, a->n, a->b, q, a->want_tuples, &a->res[q]);
    }
    return NULL;
}

void vlgo_result_free(vlgo_result* r)
{
    if (!r) return;
    free(r->match_off); free(r->tuple_off); free(r->starts); free(r->tuples); free(r->status);
    memset(r, 0, sizeof *r);
}

int vlgo_search_batch(const uint8_t* text, uint64_t n, const vlgo_batch* b,
                      int want_tuples, int threads, vlgo_result* out)
{
    memset(out, 0, sizeof *out);
    uint32_t Q = b->n_queries;
    one_result* res = calloc(Q ? Q : 1, sizeof *res);
    if (!res) return VLGO_ERR_NOMEM;
    if (threads < 1) threads = 1;
    if ((uint32_t)threads > Q && Q > 0) threads = (int)Q;
    volatile uint32_t next = 0;
    worker_arg arg = { text, n, b, want_tuples, res, &next };
    if (threads == 1) {
        worker(&arg);
    } else {
        pthread_t* th = calloc((size_t)threads, sizeof *th);
        for (int t = 0; t < threads; t++) pthread_create(&th[t], NULL, worker, &arg);
        for (int t = 0; t < threads; t++) pthread_join(th[t], NULL);
        free(th);
    }
    out->match_off = calloc((size_t)Q + 1, sizeof(uint64_t));
    out->tuple_off = calloc((size_t)Q + 1, sizeof(uint64_t));
    out->status = calloc(Q ? Q : 1, sizeof(int32_t));
    for (uint32_t q = 0; q < Q; q++) {
        uint32_t k = b->sub_off[q + 1] - b->sub_off[q];
        out->match_off[q + 1] = out->match_off[q] + res[q].count;
        out->tuple_off[q + 1] = out->tuple_off[q] + res[q].count * k;
        out->status[q] = res[q].status;
    }
    uint64_t M = out->match_off[Q], T = out->tuple_off[Q];
    out->starts = malloc((M ? M : 1) * sizeof(uint64_t));
    out->tuples = want_tuples ? malloc((T ? T : 1) * sizeof(uint64_t)) : NULL;
    for (uint32_t q = 0; q < Q; q++) {
        uint32_t k = b->sub_off[q + 1] - b->sub_off[q];
        if (res[q].count) {
            memcpy(out->starts + out->match_off[q], res[q].starts, res[q].count * sizeof(uint64_t));
            if (want_tuples) memcpy(out->tuples + out->tuple_off[q], res[q].tuples, res[q].count * k * sizeof(uint64_t));
        }
        free(res[q].starts);
        free(res[q].tuples);
    }
    free(res);
    return VLGO_OK;
}
