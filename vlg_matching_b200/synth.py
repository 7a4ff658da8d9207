"""Deterministic synthetic inputs for the five BASELINE.json configs.

Everything derives from splitmix64 used as a counter-based generator (draw i = mix(seed + (i+1)*GAMMA)),
so texts can be produced block by block, in any block size, on any box, and compare equal.  All draws
are `x % m` (the modulo bias is kept on purpose: it is part of the definition of the inputs).
Stands in for the reference's collections and frequent-phrase patterns
(benchmark/gapped-matching/Makefile:90-106, test_case.config:27-53), which are not redistributable.
"""
from __future__ import annotations

import ctypes as C
import hashlib
import os

import numpy as np

GAMMA = np.uint64(0x9E3779B97F4A7C15)
_M1 = np.uint64(0xBF58476D1CE4E5B9)
_M2 = np.uint64(0x94D049BB133111EB)


def splitmix64(seed: int, start: int, count: int) -> np.ndarray:
    """draws start .. start+count-1 of the stream `seed` (uint64)."""
    with np.errstate(over="ignore"):
        i = np.arange(start + 1, start + count + 1, dtype=np.uint64)
        z = np.uint64(seed) + i * GAMMA
        z = (z ^ (z >> np.uint64(30))) * _M1
        z = (z ^ (z >> np.uint64(27))) * _M2
        return z ^ (z >> np.uint64(31))


FUNCTION_WORDS = ("the of and to in a is that for it as was with be by on not he i this are or his from at which but "
                  "have an had they you were their one all we can her has there been if more when will would who so no "
                  "she other its may these what them than some him time into only could").split()
assert len(FUNCTION_WORDS) == 64
_LETTERS = np.frombuffer(b"etaoinshrdlcumwfgypbvkjxqz", dtype=np.uint8)
# cumulative English unigram weights (per mille), same order as _LETTERS
_LETTER_CUM = np.cumsum([127, 91, 82, 75, 70, 67, 63, 61, 60, 43, 40, 28, 28, 24, 24, 22, 20, 20, 19, 15, 10, 8, 2, 2, 1, 1])
_LEN_TABLE = np.asarray([1, 2, 2, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 6, 6, 6, 7, 7, 7, 8, 8, 9, 9, 10, 10, 11, 12], dtype=np.int64)
VOCAB = 50000


_VOCAB_CACHE = {}


def _vocabulary(seed: int):
    """-> (blob uint8, off int64[VOCAB+1]); words 0..63 are real function words, the rest pseudo-words."""
    if seed not in _VOCAB_CACHE:
        _VOCAB_CACHE[seed] = _vocabulary_build(seed)
    return _VOCAB_CACHE[seed]


def _vocabulary_build(seed: int):
    r = splitmix64(seed ^ 0x5EED0001, 0, VOCAB * 13)
    words = []
    seen = set()
    for w in FUNCTION_WORDS:
        words.append(w.encode())
        seen.add(words[-1])
    j = 0
    total = int(_LETTER_CUM[-1])
    while len(words) < VOCAB:
        if j + 13 > r.size:
            r = np.concatenate([r, splitmix64(seed ^ 0x5EED0001, r.size, VOCAB * 13)])
        ln = int(_LEN_TABLE[int(r[j] % np.uint64(len(_LEN_TABLE)))])
        idx = np.searchsorted(_LETTER_CUM, (r[j + 1:j + 1 + ln] % np.uint64(total)).astype(np.int64), side="right")
        j += 13
        w = _LETTERS[idx].tobytes()
        if w in seen:
            continue
        seen.add(w)
        words.append(w)
    off = np.zeros(VOCAB + 1, dtype=np.int64)
    off[1:] = np.cumsum([len(w) for w in words])
    return np.frombuffer(b"".join(words), dtype=np.uint8), off


_GEN = None


def _native():
    """libvlgsynth.so (csrc/host/synth_gen.c): the same generators, threaded C.  None if it was not built; the
    numpy code below is the definition and the fallback (tests compare the two)."""
    global _GEN
    if _GEN is None:
        p = os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib", "libvlgsynth.so")
        _GEN = False
        if os.path.exists(p) and not os.environ.get("VLGB_SYNTH_NUMPY"):
            try:
                g = C.CDLL(p)
                g.vlgs_english.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p,
                                           C.c_uint32, C.c_uint64, C.c_int]
                g.vlgs_dna.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_int]
                _GEN = g
            except OSError:
                _GEN = False
    return _GEN or None


def english_text(n: int, seed: int, block_words: int = 1 << 21) -> np.ndarray:
    """n bytes of Zipf(1.0) word stream over a 50k vocabulary, single spaces, occasional '. , ;';
    bytes in [0x20, 0x7e] only (so create_collection.cpp:87-90 would leave it unchanged)."""
    blob, off = _vocabulary(seed)
    wlen = np.diff(off)
    weights = (np.uint64(1 << 32) // np.arange(1, VOCAB + 1, dtype=np.uint64))
    cum = np.cumsum(weights)
    total = cum[-1]
    g = _native()
    if g is not None and n >= (1 << 16):
        out = np.empty(n, dtype=np.uint8)
        blob_c, off_c, cum_c = np.ascontiguousarray(blob), np.ascontiguousarray(off, dtype=np.int64), np.ascontiguousarray(cum, dtype=np.uint64)
        rc = g.vlgs_english(out.ctypes.data, n, seed & 0xFFFFFFFFFFFFFFFF, blob_c.ctypes.data, off_c.ctypes.data, cum_c.ctypes.data,
                            VOCAB, block_words, min(os.cpu_count() or 1, 64))
        if rc == 0:
            return out
    out = np.empty(n, dtype=np.uint8)
    filled = 0
    w0 = 0
    punct = np.frombuffer(b".,,;", dtype=np.uint8)
    while filled < n:
        r = splitmix64(seed, 2 * w0, 2 * block_words)
        wid = np.searchsorted(cum, r[0::2] % total, side="right").astype(np.int64)
        pr = (r[1::2] % np.uint64(64)).astype(np.int64)
        has_p = pr < 4
        ln = wlen[wid] + has_p + 1  # word [+ punct] + space
        end = np.cumsum(ln)
        start = end - ln
        tot = int(end[-1])
        buf = np.full(tot, 0x20, dtype=np.uint8)
        # gather the word bytes
        wl = wlen[wid]
        src0 = np.repeat(off[wid] - (np.cumsum(wl) - wl), wl)
        src = src0 + np.arange(int(wl.sum()), dtype=np.int64)
        dst = np.repeat(start - (np.cumsum(wl) - wl), wl) + np.arange(int(wl.sum()), dtype=np.int64)
        buf[dst] = blob[src]
        pidx = np.nonzero(has_p)[0]
        buf[start[pidx] + wl[pidx]] = punct[pr[pidx]]
        take = min(tot, n - filled)
        out[filled:filled + take] = buf[:take]
        filled += take
        w0 += block_words
    return out


def dna_text(n: int, seed: int, block: int = 1 << 24) -> np.ndarray:
    """n bytes iid uniform over ACGT."""
    acgt = np.frombuffer(b"ACGT", dtype=np.uint8)
    out = np.empty(n, dtype=np.uint8)
    g = _native()
    if g is not None and n >= (1 << 16):
        g.vlgs_dna(out.ctypes.data, n, seed & 0xFFFFFFFFFFFFFFFF, min(os.cpu_count() or 1, 64))
        return out
    shifts = (np.arange(32, dtype=np.uint64) * np.uint64(2))[None, :]
    filled = 0
    d0 = 0
    while filled < n:
        cnt = min(block, (n - filled + 31) // 32)
        r = splitmix64(seed, d0, cnt)
        sym = ((r[:, None] >> shifts) & np.uint64(3)).astype(np.uint8).reshape(-1)
        take = min(sym.size, n - filled)
        out[filled:filled + take] = acgt[sym[:take]]
        filled += take
        d0 += cnt
    return out


def sha256(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def sample_queries(text: np.ndarray, nq: int, k: int, len_range, gap_choices, seed: int, word_aligned: bool):
    """Subpatterns sampled FROM the text (non-empty lists, mirrors the reference's frequent-phrase patterns).
    -> list of (subs: list[bytes], gaps: list[(a, b)])."""
    n = text.size
    lo_len, hi_len = len_range
    r = splitmix64(seed, 0, nq * k * 3 + nq * k)
    j = 0
    out = []
    for _ in range(nq):
        subs, gaps = [], []
        for i in range(k):
            ln = lo_len + int(r[j] % np.uint64(hi_len - lo_len + 1))
            o = int(r[j + 1] % np.uint64(n - ln - 64))
            if word_aligned:
                w = text[o:o + 64]
                sp = np.nonzero(w == 0x20)[0]
                if sp.size:
                    o = o + int(sp[0]) + 1
            subs.append(text[o:o + ln].tobytes())
            if i + 1 < k:
                gaps.append(gap_choices[int(r[j + 2] % np.uint64(len(gap_choices)))])
            j += 3
        out.append((subs, gaps))
    return out


def to_library_syntax(q) -> bytes:
    subs, gaps = q
    s = subs[0]
    for g, p in zip(gaps, subs[1:]):
        s += b".{%d,%d}?" % g + p
    return s


def to_benchmark_syntax(q) -> bytes:
    subs, gaps = q
    s = subs[0]
    for g, p in zip(gaps, subs[1:]):
        s += b".{%d,%d}" % g + p
    return s


# ---- the five configs of BASELINE.json.  `scale` shrinks text and query count for
# tests; scale=1 is the full configuration.
CONFIGS = {
    1: dict(kind="english", n=1_000_000, seed=1, nq=1, k=2, lens=(3, 7), gaps=[(10, 100)], qseed=11),
    2: dict(kind="english", n=200 << 20, seed=2, nq=1000, k=2, lens=(3, 7), gaps=[(100, 1000)], qseed=12),
    3: dict(kind="dna", n=1 << 30, seed=3, nq=1000, k=3, lens=(5, 8), gaps=[(0, 50), (100, 1000), (1000, 10000)], qseed=13),
    4: dict(kind="english", n=8 << 30, seed=4, nq=10000, k=2, lens=(3, 7), gaps=[(100, 1000)], qseed=14),
    5: dict(kind="dna", n=1 << 30, seed=3, nq=64, k=2, lens=(1, 2), gaps=[(0, 16), (8, 64)], qseed=15),
}


def make_text(cfg: int, n: int | None = None) -> np.ndarray:
    c = CONFIGS[cfg]
    n = c["n"] if n is None else n
    return english_text(n, c["seed"]) if c["kind"] == "english" else dna_text(n, c["seed"])


def make_queries(cfg: int, text: np.ndarray, nq: int | None = None):
    c = CONFIGS[cfg]
    return sample_queries(text, c["nq"] if nq is None else nq, c["k"], c["lens"], c["gaps"], c["qseed"],
                          word_aligned=c["kind"] == "english")


# ---- the reference's experiment grid -------------------------------------------------
# benchmark/gapped-matching/patterns.config:2-16 (subpattern length x gap), num_subpatterns.config:1-5
GRID_LENGTHS = (3, 5, 7)
GRID_GAPS = ((100, 110), (1000, 1100), (10000, 11000), (10, 20), (0, 100))
GRID_K = (2, 4, 8, 16, 32)


def frequent_phrases(text: np.ndarray, count: int, length: int):
    """What examples/generate-pattern.cpp:1-120 prints for `file count length 1 0`: the `count` most frequent
    substrings of `length` alpha-numerical bytes (isalnum or '_').  The reference walks its suffix tree in
    lexicographic order and keeps a min-heap of (occurrences, phrase): a phrase enters while the heap is not full
    or when it occurs MORE often than the heap's minimum (which is then dropped); the heap is finally popped in
    ascending (occurrences, phrase) order, and short results are padded by repetition.  numpy definition (the
    device generator, tools built on the SA index, is tested against this)."""
    import heapq
    n = text.size
    if n < length:
        return [b"x" * length] * count
    ok = np.zeros(256, dtype=bool)
    for c in range(256):
        ch = bytes([c])
        ok[c] = (ch.isalnum() and c < 128) or ch == b"_"
    good = ok[text]
    # windows whose `length` bytes are all in the charset
    run = np.cumsum(~good)
    win_ok = (run[length - 1:] - np.concatenate([[0], run[:n - length]])) == 0
    pos = np.nonzero(win_ok)[0]
    key = np.zeros(pos.size, dtype=np.uint64)
    for j in range(length):
        key = (key << np.uint64(8)) | text[pos + j].astype(np.uint64)
    grams, occ = np.unique(key, return_counts=True)  # ascending key = lexicographic (big-endian packing)
    heap = []
    for g, o in zip(grams.tolist(), occ.tolist()):
        if len(heap) == count and o <= heap[0][0]:
            continue
        heapq.heappush(heap, (o, g))
        if len(heap) > count:
            heapq.heappop(heap)
    res = []
    while heap and len(res) < count:
        o, g = heapq.heappop(heap)
        res.append(int(g).to_bytes(8, "big")[8 - length:])
    if not res:
        res.append(b"x" * length)
    si = len(res)
    while len(res) < count:
        res.append(res[len(res) % si])
    return res


def grid_queries(words, k: int, gap, samples: int):
    """benchmark/gapped-matching/Makefile:98-106: the word list cut into lines of k words (`xargs -n k`), joined
    by the gap, first `samples` lines."""
    out = []
    for i in range(0, len(words), k):
        line = words[i:i + k]
        out.append((list(line), [gap] * (len(line) - 1)))
        if len(out) == samples:
            break
    return out
