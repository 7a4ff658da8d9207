"""vlg_matching_b200 — B200-native variable-length-gap matcher behind the reference's matcher API.

Python mirror of the library surface of olydis/vlg_matching (include/sdsl/vlg_index.hpp):
    idx = construct(text_or_path)        construct(vlg_index&, file, num_bytes)  vlg_index.hpp:375-392
    for m in locate(idx, "P1.{a,b}?P2"): locate(idx, query)                      vlg_index.hpp:395-401
        m[0], m[1] ...                   it[i]                                   vlg_index.hpp:337-340
    count(idx, query)                    count(idx, query)                       vlg_index.hpp:404-411
and of the benchmark surface (benchmark/gapped-matching/include/index_wcsearch3.hpp:349-423): IndexB200.
All searching happens in libvlgb200.so (sm_100a kernels); nothing here computes matches.
"""
from __future__ import annotations

import numpy as np

from . import capi
from .capi import (SYNTAX_BENCHMARK, SYNTAX_LIBRARY, WANT_COUNT, WANT_STARTS, WANT_TUPLES, Index, PackedBatch,
                   VlgbError)

__all__ = ["construct", "locate", "count", "GappedPatternQuery", "IndexB200", "Index", "PackedBatch", "VlgbError"]


class GappedPatternQuery:
    """gapped_pattern_query (vlg_index.hpp:43-106): subpatterns + start-to-start gap bounds."""

    def __init__(self, s: bytes | str, syntax: int = SYNTAX_LIBRARY):
        try:
            self.subpatterns, self.raw_gaps = capi.parse_query(s, syntax)
        except VlgbError as e:
            if e.code in (capi.ERR_LAZY, capi.ERR_MINMAX, capi.ERR_NUMBER):
                raise ValueError(capi.lib.vlgb_last_error().decode("latin-1")) from e  # std::runtime_error in the reference
            raise
        # vlg_index.hpp:95 stores (min + |P_i|, max + |P_i|)
        self.gaps = [(a + len(p), b + len(p)) for (a, b), p in zip(self.raw_gaps, self.subpatterns)]


def construct(text, num_bytes: int = 1) -> Index:
    """construct(idx, file, 1) / construct_im(idx, data, 1) for the byte alphabet; num_bytes = 0 with a
    sequence of non-zero integers builds the int-alphabet index (vlg_index<int_alphabet_tag>,
    examples/vlg_matching.cpp:45-61)."""
    if num_bytes not in (0, 1):
        raise ValueError("num_bytes must be 1 (bytes) or 0 (integer symbols)")
    try:
        if num_bytes == 0:
            return Index.build_symbols(np.asarray(text, dtype=np.uint64))
        if isinstance(text, str):
            return Index.build_file(text)
        return Index.build(text)
    except VlgbError as e:
        if e.code == capi.ERR_ZERO_SYMBOL:
            raise ValueError(str(e)) from e  # std::logic_error in the reference (construct.hpp:35-45)
        raise


def _query(q):
    if isinstance(q, GappedPatternQuery):
        return q
    return GappedPatternQuery(q)


def _int_symbols(sub: bytes):
    """Subpattern of the int alphabet: whitespace-separated decimals, read like `while (ss >> symbol)`
    (vlg_index.hpp:63-69): stops at the first token that is not a number."""
    out = []
    for tok in sub.split():
        if not tok.isdigit():
            break
        out.append(int(tok))
    return out


def _batch_for(idx: Index, q: GappedPatternQuery):
    """-> PackedBatch, or None when a subpattern holds a symbol that cannot occur in an int-alphabet index."""
    if idx.symbol_width() == 1 and not getattr(idx, "int_alphabet", False):
        return PackedBatch([(q.subpatterns, q.raw_gaps)])
    subs = []
    for s in q.subpatterns:
        e = idx.encode(_int_symbols(s))
        if e is None:
            return None
        subs.append(e)
    return PackedBatch([(subs, q.raw_gaps)])


def locate(idx: Index, query) -> np.ndarray:
    """-> uint64 array [count, k]: row m holds it[0..k) of the m-th match (ascending start)."""
    q = _query(query)
    b = _batch_for(idx, q)
    if b is None:
        return np.zeros((0, len(q.subpatterns)), dtype=np.uint64)
    r = idx.search(b, WANT_TUPLES)
    st = r.status(0)
    if st != 0:
        raise ValueError(f"query rejected with status {st}")
    return r.tuples(0)


def count(idx: Index, query) -> int:
    q = _query(query)
    b = _batch_for(idx, q)
    if b is None:
        return 0
    r = idx.search(b, WANT_COUNT)
    if r.status(0) != 0:
        raise ValueError(f"query rejected with status {r.status(0)}")
    return int(r.counts()[0])


class IndexB200:
    """The 7-method benchmark index interface (gm_index.cpp:42-56, gm_search.cpp:71-104)."""

    def __init__(self, text=None):
        self.idx = None if text is None else construct(text)

    def name(self) -> str:
        return "VLG-B200"

    def serialize(self, path: str):
        self.idx.save(path)

    def load(self, path: str):
        self.idx = Index.load(path)

    def swap(self, other: "IndexB200"):
        self.idx, other.idx = other.idx, self.idx

    @staticmethod
    def _batch(patterns):
        qs = [GappedPatternQuery(p, SYNTAX_BENCHMARK) if not isinstance(p, GappedPatternQuery) else p for p in patterns]
        return PackedBatch([(q.subpatterns, q.raw_gaps) for q in qs])

    def info(self, pattern) -> str:
        """index_wcsearch3::info (index_wcsearch3.hpp:388-397): total SA-range size of the subpatterns."""
        r = self.idx.search(self._batch([pattern]), WANT_COUNT)
        return str(r.occurrences(0))

    def prepare(self, pattern):
        pass

    def search(self, pattern) -> np.ndarray:
        """-> match starts, ascending (gapped_search_result::positions, utils.hpp:73-80)."""
        return self.search_batch([pattern])[0]

    def search_batch(self, patterns):
        b = self._batch(patterns)
        r = self.idx.search(b, WANT_STARTS)
        return [r.starts(q) for q in range(len(b))]
