"""ctypes binding of libvlgb200.so (include/vlgb200.h).

This is the binding the tests and bench.py use; the reference-facing host layer is the C++ header
vlg_matching_b200/csrc/host/vlg_b200.hpp.  There is no CPU fallback: importing works everywhere (so
the symbol table can be checked on a CPU box), every compute call needs a CUDA device and raises
VlgbError otherwise; a missing library raises at import time.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libvlgb200.so")

OK, ERR_CUDA, ERR_ARG, ERR_LAZY, ERR_MINMAX, ERR_NUMBER, ERR_EMPTY_SUB, ERR_ZERO_SYMBOL, ERR_NOMEM, ERR_IO, \
    ERR_TOO_LARGE = 0, -1, -2, -3, -4, -5, -6, -7, -8, -9, -10
SYNTAX_LIBRARY, SYNTAX_BENCHMARK = 0, 1
WANT_COUNT, WANT_STARTS, WANT_TUPLES, KEEP_ON_DEVICE, ASYNC_COPY = 0, 1, 2, 16, 32
RANK_KEEP_SA, RANK_COMPACT = 0, 1
PATH_AUTO, PATH_SCAN, PATH_RANK = 0, 1, 2
MAX_STAGES = 16

# every symbol include/vlgb200.h declares
SYMBOLS = [
    "vlgb_device_count", "vlgb_init", "vlgb_last_error", "vlgb_version", "vlgb_parse_query", "vlgb_batch_free",
    "vlgb_index_build", "vlgb_index_build_file", "vlgb_index_build_device", "vlgb_index_save", "vlgb_index_load",
    "vlgb_index_size", "vlgb_index_extract", "vlgb_index_free", "vlgb_search_batch", "vlgb_result_queries",
    "vlgb_result_status", "vlgb_result_count", "vlgb_result_total", "vlgb_result_counts", "vlgb_result_starts",
    "vlgb_result_tuples", "vlgb_result_exit_cursors", "vlgb_result_first_starts", "vlgb_result_occurrences",
    "vlgb_result_fixup", "vlgb_result_wait", "vlgb_result_free", "vlgb_last_stats", "vlgb_stage_name", "vlgb_set_tuning", "vlgb_set_stage_timing",
    "vlgb_stream_handle", "vlgb_index_build_symbols", "vlgb_index_symbol_width", "vlgb_encode_symbols",
    "vlgb_parse_batch", "vlgb_parse_lines", "vlgb_index_attach_rank", "vlgb_index_rank_info", "vlgb_index_sa_extract", "vlgb_index_sa_extract_static",
    "vlgb_set_path", "vlgb_result_device_starts", "vlgb_result_device_tuples", "vlgb_comm_unique_id", "vlgb_comm_init",
    "vlgb_comm_destroy", "vlgb_search_sharded", "vlgb_chain_plan", "vlgb_index_frequent_phrases",
]


class VlgbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"vlgb error {code}: {msg}")
        self.code = code


class Batch(C.Structure):
    _fields_ = [("n_queries", C.c_uint32), ("sub_off", C.c_void_p), ("pat_off", C.c_void_p),
                ("pat_bytes", C.c_void_p), ("gap_min", C.c_void_p), ("gap_max", C.c_void_p)]


class Stats(C.Structure):
    _fields_ = [("n_stages", C.c_uint32), ("ms", C.c_float * MAX_STAGES), ("bytes", C.c_uint64 * MAX_STAGES),
                ("launches", C.c_uint32 * MAX_STAGES), ("h2d_bytes", C.c_uint64), ("d2h_bytes", C.c_uint64),
                ("distinct_patterns", C.c_uint32), ("scan_passes", C.c_uint32), ("segments", C.c_uint32),
                ("query_chunks", C.c_uint32), ("total_occurrences", C.c_uint64), ("join_slots", C.c_uint64),
                ("list_path", C.c_uint32), ("host_syncs", C.c_uint32)]


class RankInfo(C.Structure):
    _fields_ = [("present", C.c_uint32), ("levels", C.c_uint32), ("compact", C.c_uint32), ("doubling_rounds", C.c_uint32),
                ("bytes", C.c_uint64), ("build_ms", C.c_double), ("sa_ms", C.c_double)]


if not os.path.exists(LIB_PATH):
    raise ImportError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                      "(there is no CPU fallback)")
lib = C.CDLL(LIB_PATH)

_P = C.c_void_p
lib.vlgb_device_count.restype = C.c_int
lib.vlgb_init.argtypes = [C.c_int]
lib.vlgb_last_error.restype = C.c_char_p
lib.vlgb_version.restype = C.c_char_p
lib.vlgb_parse_query.argtypes = [C.c_char_p, C.c_size_t, C.c_int, C.POINTER(C.POINTER(Batch))]
lib.vlgb_parse_batch.argtypes = [C.POINTER(C.c_char_p), C.POINTER(C.c_size_t), C.c_uint32, C.c_int, C.c_void_p,
                                 C.POINTER(C.POINTER(Batch))]
lib.vlgb_parse_lines.argtypes = [C.c_char_p, C.c_size_t, C.c_int, C.POINTER(C.POINTER(Batch))]
lib.vlgb_batch_free.argtypes = [C.POINTER(Batch)]
lib.vlgb_index_build.argtypes = [_P, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, C.POINTER(_P)]
lib.vlgb_index_build_file.argtypes = [C.c_char_p, C.POINTER(_P)]
lib.vlgb_index_build_device.argtypes = [_P, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, C.POINTER(_P)]
lib.vlgb_index_save.argtypes = [_P, C.c_char_p]
lib.vlgb_index_load.argtypes = [C.c_char_p, C.POINTER(_P)]
lib.vlgb_index_size.argtypes = [_P]
lib.vlgb_index_size.restype = C.c_uint64
lib.vlgb_index_extract.argtypes = [_P, C.c_uint64, C.c_uint64, _P]
lib.vlgb_index_free.argtypes = [_P]
lib.vlgb_search_batch.argtypes = [_P, C.POINTER(Batch), _P, C.c_uint32, C.POINTER(_P)]
lib.vlgb_result_queries.argtypes = [_P]
lib.vlgb_result_queries.restype = C.c_uint32
lib.vlgb_result_status.argtypes = [_P, C.c_uint32]
lib.vlgb_result_status.restype = C.c_int32
lib.vlgb_result_count.argtypes = [_P, C.c_uint32]
lib.vlgb_result_count.restype = C.c_uint64
lib.vlgb_result_total.argtypes = [_P]
lib.vlgb_result_total.restype = C.c_uint64
for _f in ("vlgb_result_counts", "vlgb_result_exit_cursors", "vlgb_result_first_starts"):
    getattr(lib, _f).argtypes = [_P]
    getattr(lib, _f).restype = C.POINTER(C.c_uint64)
for _f in ("vlgb_result_starts", "vlgb_result_tuples"):
    getattr(lib, _f).argtypes = [_P, C.c_uint32]
    getattr(lib, _f).restype = C.POINTER(C.c_uint64)
lib.vlgb_result_occurrences.argtypes = [_P, C.c_uint32]
lib.vlgb_result_occurrences.restype = C.c_uint64
lib.vlgb_result_fixup.argtypes = [_P, _P, C.POINTER(C.c_uint32)]
lib.vlgb_result_wait.argtypes = [_P]
lib.vlgb_result_free.argtypes = [_P]
lib.vlgb_last_stats.argtypes = [C.POINTER(Stats)]
lib.vlgb_stage_name.argtypes = [C.c_uint32]
lib.vlgb_stage_name.restype = C.c_char_p
lib.vlgb_set_tuning.argtypes = [C.c_uint64, C.c_uint64]
lib.vlgb_set_stage_timing.argtypes = [C.c_int]
lib.vlgb_stream_handle.restype = C.c_void_p
lib.vlgb_index_build_symbols.argtypes = [_P, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, C.POINTER(_P)]
lib.vlgb_index_symbol_width.argtypes = [_P]
lib.vlgb_index_symbol_width.restype = C.c_uint32
lib.vlgb_encode_symbols.argtypes = [C.c_uint32, _P, C.c_uint64, _P]
lib.vlgb_index_attach_rank.argtypes = [_P, C.c_uint32]
lib.vlgb_index_rank_info.argtypes = [_P, C.POINTER(RankInfo)]
lib.vlgb_index_sa_extract.argtypes = [_P, C.c_uint64, C.c_uint64, _P]
lib.vlgb_index_sa_extract_static.argtypes = [_P, C.c_uint64, C.c_uint64, _P]
SA_EXTRACT_STATIC_MAX = 1 << 17
lib.vlgb_set_path.argtypes = [C.c_int]
lib.vlgb_index_frequent_phrases.argtypes = [_P, C.c_uint32, C.c_uint32, C.c_uint32, _P, _P, C.POINTER(C.c_uint32)]
for _f in ("vlgb_result_device_starts", "vlgb_result_device_tuples"):
    getattr(lib, _f).argtypes = [_P, C.c_uint32]
    getattr(lib, _f).restype = C.c_void_p
COMM_ID_BYTES = 128


class ShardStats(C.Structure):
    _fields_ = [("n_ranks", C.c_uint32), ("rounds", C.c_uint32), ("reentered_shards", C.c_uint32),
                ("reentered_queries", C.c_uint32), ("exchange_ms", C.c_float), ("global_matches", C.c_uint64)]


lib.vlgb_comm_unique_id.argtypes = [_P]
lib.vlgb_comm_init.argtypes = [_P, C.c_int, C.c_int]
lib.vlgb_search_sharded.argtypes = [C.POINTER(_P), C.c_uint32, C.POINTER(Batch), C.c_uint32, C.POINTER(_P), _P,
                                    C.POINTER(ShardStats)]
lib.vlgb_chain_plan.argtypes = [_P, _P, C.c_uint32, C.c_uint32, _P, _P]


def check(rc):
    if rc != 0:
        raise VlgbError(rc, lib.vlgb_last_error().decode("latin-1"))


def device_count() -> int:
    return lib.vlgb_device_count()


def init(device: int = 0):
    check(lib.vlgb_init(device))


def last_stats() -> dict:
    s = Stats()
    check(lib.vlgb_last_stats(C.byref(s)))
    stages = {}
    for i in range(s.n_stages):
        stages[lib.vlgb_stage_name(i).decode()] = {"ms": float(s.ms[i]), "bytes": int(s.bytes[i]),
                                                   "launches": int(s.launches[i])}
    return {"stages": stages, "h2d_bytes": int(s.h2d_bytes), "d2h_bytes": int(s.d2h_bytes),
            "distinct_patterns": int(s.distinct_patterns), "scan_passes": int(s.scan_passes),
            "segments": int(s.segments), "query_chunks": int(s.query_chunks),
            "total_occurrences": int(s.total_occurrences), "join_slots": int(s.join_slots),
            "list_path": {PATH_SCAN: "scan", PATH_RANK: "rank"}.get(int(s.list_path), "none"), "host_syncs": int(s.host_syncs)}


def comm_unique_id() -> bytes:
    buf = (C.c_uint8 * COMM_ID_BYTES)()
    check(lib.vlgb_comm_unique_id(buf))
    return bytes(buf)


def comm_init(comm_id: bytes | None, n_ranks: int, rank: int):
    """ncclCommInitRank on this thread's device (vlgb_init first).  n_ranks = 1 needs no id."""
    buf = (C.c_uint8 * COMM_ID_BYTES).from_buffer_copy(comm_id) if comm_id else None
    check(lib.vlgb_comm_init(buf, n_ranks, rank))


def comm_destroy():
    lib.vlgb_comm_destroy()


def chain_plan(rows: np.ndarray, used: np.ndarray):
    """vlgb_chain_plan: rows [G, Q, 3], used [G, Q] -> (entries [G, Q], redo bool [G]).  Pure host logic."""
    rows = np.ascontiguousarray(rows, dtype=np.uint64)
    used = np.ascontiguousarray(used, dtype=np.uint64)
    g, q = used.shape
    entries = np.zeros((g, q), dtype=np.uint64)
    redo = np.zeros(max(g, 1), dtype=np.uint8)
    rc = lib.vlgb_chain_plan(rows.ctypes.data, used.ctypes.data, g, q, entries.ctypes.data, redo.ctypes.data)
    if rc < 0:
        check(rc)
    return entries, redo[:g].astype(bool)


def search_sharded(shards, batch, want=WANT_STARTS):
    """vlgb_search_sharded on this rank's shards (consecutive pieces of the text, left to right).
    -> (results [one per local shard], global per-query counts over all ranks, stats dict)."""
    n = len(shards)
    hs = (_P * n)(*[s.h for s in shards])
    rs = (_P * n)()
    gc = np.zeros(max(len(batch), 1), dtype=np.uint64)
    st = ShardStats()
    check(lib.vlgb_search_sharded(hs, n, C.byref(batch.c), want, rs, gc.ctypes.data, C.byref(st)))
    return ([Result(_P(rs[j]), batch, want) for j in range(n)], gc[:len(batch)],
            {k: getattr(st, k) for k, _ in ShardStats._fields_})


def set_path(path: int):
    check(lib.vlgb_set_path(path))


def parse_query(q: bytes | str, syntax: int = SYNTAX_LIBRARY):
    """-> (subpatterns: list[bytes], gaps: list[(a, b)]); raises VlgbError with the reference's codes."""
    if isinstance(q, str):
        q = q.encode("latin-1")
    pb = C.POINTER(Batch)()
    check(lib.vlgb_parse_query(q, len(q), syntax, C.byref(pb)))
    b = pb.contents
    k = C.cast(b.sub_off, C.POINTER(C.c_uint32))[1]
    po = C.cast(b.pat_off, C.POINTER(C.c_uint64))
    by = C.cast(b.pat_bytes, C.POINTER(C.c_uint8))
    gmin = C.cast(b.gap_min, C.POINTER(C.c_uint64))
    gmax = C.cast(b.gap_max, C.POINTER(C.c_uint64))
    subs = [bytes(bytearray(by[po[i]:po[i + 1]])) for i in range(k)]
    gaps = [(int(gmin[i]), int(gmax[i])) for i in range(k - 1)]
    lib.vlgb_batch_free(pb)
    return subs, gaps


class PackedBatch:
    """Queries [(subs, gaps), ...] in the CSR layout of vlgb_batch (numpy arrays kept alive here)."""

    def __init__(self, queries):
        sub_off, pat_off, gmin, gmax = [0], [0], [], []
        blob = bytearray()
        self.k = []
        for subs, gaps in queries:
            self.k.append(len(subs))
            for i, s in enumerate(subs):
                blob += s
                pat_off.append(len(blob))
                a, b = gaps[i] if i < len(gaps) else (0, 0)
                gmin.append(a)
                gmax.append(b)
            sub_off.append(len(pat_off) - 1)
        self.sub_off = np.asarray(sub_off, dtype=np.uint32)
        self.pat_off = np.asarray(pat_off, dtype=np.uint64)
        self.blob = np.frombuffer(bytes(blob) or b"\0", dtype=np.uint8).copy()
        self.gmin = np.asarray(gmin or [0], dtype=np.uint64)
        self.gmax = np.asarray(gmax or [0], dtype=np.uint64)
        self.c = Batch(len(queries), self.sub_off.ctypes.data, self.pat_off.ctypes.data, self.blob.ctypes.data,
                       self.gmin.ctypes.data, self.gmax.ctypes.data)

    def __len__(self):
        return len(self.k)


class ParsedBatch:
    """n query strings parsed by the library in ONE call (vlgb_parse_batch); the batch is owned by the library.
    status[i] != 0 marks a query that failed to parse (it stays in the batch as an empty query)."""

    def __init__(self, strings, syntax: int = SYNTAX_LIBRARY, strict: bool = True):
        if isinstance(strings, (bytes, bytearray)):  # a pattern-file image: one query per line (vlgb_parse_lines)
            self._pb = C.POINTER(Batch)()
            check(lib.vlgb_parse_lines(bytes(strings), len(strings), syntax, C.byref(self._pb)))
            self.c = self._pb.contents
            n = int(self.c.n_queries)
            self.status = np.zeros(max(n, 1), dtype=np.int32)
            so = np.ctypeslib.as_array(C.cast(self.c.sub_off, C.POINTER(C.c_uint32)), shape=(n + 1,))
            self.k = np.diff(so).astype(np.int64).tolist()
            return
        raw = [q.encode("latin-1") if isinstance(q, str) else bytes(q) for q in strings]
        n = len(raw)
        arr = (C.c_char_p * max(n, 1))(*raw)
        lens = (C.c_size_t * max(n, 1))(*[len(r) for r in raw])
        self.status = np.zeros(max(n, 1), dtype=np.int32)
        self._pb = C.POINTER(Batch)()
        check(lib.vlgb_parse_batch(arr, lens, n, syntax, None if strict else self.status.ctypes.data, C.byref(self._pb)))
        self.c = self._pb.contents
        so = np.ctypeslib.as_array(C.cast(self.c.sub_off, C.POINTER(C.c_uint32)), shape=(n + 1,))
        self.k = np.diff(so).astype(np.int64).tolist()

    def __len__(self):
        return len(self.k)

    def __del__(self):
        if getattr(self, "_pb", None):
            lib.vlgb_batch_free(self._pb)
            self._pb = None


class Result:
    def __init__(self, handle, batch: PackedBatch, want: int):
        self.h = handle
        self.batch = batch
        self.want = want

    def __del__(self):
        self.free()

    def free(self):
        if self.h and lib is not None:  # `lib` is gone during interpreter shutdown
            lib.vlgb_result_free(self.h)
        self.h = None

    def wait(self):
        """ASYNC_COPY results: block until the host lists have landed (starts() / tuples() do it implicitly)."""
        check(lib.vlgb_result_wait(self.h))

    @property
    def n(self):
        return lib.vlgb_result_queries(self.h)

    def status(self, q):
        return lib.vlgb_result_status(self.h, q)

    def _u64(self, ptr) -> np.ndarray:
        # copy of n uint64 at a library-owned address (np.ctypeslib.as_array builds a new ctypes type per call: slow)
        if not self.n:
            return np.zeros(0, dtype=np.uint64)
        addr = C.cast(ptr, C.c_void_p).value
        return np.frombuffer((C.c_uint64 * self.n).from_address(addr), dtype=np.uint64).copy()

    def counts(self) -> np.ndarray:
        return self._u64(lib.vlgb_result_counts(self.h))

    def total(self) -> int:
        return lib.vlgb_result_total(self.h)

    def exit_cursors(self) -> np.ndarray:
        return self._u64(lib.vlgb_result_exit_cursors(self.h))

    def first_starts(self) -> np.ndarray:
        return self._u64(lib.vlgb_result_first_starts(self.h))

    def occurrences(self, q) -> int:
        return lib.vlgb_result_occurrences(self.h, q)

    def starts(self, q) -> np.ndarray:
        c = lib.vlgb_result_count(self.h, q)
        p = lib.vlgb_result_starts(self.h, q)
        if not p or c == 0:
            return np.zeros(0, dtype=np.uint64)
        return np.ctypeslib.as_array(p, shape=(c,)).copy()

    def tuples(self, q) -> np.ndarray:
        c = lib.vlgb_result_count(self.h, q)
        k = self.batch.k[q]
        p = lib.vlgb_result_tuples(self.h, q)
        if not p or c == 0:
            return np.zeros((0, k), dtype=np.uint64)
        return np.ctypeslib.as_array(p, shape=(c * k,)).copy().reshape(c, k)

    def device_starts(self, q) -> int:
        """Device address of query q's match starts (u64, KEEP_ON_DEVICE results), 0 if none."""
        return lib.vlgb_result_device_starts(self.h, q) or 0

    def device_tuples(self, q) -> int:
        return lib.vlgb_result_device_tuples(self.h, q) or 0

    def fixup(self, entry_cursor: np.ndarray) -> int:
        e = np.ascontiguousarray(entry_cursor, dtype=np.uint64)
        n = C.c_uint32(0)
        check(lib.vlgb_result_fixup(self.h, e.ctypes.data, C.byref(n)))
        return n.value


class Index:
    """A text shard resident in HBM (vlgb_index)."""

    def __init__(self, handle):
        self.h = handle

    @classmethod
    def build(cls, text, global_offset=0, owned_len=None, total_len=None):
        if isinstance(text, (bytes, bytearray)):
            text = np.frombuffer(bytes(text), dtype=np.uint8)
        text = np.ascontiguousarray(text, dtype=np.uint8)
        n = text.size
        h = _P()
        check(lib.vlgb_index_build(text.ctypes.data if n else None, n, global_offset, n if owned_len is None else owned_len,
                                   (global_offset + n) if total_len is None else total_len, C.byref(h)))
        return cls(h)

    @classmethod
    def build_symbols(cls, symbols, global_offset=0, owned_len=None, total_len=None):
        """Integer alphabet (vlg_index<int_alphabet_tag>): non-zero integers, positions count symbols."""
        sym = np.ascontiguousarray(symbols, dtype=np.uint64)
        n = sym.size
        h = _P()
        check(lib.vlgb_index_build_symbols(sym.ctypes.data if n else None, n, global_offset, n if owned_len is None else owned_len,
                                           (global_offset + n) if total_len is None else total_len, C.byref(h)))
        r = cls(h)
        r.int_alphabet = True
        return r

    def symbol_width(self) -> int:
        return lib.vlgb_index_symbol_width(self.h)

    def encode(self, symbols):
        """-> bytes of a subpattern given as integers, or None if a symbol cannot occur in this index."""
        sym = np.ascontiguousarray(symbols, dtype=np.uint64)
        w = self.symbol_width()
        out = np.zeros(max(sym.size * w, 1), dtype=np.uint8)
        rc = lib.vlgb_encode_symbols(w, sym.ctypes.data if sym.size else None, sym.size, out.ctypes.data)
        if rc in (ERR_TOO_LARGE, ERR_ZERO_SYMBOL):
            return None
        check(rc)
        return out[:sym.size * w].tobytes()

    @classmethod
    def build_device(cls, ptr: int, n: int, global_offset=0, owned_len=None, total_len=None):
        h = _P()
        check(lib.vlgb_index_build_device(ptr, n, global_offset, n if owned_len is None else owned_len,
                                          (global_offset + n) if total_len is None else total_len, C.byref(h)))
        return cls(h)

    @classmethod
    def build_file(cls, path: str):
        h = _P()
        check(lib.vlgb_index_build_file(path.encode(), C.byref(h)))
        return cls(h)

    @classmethod
    def load(cls, path: str):
        h = _P()
        check(lib.vlgb_index_load(path.encode(), C.byref(h)))
        return cls(h)

    def save(self, path: str):
        check(lib.vlgb_index_save(self.h, path.encode()))

    def size(self) -> int:
        return lib.vlgb_index_size(self.h)

    def extract(self, pos: int, length: int) -> bytes:
        out = np.zeros(max(length, 1), dtype=np.uint8)
        check(lib.vlgb_index_extract(self.h, pos, length, out.ctypes.data))
        return out[:length].tobytes()

    def attach_rank(self, flags: int = RANK_KEEP_SA) -> dict:
        """Build suffix array + wavelet tree (flat rank blocks) on the device; -> rank_info()."""
        check(lib.vlgb_index_attach_rank(self.h, flags))
        return self.rank_info()

    def rank_info(self) -> dict:
        r = RankInfo()
        check(lib.vlgb_index_rank_info(self.h, C.byref(r)))
        return {k: getattr(r, k) for k, _ in RankInfo._fields_}

    def frequent_phrases(self, count: int, length: int, charset: int = 0):
        """generate-pattern.cpp: -> (list of `count` phrases (bytes), occurrences uint64[count], distinct found)."""
        out = np.zeros(max(count * length, 1), dtype=np.uint8)
        occ = np.zeros(max(count, 1), dtype=np.uint64)
        nd = C.c_uint32(0)
        check(lib.vlgb_index_frequent_phrases(self.h, length, count, charset, out.ctypes.data, occ.ctypes.data, C.byref(nd)))
        return [out[i * length:(i + 1) * length].tobytes() for i in range(count)], occ[:count], nd.value

    def sa_extract(self, first: int, count: int) -> np.ndarray:
        out = np.zeros(max(count, 1), dtype=np.uint64)
        check(lib.vlgb_index_sa_extract(self.h, first, count, out.ctypes.data))
        return out[:count]

    def search(self, batch: PackedBatch, want=WANT_TUPLES, entry_cursor=None) -> Result:
        e = None
        if entry_cursor is not None:
            e = np.ascontiguousarray(entry_cursor, dtype=np.uint64)
        r = _P()
        check(lib.vlgb_search_batch(self.h, C.byref(batch.c), e.ctypes.data if e is not None else None, want, C.byref(r)))
        return Result(r, batch, want)

    def free(self):
        if self.h and lib is not None:
            lib.vlgb_index_free(self.h)
        self.h = None

    def __del__(self):
        self.free()
