"""TEST INFRASTRUCTURE ONLY — ctypes wrapper around oracle/libvlg_oracle.so (the CPU restatement of
the reference's VLG result set, see oracle/vlg_oracle.h) and around oracle/_ref/ref_vlg (the
unmodified reference matcher built by oracle/Makefile).

Importable only from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs.  The product package (vlg_matching_b200/) never imports this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import tempfile

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libvlg_oracle.so")
REF_BIN = os.path.join(_HERE, "_ref", "ref_vlg")

ERR_NAMES = {0: "OK", -1: "LAZY", -2: "MINMAX", -3: "NUMBER", -4: "EMPTY_SUB", -5: "NOMEM", -6: "ARG"}


class _Parsed(C.Structure):
    _fields_ = [("k", C.c_uint32), ("pat_off", C.POINTER(C.c_uint64)), ("pat_bytes", C.POINTER(C.c_uint8)),
                ("gap_min", C.POINTER(C.c_uint64)), ("gap_max", C.POINTER(C.c_uint64))]


class _Batch(C.Structure):
    _fields_ = [("n_queries", C.c_uint32), ("sub_off", C.c_void_p), ("pat_off", C.c_void_p),
                ("pat_bytes", C.c_void_p), ("gap_min", C.c_void_p), ("gap_max", C.c_void_p)]


class _Result(C.Structure):
    _fields_ = [("match_off", C.POINTER(C.c_uint64)), ("tuple_off", C.POINTER(C.c_uint64)),
                ("starts", C.POINTER(C.c_uint64)), ("tuples", C.POINTER(C.c_uint64)),
                ("status", C.POINTER(C.c_int32))]


_lib = None


def build():
    """Compile the C restatement (and, when /root/reference is mounted, oracle/_ref)."""
    subprocess.run(["make", "-s", "-C", _HERE, "all"], check=True)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            subprocess.run(["make", "-s", "-C", _HERE, "libvlg_oracle.so"], check=True)
        _lib = C.CDLL(LIB_PATH)
        _lib.vlgo_parse.argtypes = [C.c_char_p, C.c_size_t, C.c_int, C.POINTER(_Parsed)]
        _lib.vlgo_parse.restype = C.c_int
        _lib.vlgo_parsed_free.argtypes = [C.POINTER(_Parsed)]
        _lib.vlgo_search_batch.argtypes = [C.c_void_p, C.c_uint64, C.POINTER(_Batch), C.c_int, C.c_int,
                                           C.POINTER(_Result)]
        _lib.vlgo_search_batch.restype = C.c_int
        _lib.vlgo_result_free.argtypes = [C.POINTER(_Result)]
    return _lib


def parse(query: bytes | str, syntax: int = 0):
    """-> (status, [subpattern bytes...], [(a,b)...]) following vlg_index.hpp:54-105 / utils.hpp:40-66."""
    if isinstance(query, str):
        query = query.encode("latin-1")
    p = _Parsed()
    rc = lib().vlgo_parse(query, len(query), syntax, C.byref(p))
    if rc != 0:
        return rc, [], []
    subs, gaps = [], []
    for i in range(p.k):
        a, b = p.pat_off[i], p.pat_off[i + 1]
        subs.append(bytes(bytearray(p.pat_bytes[a:b])))
        if i + 1 < p.k:
            gaps.append((p.gap_min[i], p.gap_max[i]))
    lib().vlgo_parsed_free(C.byref(p))
    return 0, subs, gaps


def pack(queries):
    """queries: list of (subs: list[bytes], gaps: list[(a,b)]) -> numpy CSR arrays."""
    sub_off = [0]
    pat_off = [0]
    blob = bytearray()
    gmin, gmax = [], []
    for subs, gaps in queries:
        for i, s in enumerate(subs):
            blob += s
            pat_off.append(len(blob))
            if i < len(gaps):
                gmin.append(gaps[i][0]); gmax.append(gaps[i][1])
            else:
                gmin.append(0); gmax.append(0)
        sub_off.append(len(pat_off) - 1)
    return (np.asarray(sub_off, dtype=np.uint32), np.asarray(pat_off, dtype=np.uint64),
            np.frombuffer(bytes(blob) or b"\0", dtype=np.uint8).copy(),
            np.asarray(gmin or [0], dtype=np.uint64), np.asarray(gmax or [0], dtype=np.uint64))


def search(text: np.ndarray, queries, want_tuples=True, threads=1):
    """Oracle search.  text: uint8 array; queries: list of (subs, gaps) or query strings (library syntax).
    Returns list per query: (status, starts ndarray, tuples ndarray[count,k] or None)."""
    parsed = []
    pre_status = []
    for q in queries:
        if isinstance(q, (str, bytes)):
            rc, subs, gaps = parse(q, 0)
            pre_status.append(rc)
            parsed.append((subs, gaps) if rc == 0 else ([b"x"], []))
        else:
            pre_status.append(0)
            parsed.append(q)
    text = np.ascontiguousarray(text, dtype=np.uint8)
    sub_off, pat_off, blob, gmin, gmax = pack(parsed)
    b = _Batch(len(parsed), sub_off.ctypes.data, pat_off.ctypes.data, blob.ctypes.data,
               gmin.ctypes.data, gmax.ctypes.data)
    r = _Result()
    rc = lib().vlgo_search_batch(text.ctypes.data, text.size, C.byref(b), int(want_tuples), threads, C.byref(r))
    if rc != 0:
        raise RuntimeError(f"oracle search failed: {rc}")
    out = []
    for qi, (subs, _) in enumerate(parsed):
        k = len(subs)
        m0, m1 = r.match_off[qi], r.match_off[qi + 1]
        starts = np.ctypeslib.as_array(r.starts, shape=(max(r.match_off[len(parsed)], 1),))[m0:m1].copy()
        tup = None
        if want_tuples:
            t0, t1 = r.tuple_off[qi], r.tuple_off[qi + 1]
            tup = np.ctypeslib.as_array(r.tuples, shape=(max(r.tuple_off[len(parsed)], 1),))[t0:t1].copy().reshape(-1, k)
        st = pre_status[qi] if pre_status[qi] != 0 else r.status[qi]
        if pre_status[qi] != 0:
            starts = starts[:0]
            tup = None if tup is None else tup[:0]
        out.append((st, starts, tup))
    lib().vlgo_result_free(C.byref(r))
    return out


def search_timed(text: np.ndarray, queries, threads: int):
    """(seconds, total matches, checksum) for the CPU baseline leg; starts only (gm_search.cpp:110-114)."""
    import time
    parsed = [q if not isinstance(q, (str, bytes)) else parse(q, 0)[1:] for q in queries]
    text = np.ascontiguousarray(text, dtype=np.uint8)
    sub_off, pat_off, blob, gmin, gmax = pack(parsed)
    b = _Batch(len(parsed), sub_off.ctypes.data, pat_off.ctypes.data, blob.ctypes.data,
               gmin.ctypes.data, gmax.ctypes.data)
    r = _Result()
    t0 = time.perf_counter()
    rc = lib().vlgo_search_batch(text.ctypes.data, text.size, C.byref(b), 0, threads, C.byref(r))
    dt = time.perf_counter() - t0
    if rc != 0:
        raise RuntimeError(f"oracle search failed: {rc}")
    M = r.match_off[len(parsed)]
    starts = np.ctypeslib.as_array(r.starts, shape=(max(M, 1),))[:M]
    csum = int(starts.sum(dtype=np.uint64)) if M else 0
    lib().vlgo_result_free(C.byref(r))
    return dt, int(M), csum


# ----------------------------------------------------------------------------- reference binary

def have_reference() -> bool:
    return os.path.exists(REF_BIN) and os.access(REF_BIN, os.X_OK)


def reference_locate(text: bytes, queries: list[str | bytes], timeout=600):
    """Run the UNMODIFIED reference (oracle/_ref/ref_vlg locate).  Returns per query
    ("Q", count, [tuple,...]) or ("E", message)."""
    with tempfile.TemporaryDirectory() as d:
        tf, qf = os.path.join(d, "t.bin"), os.path.join(d, "q.txt")
        with open(tf, "wb") as f:
            f.write(bytes(text))
        with open(qf, "wb") as f:
            for q in queries:
                f.write((q.encode("latin-1") if isinstance(q, str) else q) + b"\n")
        out = subprocess.run([REF_BIN, "locate", tf, qf], check=True, capture_output=True, timeout=timeout).stdout
    res = []
    lines = out.decode("latin-1").split("\n")
    i = 0
    while i < len(lines):
        ln = lines[i]
        if ln.startswith("Q "):
            _, qi, cnt = ln.split()
            cnt = int(cnt)
            tups = [tuple(int(v) for v in lines[i + 1 + j].split()) for j in range(cnt)]
            res.append(("Q", cnt, tups))
            i += 1 + cnt
        elif ln.startswith("E "):
            res.append(("E", ln.split(" ", 2)[2]))
            i += 1
        else:
            i += 1
    return res
