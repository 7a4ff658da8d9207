"""CPU: the C-ABI library loads and exports every symbol include/vlgb200.h declares; host-only entry
points (parser, argument checks) behave like the reference parser.  No compute call is made."""
import os
import re

import pytest

from vlg_matching_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "vlgb200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(vlgb_[a-z_0-9]+)\s*\(", src)))


def test_every_declared_symbol_is_exported():
    decl = declared_symbols()
    assert len(decl) >= 28
    for s in decl:
        assert hasattr(capi.lib, s), s
    assert sorted(capi.SYMBOLS) == decl


def test_parser_matches_reference_rules():
    # vlg_index.hpp:54-105
    subs, gaps = capi.parse_query("ac.{2,5}?a.{4,8}?b")
    assert subs == [b"ac", b"a", b"b"] and gaps == [(2, 5), (4, 8)]
    for q, code in [("a.{1,2}b", capi.ERR_LAZY), ("a.{3,2}?b", capi.ERR_MINMAX), ("a.{x,2}?b", capi.ERR_NUMBER),
                    ("a.{3,2}b", capi.ERR_MINMAX)]:  # min/max is checked before the '?' (vlg_index.hpp:92-99)
        with pytest.raises(capi.VlgbError) as e:
            capi.parse_query(q)
        assert e.value.code == code, q
    # benchmark parser (utils.hpp:40-66): no '?', a '?' belongs to the next subpattern
    subs, gaps = capi.parse_query("the.{10,100}?and", capi.SYNTAX_BENCHMARK)
    assert subs == [b"the", b"?and"] and gaps == [(10, 100)]
    subs, gaps = capi.parse_query("plain")
    assert subs == [b"plain"] and gaps == []


def test_parser_agrees_with_oracle_parser():
    from oracle import pyoracle as O
    import random
    rng = random.Random(3)
    alphabet = "ab.{}?,0123456789 -+x"
    code_map = {O_code: c for O_code, c in [(-1, capi.ERR_LAZY), (-2, capi.ERR_MINMAX), (-3, capi.ERR_NUMBER)]}
    for _ in range(3000):
        q = "".join(rng.choice(alphabet) for _ in range(rng.randrange(0, 18)))
        if rng.random() < 0.5:
            q = "a.{%s,%s}%sb" % (rng.choice(["1", " 2", "x", "", "-1", "+3", "9" * 25]), rng.choice(["5", "0", "", "7x"]),
                                  rng.choice(["?", ""])) + q
        for syntax in (0, 1):
            rc, subs, gaps = O.parse(q, syntax)
            try:
                s2, g2 = capi.parse_query(q, syntax)
                assert rc == 0 and s2 == subs and g2 == gaps, q
            except capi.VlgbError as e:
                assert rc != 0 and code_map[rc] == e.code, (q, rc, e.code)


def test_no_gpu_means_loud_failure():
    if capi.device_count() > 0:
        pytest.skip("GPU present")
    with pytest.raises(capi.VlgbError) as e:
        capi.Index.build(b"abracadabra")
    assert e.value.code == capi.ERR_CUDA


def test_zero_symbol_rejected_before_any_device_work():
    with pytest.raises(capi.VlgbError) as e:
        capi.Index.build(b"abc\0def")
    assert e.value.code == capi.ERR_ZERO_SYMBOL  # construct.hpp:35-45


def test_parse_batch_equals_per_query_parse():
    """vlgb_parse_batch = the per-string parser applied to every string, in one call, same codes."""
    qs = [b"ab.{1,2}cd", b"x", b"a.{0,10}b.{5,7}c", b"a.{3,2}b", b"q.{7,9}"]
    pb = capi.ParsedBatch(qs, capi.SYNTAX_BENCHMARK, strict=False)
    assert len(pb) == len(qs)
    assert pb.status[3] == capi.ERR_MINMAX if hasattr(capi, "ERR_MINMAX") else pb.status[3] == -4
    assert pb.k[3] == 0
    for i, q in enumerate(qs):
        if i == 3:
            continue
        subs, gaps = capi.parse_query(q, capi.SYNTAX_BENCHMARK)
        assert pb.status[i] == 0 and pb.k[i] == len(subs)
    # the arrays agree with the Python-side packer
    good = [q for i, q in enumerate(qs) if i != 3]
    pb2 = capi.ParsedBatch(good, capi.SYNTAX_BENCHMARK)
    ref = capi.PackedBatch([capi.parse_query(q, capi.SYNTAX_BENCHMARK) for q in good])
    import ctypes as C
    import numpy as np
    n_sub = int(ref.sub_off[-1])
    so = np.ctypeslib.as_array(C.cast(pb2.c.sub_off, C.POINTER(C.c_uint32)), shape=(len(good) + 1,))
    po = np.ctypeslib.as_array(C.cast(pb2.c.pat_off, C.POINTER(C.c_uint64)), shape=(n_sub + 1,))
    gmin = np.ctypeslib.as_array(C.cast(pb2.c.gap_min, C.POINTER(C.c_uint64)), shape=(n_sub,))
    gmax = np.ctypeslib.as_array(C.cast(pb2.c.gap_max, C.POINTER(C.c_uint64)), shape=(n_sub,))
    by = np.ctypeslib.as_array(C.cast(pb2.c.pat_bytes, C.POINTER(C.c_uint8)), shape=(int(po[-1]),))
    assert (so == ref.sub_off).all() and (po == ref.pat_off).all() and (by == ref.blob[: len(by)]).all()
    # gaps after a query's last subpattern are ignored by the library; compare the meaningful ones
    for q in range(len(good)):
        for s_ in range(int(so[q]), int(so[q + 1]) - 1):
            assert gmin[s_] == ref.gmin[s_] and gmax[s_] == ref.gmax[s_]
    with pytest.raises(capi.VlgbError):
        capi.ParsedBatch(qs, capi.SYNTAX_BENCHMARK)  # strict: the bad query fails the call
    # a pattern-file image gives the same batch (with and without the trailing newline)
    for image in (b"\n".join(good) + b"\n", b"\n".join(good)):
        pb3 = capi.ParsedBatch(image, capi.SYNTAX_BENCHMARK)
        assert pb3.k == pb2.k
        so3 = np.ctypeslib.as_array(C.cast(pb3.c.sub_off, C.POINTER(C.c_uint32)), shape=(len(good) + 1,))
        po3 = np.ctypeslib.as_array(C.cast(pb3.c.pat_off, C.POINTER(C.c_uint64)), shape=(n_sub + 1,))
        by3 = np.ctypeslib.as_array(C.cast(pb3.c.pat_bytes, C.POINTER(C.c_uint8)), shape=(int(po3[-1]),))
        assert (so3 == so).all() and (po3 == po).all() and (by3 == by).all()


def test_product_never_touches_the_oracle():
    """oracle/ is test infrastructure: nothing under vlg_matching_b200/ (Python, CUDA, C++ host code) may import,
    include, link or execute it, and the Python package must fail loudly without the CUDA library."""
    pkg = os.path.join(ROOT, "vlg_matching_b200")
    offenders = []
    for d, _, files in os.walk(pkg):
        if "__pycache__" in d:
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".hpp", ".cpp", ".h", "Makefile")):
                text = open(os.path.join(d, f), errors="replace").read()
                if re.search(r"\boracle\b|pyoracle|vlgo_|_ref/", text):
                    offenders.append(os.path.relpath(os.path.join(d, f), ROOT))
    assert offenders == [], offenders
    src = open(os.path.join(pkg, "capi.py")).read()
    assert "raise" in src and "libvlgb200.so" in src  # import fails when the extension is missing
