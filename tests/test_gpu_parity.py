"""GPU: the CUDA path, called through the C ABI (libvlgb200.so), against
  * the vectors generated from the unmodified reference (tests/golden/*.json), and
  * the CPU oracle (oracle/vlg_oracle.c) on seeded synthetic inputs.
Integer positions: the bar is bit-exact equality of every tuple."""
import json
import os

import numpy as np
import pytest

from oracle import pyoracle as O

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def V():
    import vlg_matching_b200 as V
    from vlg_matching_b200 import capi
    assert capi.device_count() > 0
    capi.init(0)
    return V


def golden(name):
    with open(os.path.join(HERE, "golden", name)) as f:
        return json.load(f)["vectors"]


def run_vectors(V, vectors):
    from vlg_matching_b200 import capi
    n_checked = 0
    for v in vectors:
        text = v["text"].encode("latin-1")
        idx = V.Index.build(text)
        parsed, cases = [], []
        for c in v["cases"]:
            try:
                subs, gaps = capi.parse_query(c["query"])
            except capi.VlgbError as e:
                assert "error" in c, c
                want = capi.ERR_LAZY if "lazy" in c["error"] else (capi.ERR_MINMAX if "min-gap" in c["error"] else capi.ERR_NUMBER)
                assert e.code == want
                continue
            assert "error" not in c, c
            parsed.append((subs, gaps))
            cases.append(c)
        if not parsed:
            continue
        b = V.PackedBatch(parsed)
        r = idx.search(b, capi.WANT_TUPLES)
        rc = idx.search(b, capi.WANT_COUNT)
        for q, c in enumerate(cases):
            if any(len(s) == 0 for s in parsed[q][0]):
                assert r.status(q) == capi.ERR_EMPTY_SUB  # fenced, SURVEY A.4
                continue
            assert r.status(q) == 0
            assert int(r.counts()[q]) == c["count"] == int(rc.counts()[q]), (v["text"], c["query"])
            assert r.tuples(q).tolist() == c["tuples"], (v["text"], c["query"])
            assert r.starts(q).tolist() == [t[0] for t in c["tuples"]]
            n_checked += 1
        idx.free()
    return n_checked


def test_reference_appendix_b(V):
    assert run_vectors(V, golden("appendix_b.json")) >= 40


def test_reference_random_vectors(V):
    assert run_vectors(V, golden("random_ref.json")) >= 1500


def test_reference_experiment_grid(V):
    """k in {2,4,8,16,32} x the subpattern-length / gap grid of patterns.config, answered by the
    unmodified reference (tests/golden/grid_ref.json, made by make_golden_grid.py)."""
    from test_oracle_golden import check_grid_case, grid_sections
    from vlg_matching_b200 import capi
    n = 0
    for text, cells, cases, hashlib in grid_sections():
        idx = V.Index.build(text)
        qs = [capi.parse_query(c["query"]) for c in cases]
        r = idx.search(V.PackedBatch(qs), capi.WANT_TUPLES)
        for q, c in enumerate(cases):
            assert r.status(q) == 0
            check_grid_case(c, r.tuples(q), hashlib)
            n += 1
        idx.free()
    assert n >= 600


def compare_with_oracle(V, text, queries, want_tuples=True, **build_kw):
    from vlg_matching_b200 import capi
    idx = V.Index.build(text, **build_kw)
    r = idx.search(V.PackedBatch(queries), capi.WANT_TUPLES if want_tuples else capi.WANT_STARTS)
    ref = O.search(text, queries, want_tuples=want_tuples, threads=8)
    total = 0
    for q, (st, starts, tup) in enumerate(ref):
        assert st == 0 and r.status(q) == 0
        assert np.array_equal(r.starts(q), starts), (q, queries[q])
        if want_tuples:
            assert np.array_equal(r.tuples(q), tup), (q, queries[q])
        total += len(starts)
    idx.free()
    return total


@pytest.mark.parametrize("cfg,n,nq", [(1, 1_000_000, 1), (2, 4 << 20, 300), (3, 4 << 20, 200), (5, 1 << 20, 64)])
def test_configs_scaled_vs_oracle(V, cfg, n, nq):
    from vlg_matching_b200 import synth
    text = synth.make_text(cfg, n)
    queries = synth.make_queries(cfg, text, nq)
    assert compare_with_oracle(V, text, queries) > 0


def test_many_patterns_multi_pass(V):
    # > 1024 distinct subpatterns -> several scan passes; long (> 8 byte) subpatterns share prefixes
    from vlg_matching_b200 import synth
    text = synth.make_text(2, 2 << 20)
    queries = synth.sample_queries(text, 900, 2, (2, 14), [(0, 30), (5, 200)], 77, True)
    queries += synth.sample_queries(text, 200, 4, (1, 4), [(0, 10), (3, 60)], 78, False)
    queries += synth.sample_queries(text, 100, 1, (1, 20), [(0, 0)], 79, False)
    assert compare_with_oracle(V, text, queries) > 0


def test_small_alphabet_dense_lists(V):
    rng = np.random.default_rng(5)
    text = rng.integers(97, 99, size=300_000, dtype=np.uint8)  # 'a','b' only: every list is huge
    queries = [([b"a"], []), ([b"b"], []), ([b"ab"], []), ([b"a", b"b"], [(0, 0)]), ([b"a", b"a", b"a"], [(0, 1), (0, 3)]),
               ([b"ab", b"ba", b"ab", b"b"], [(0, 2), (1, 1), (0, 7)]), ([b"aaaaaaaaaaaa"], []),
               ([b"abababab", b"b"], [(5, 9)]), ([b"b", b"a"], [(1000, 2000)])]
    assert compare_with_oracle(V, text, queries) > 0


def test_edge_cases(V):
    from vlg_matching_b200 import capi
    idx = V.Index.build(b"abracadabrasimsalabim")
    b = V.PackedBatch([([b"a"], []), ([b""], []), ([b"a", b""], [(0, 3)]), ([b"zzz", b"a"], [(0, 5)]),
                       ([b"abracadabrasimsalabimX"], []), ([b"a", b"a"], [(2**63, 2**64 - 1)]), ([b"abracadabrasimsalabim"], [])])
    r = idx.search(b, capi.WANT_TUPLES)
    assert r.status(0) == 0 and r.tuples(0)[:, 0].tolist() == [0, 3, 5, 7, 10, 15, 17]
    assert r.status(1) == capi.ERR_EMPTY_SUB and r.status(2) == capi.ERR_EMPTY_SUB
    assert r.status(3) == 0 and r.counts()[3] == 0
    assert r.status(4) == 0 and r.counts()[4] == 0       # longer than the text
    assert r.status(5) == 0 and r.counts()[5] == 0       # saturating gap arithmetic
    assert r.counts()[6] == 1 and r.tuples(6).tolist() == [[0]]
    assert r.occurrences(0) == 7
    # empty batch, one-byte text
    r0 = idx.search(V.PackedBatch([]), capi.WANT_TUPLES)
    assert r0.n == 0 and r0.total() == 0
    one = V.Index.build(b"a")
    r1 = one.search(V.PackedBatch([([b"a"], []), ([b"b"], []), ([b"a", b"a"], [(0, 0)])]), capi.WANT_TUPLES)
    assert r1.counts().tolist() == [1, 0, 0]
    assert idx.extract(4, 7) == b"cadabra" and idx.size() == 21


def test_library_and_benchmark_surfaces(V, tmp_path):
    idx = V.construct(b"abracadabrasimsalabim")
    assert V.count(idx, "ac.{2,5}?a.{4,8}?b") == 1                      # examples/vlg_matching.cpp:40
    assert V.locate(idx, "a.{0,10}?a.{0,10}?a").tolist() == [[0, 3, 5], [7, 10, 15]]
    assert V.count(idx, "foo.{0,10}?bar") == 0
    with pytest.raises(ValueError, match="lazy"):
        V.locate(idx, "a.{1,2}b")
    with pytest.raises(ValueError, match="min-gap > max-gap"):
        V.locate(idx, "a.{3,2}?b")
    with pytest.raises(ValueError, match="zero symbol"):
        V.construct(b"ab\0c")
    bidx = V.IndexB200(b"abracadabrasimsalabim")
    assert bidx.search("a.{0,10}a").tolist() == [0, 5, 10]
    assert bidx.info("a.{0,10}b") == "10"
    p = str(tmp_path / "index.vlgb")
    bidx.serialize(p)
    other = V.IndexB200()
    other.load(p)
    assert other.search("a.{0,10}a").tolist() == [0, 5, 10]


@pytest.mark.parametrize("path", ["device_state", "search_again"])
def test_entry_cursor_and_sharded_equivalence(V, path):
    """Text cut into overlapping shards (owned part + halo >= max span): per-shard results, chained
    through exit cursors with vlgb_result_fixup, equal the single-shard result."""
    from vlg_matching_b200 import capi, synth
    text = synth.make_text(2, 1 << 20)
    queries = synth.make_queries(2, text, 120) + synth.sample_queries(text, 60, 3, (1, 3), [(0, 40), (100, 1000)], 5, False)
    queries += [([b"e"], []), ([b" "], []), ([b"th"], [])]
    span = max(sum(len(s) for s in subs) + sum(g[1] for g in gaps) for subs, gaps in queries)
    whole = V.Index.build(text)
    b = V.PackedBatch(queries)
    rw = whole.search(b, capi.WANT_TUPLES)
    G = 5
    n = text.size
    cuts = [n * g // G for g in range(G + 1)]
    cuts[2] += 1  # odd boundary: the k = 1 parity quirk must use whole-text parity
    cursor = np.zeros(len(queries), dtype=np.uint64)
    parts = [[] for _ in queries]
    changed = 0
    for g in range(G):
        lo, hi = cuts[g], cuts[g + 1]
        end = min(n, hi + span)
        shard = V.Index.build(text[lo:end], global_offset=lo, owned_len=hi - lo, total_len=n)
        r = shard.search(b, capi.WANT_TUPLES)          # speculative: entry cursor 0
        if path == "search_again":                     # another search in between: the kept selection state is gone
            shard.search(V.PackedBatch(queries[:1]), capi.WANT_COUNT).free()
        changed += r.fixup(cursor)                     # true cursors from the left neighbour
        cursor = r.exit_cursors()
        for q in range(len(queries)):
            assert r.status(q) == 0
            parts[q].append(r.tuples(q))
        r2 = shard.search(b, capi.WANT_TUPLES, entry_cursor=np.asarray([c for c in r.exit_cursors()]))
        del r2
    for q in range(len(queries)):
        got = np.concatenate(parts[q], axis=0)
        assert np.array_equal(got, rw.tuples(q)), (q, queries[q])
    assert changed > 0  # the single-byte queries straddle every boundary


def test_repeated_reentry_equals_fresh_search(V):
    """vlgb_result_fixup applied again and again to ONE result (cursors moving right and left) always equals a
    fresh search entered at those cursors: the re-entry works on the selection state kept on the device."""
    from vlg_matching_b200 import capi, synth
    text = synth.make_text(2, 1 << 19)
    queries = synth.make_queries(2, text, 80) + synth.sample_queries(text, 40, 3, (1, 3), [(0, 40), (50, 400)], 7, False)
    queries += [([b"e"], []), ([b"th"], []), ([b"zzzzqq"], []), ([b"e", b"zzzzqq"], [(0, 50)])]
    idx = V.Index.build(text)
    b = V.PackedBatch(queries)
    rng = np.random.default_rng(5)
    cursors = [rng.integers(0, text.size, len(queries)).astype(np.uint64) for _ in range(4)]
    cursors.insert(2, np.zeros(len(queries), dtype=np.uint64))                 # back to the beginning
    cursors.append(np.full(len(queries), text.size + 5, dtype=np.uint64))      # beyond the text: nothing left
    fresh = []
    for e in cursors:
        rf = idx.search(b, capi.WANT_TUPLES, entry_cursor=e)
        fresh.append(([rf.tuples(q).copy() for q in range(len(queries))], rf.exit_cursors().copy(), rf.counts().copy()))
    r = idx.search(b, capi.WANT_TUPLES)
    for e, (tup, ext, cnt) in zip(cursors, fresh):
        r.fixup(e)
        assert np.array_equal(r.counts(), cnt)
        assert np.array_equal(r.exit_cursors(), ext)
        for q in range(len(queries)):
            assert np.array_equal(r.tuples(q), tup[q]), (q, queries[q])


def test_reentry_over_many_tiles_equals_fresh_search(V):
    """Re-entry walks tile by tile (select.cuh, sel_refix_kernel).  Periodic texts make orbits that NEVER coalesce
    (`ab.{0,0}ab` on abab...: entered at 0 it reports 0, 4, 8, ..., entered at 2 it reports 2, 6, 10, ...), so every
    one of the ~1500 tiles of the query changes; a cursor in the middle changes only the tiles behind it; tiles the
    new orbit jumps over are emptied (sparse second pattern); cursors move right, left, beyond the end and back."""
    from vlg_matching_b200 import capi
    n = 3_200_000
    text = np.frombuffer((b"ab" * (n // 2))[:n], dtype=np.uint8).copy()
    text[1_000_001:1_000_400] = ord("c")            # a hole: orbits entered on either side of it differ
    text[2_500_000:2_500_003] = np.frombuffer(b"xyz", dtype=np.uint8)
    queries = [([b"ab", b"ab"], [(0, 0)]), ([b"a"], []), ([b"ab"], []), ([b"aba", b"b"], [(0, 2)]), ([b"b", b"xyz"], [(0, 3_000_000)]),
               ([b"ab", b"a", b"b"], [(1, 5), (0, 8)]), ([b"c", b"ab"], [(0, 600)])]
    idx = V.Index.build(text)
    b = V.PackedBatch(queries)
    nq = len(queries)
    cursors = [np.full(nq, c, dtype=np.uint64) for c in (2, 0, 1_600_001, 3, 999_990, n + 7, 1, 1_000_200, 0)]
    cursors.append(np.arange(nq, dtype=np.uint64) * np.uint64(333_333) + np.uint64(1))
    r = idx.search(b, capi.WANT_STARTS)
    for e in cursors:
        rf = idx.search(b, capi.WANT_STARTS, entry_cursor=e)
        want = ([rf.starts(q).copy() for q in range(nq)], rf.exit_cursors().copy(), rf.counts().copy())
        rf.free()
        r2 = idx.search(b, capi.WANT_STARTS)     # the retained selection state now belongs to r2 ...
        r2.fixup(e)                              # ... which re-enters on the device (tile walk)
        r.fixup(e)                               # and r through the general path
        for res in (r2, r):
            assert np.array_equal(res.counts(), want[2]), e[0]
            assert np.array_equal(res.exit_cursors(), want[1]), e[0]
            for q in range(nq):
                assert np.array_equal(res.starts(q), want[0][q]), (int(e[0]), q)
        # again on the same result, another cursor right away (the state left by the first re-entry is the old orbit now)
        e2 = e + np.uint64(2)
        rf = idx.search(b, capi.WANT_STARTS, entry_cursor=e2)
        r2.fixup(e2)
        assert np.array_equal(r2.counts(), rf.counts())
        for q in range(nq):
            assert np.array_equal(r2.starts(q), rf.starts(q)), (int(e2[0]), q)
        rf.free()
        r2.free()
    assert idx.search(b, capi.WANT_COUNT).counts()[0] > 700_000   # ~1500 selection tiles for the first query
    idx.free()


@pytest.mark.parametrize("cfg", [3, 5])
def test_large_dna_configs_vs_oracle(V, cfg):
    """Configs 3 (three subpatterns, mixed gaps, two scan passes) and 5 (occurrence lists > 10^7, several query
    chunks) on a 256 MiB slice of the DNA text: counts of ALL queries are consistent between count-only and
    start-producing runs, and the match starts of a sample equal the oracle's.  (tools/full_config_check.py runs
    the same at the full 1 GiB.)"""
    from vlg_matching_b200 import capi, synth
    text = synth.make_text(cfg, 256 << 20)
    queries = synth.make_queries(cfg, text)
    idx = V.Index.build(text)
    r = idx.search(V.PackedBatch(queries), capi.WANT_COUNT)
    gc = r.counts().copy()
    st = capi.last_stats()
    assert st["total_occurrences"] > (10**7 if cfg == 5 else 10**6)
    r.free()
    sample = list(range(0, len(queries), len(queries) // 4))[:4]
    qs = [queries[i] for i in sample]
    oc = O.search(text, qs, want_tuples=False, threads=2)
    rb = idx.search(V.PackedBatch(qs), capi.WANT_STARTS)
    for k, i in enumerate(sample):
        assert int(gc[i]) == len(oc[k][1]) == int(rb.counts()[k]), (cfg, i)
        assert np.array_equal(rb.starts(k), oc[k][1]), (cfg, i)


def test_full_size_properties_config2(V):
    """Config 2 at full size (200 MiB, 1000 queries): size-independent properties instead of the oracle:
    every reported tuple really occurs in the text with gaps inside bounds, starts strictly ascending and
    non-overlapping, count-only and tuple runs agree, and a 1 MiB prefix agrees with the oracle."""
    from vlg_matching_b200 import capi, synth
    text = synth.make_text(2)
    queries = synth.make_queries(2, text)
    idx = V.Index.build(text)
    b = V.PackedBatch(queries)
    r = idx.search(b, capi.WANT_TUPLES)
    rc = idx.search(b, capi.WANT_COUNT)
    assert np.array_equal(r.counts(), rc.counts())
    assert r.total() > 0
    for q, (subs, gaps) in enumerate(queries[:200]):
        t = r.tuples(q)
        if t.shape[0] == 0:
            continue
        t = t.astype(np.int64)
        for i, s in enumerate(subs):
            pos = t[:, i]
            for j, ch in enumerate(s):
                assert (text[pos + j] == ch).all()
        for i, (a, bb) in enumerate(gaps):
            d = t[:, i + 1] - t[:, i] - len(subs[i])
            assert (d >= a).all() and (d <= bb).all()
        assert (t[1:, 0] >= t[:-1, -1] + len(subs[-1])).all()
    idx.free()
    small = text[:1 << 20]
    assert compare_with_oracle(V, small, queries[:100]) >= 0


def test_int_alphabet_example_block(V):
    """examples/vlg_matching.cpp:45-61 — the int-alphabet block prints the same tuples as the byte block."""
    sym = [97, 98, 114, 97, 99, 97, 100, 97, 98, 114, 97, 115, 105, 109, 115, 97, 108, 97, 98, 105, 109]
    idx = V.construct(sym, 0)
    assert idx.symbol_width() == 1 and idx.size() == 21
    assert V.locate(idx, "97 99 .{2,5}? 97 .{4,8}? 98").tolist() == [[3, 10, 18]]
    assert V.locate(idx, "97 .{0,10}? 97 .{0,10}? 97").tolist() == [[0, 3, 5], [7, 10, 15]]
    assert V.count(idx, "1337 .{0,10}? 42") == 0          # 1337 cannot occur in a 1-byte-per-symbol index
    with pytest.raises(ValueError, match="zero symbol"):
        V.construct([5, 0, 7], 0)


@pytest.mark.parametrize("scale,width", [(1, 1), (300, 2), (70000, 4), (1 << 40, 8)])
def test_int_alphabet_wide_symbols_vs_oracle(V, scale, width):
    """Symbols of 1/2/4/8 encoded bytes: matching depends on symbol equality only, so the oracle run on
    the byte text (symbol = byte * scale + 7) must give the same tuples."""
    from vlg_matching_b200 import capi, synth
    rng = np.random.default_rng(5 + width)
    text = rng.integers(1, 5, size=60000, dtype=np.uint8)
    queries = synth.sample_queries(text, 40, 3, (1, 3), [(0, 5), (2, 40)], 77, False) + \
        synth.sample_queries(text, 10, 1, (1, 2), [(0, 0)], 78, False)
    sym = text.astype(np.uint64) * np.uint64(scale) + np.uint64(7)
    idx = capi.Index.build_symbols(sym)
    assert idx.symbol_width() == width and idx.size() == text.size
    enc = [([idx.encode(np.frombuffer(s, dtype=np.uint8).astype(np.uint64) * np.uint64(scale) + np.uint64(7)) for s in subs], gaps)
           for subs, gaps in queries]
    r = idx.search(capi.PackedBatch(enc), capi.WANT_TUPLES)
    ref = O.search(text, queries, want_tuples=True, threads=8)
    total = 0
    for q, (st, starts, tup) in enumerate(ref):
        assert r.status(q) == 0
        assert r.tuples(q).tolist() == tup.tolist(), (q, queries[q])
        total += len(starts)
    assert total > 100
    # a subpattern that is not a whole number of symbols is an argument error for that query only
    if width > 1:
        bad = capi.PackedBatch([([b"\x01"], [])])
        assert idx.search(bad, capi.WANT_COUNT).status(0) == capi.ERR_ARG


def test_scan_hit_stream_paths(V):
    """Pass 2 of the scan has three routes: scatter the hit stream recorded by pass 1 (default), re-scan rows
    whose occurrences overflow the stream (> 2 per text position), re-scan everything (no stream budget)."""
    from vlg_matching_b200 import capi, synth
    text = synth.make_text(2, 1 << 19)
    queries = synth.make_queries(2, text, 150)
    dense = np.frombuffer(b"a" * 300000 + b"ab" * 50000, dtype=np.uint8)
    dq = [([b"a"], []), ([b"aa", b"aaa"], [(0, 3)]), ([b"aaaa", b"a", b"ab"], [(1, 2), (0, 9)]), ([b"aaaaa", b"ba"], [(0, 100)])]
    try:
        for budget in (0, 1):
            capi.check(capi.lib.vlgb_set_tuning(budget, 0))
            compare_with_oracle(V, text, queries)
            compare_with_oracle(V, dense, dq)     # 5 occurrences per position: every row overflows the stream
    finally:
        capi.check(capi.lib.vlgb_set_tuning(1 << 62, 0))  # back to "as much as the default rule allows"


def test_async_copy_results_equal_blocking_results(V):
    """VLGB_ASYNC_COPY: the call returns once counts and cursors are known, starts / tuples travel on a second stream.
    Several searches in flight at once (the copy of batch i overlaps the search of batch i + 1) return exactly what the
    blocking calls return; counts are readable at once; re-entry (vlgb_result_fixup) works on a result whose copy is
    still pending; freeing without ever reading is safe."""
    from vlg_matching_b200 import capi, synth
    text = synth.make_text(2, 6 << 20)
    idx = V.Index.build(text)
    batches = [V.PackedBatch(synth.sample_queries(text, 300, k, (3, 7), [(5, 60), (100, 1000)], 500 + k, True)) for k in (2, 3, 2, 4)]
    blocking = [idx.search(b, capi.WANT_TUPLES) for b in batches]
    for want in (capi.WANT_STARTS, capi.WANT_TUPLES):
        pending = [idx.search(b, want | capi.ASYNC_COPY) for b in batches]     # four results in flight
        for r, ref in zip(pending, blocking):
            assert np.array_equal(r.counts(), ref.counts()) and np.array_equal(r.exit_cursors(), ref.exit_cursors())
        for r, ref, b in zip(reversed(pending), reversed(blocking), reversed(batches)):   # read in another order
            for q in range(0, len(b), 7):
                assert np.array_equal(r.starts(q), ref.starts(q))
                if want == capi.WANT_TUPLES:
                    assert np.array_equal(r.tuples(q), ref.tuples(q))
        for r in pending:
            r.wait()
            r.free()
    # never read: the blocks go back to the pools only after the copy has finished
    for _ in range(3):
        idx.search(batches[0], capi.WANT_STARTS | capi.ASYNC_COPY).free()
    again = idx.search(batches[0], capi.WANT_STARTS)
    for q in range(len(batches[0])):
        assert np.array_equal(again.starts(q), blocking[0].starts(q))
    # re-entry while the copy is pending: same as on a blocking result
    ra = idx.search(batches[1], capi.WANT_STARTS | capi.ASYNC_COPY)
    rb = idx.search(batches[1], capi.WANT_STARTS)   # (the retained state now belongs to rb: ra takes the slow path)
    first = rb.first_starts()
    entry = np.where(first != np.uint64(0xFFFFFFFFFFFFFFFF), first + np.uint64(1), np.uint64(0)).astype(np.uint64)
    na, nb = ra.fixup(entry), rb.fixup(entry)
    assert na == nb and nb > 0
    for q in range(len(batches[1])):
        assert np.array_equal(ra.starts(q), rb.starts(q)), q
    assert np.array_equal(ra.counts(), rb.counts())
    idx.free()
