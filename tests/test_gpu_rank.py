"""GPU: the rank path — suffix array + wavelet tree in flat rank blocks built on the device (sa_build.cu), SA ranges
by warp-cooperative search and in-order enumeration through the tree (wtsa.cuh) — against
  * a plain sorted-suffixes suffix array (what construct_sa gives the reference, construct_sa.hpp:145-171),
  * the scan path of the same library (lists must be identical: same tuples, same occurrence counts),
  * the CPU oracle.
Reference rows: suffix_array_algorithm.hpp:48-115 (SA range), wt_int.hpp:339-361 (SA[i] through the tree),
wt_helper.hpp:726-785 (in-order range walk), bit_vector_il.hpp:264-278 (rank1)."""
import numpy as np
import pytest

from oracle import pyoracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def V():
    import vlg_matching_b200 as V
    from vlg_matching_b200 import capi
    assert capi.device_count() > 0
    capi.init(0)
    yield V
    capi.set_path(capi.PATH_AUTO)


def plain_suffix_array(text: bytes):
    return np.asarray(sorted(range(len(text)), key=lambda i: text[i:]), dtype=np.uint64)


@pytest.mark.parametrize("flags", [0, 1])
def test_suffix_array_through_the_tree(V, flags):
    """vlgb_index_sa_extract reads SA[i] through the wavelet tree (wt_int::operator[]); it must be the suffix array."""
    from vlg_matching_b200 import capi, synth
    rng = np.random.default_rng(3)
    texts = [b"ab", b"ba", b"aaa", b"abracadabrasimsalabim", b"a" * 1000, b"ab" * 777 + b"a", b"mississippi" * 90,
             synth.english_text(20000, 5).tobytes(), synth.dna_text(30011, 6).tobytes(),
             rng.integers(1, 256, size=4097, dtype=np.uint8).tobytes(), (b"abcdefgh" * 64 + b"x") * 7]
    for t in texts:
        idx = V.Index.build(t)
        info = idx.attach_rank(flags)
        assert info["present"] == 1 and info["compact"] == flags and info["levels"] == max(1, (len(t) - 1).bit_length())
        got = idx.sa_extract(0, len(t))
        assert np.array_equal(got, plain_suffix_array(t)), t[:40]
        mid = idx.sa_extract(len(t) // 3, len(t) // 2)
        assert np.array_equal(mid, got[len(t) // 3:len(t) // 3 + len(t) // 2])
        idx.free()


def both_paths(V, idx, queries, want):
    from vlg_matching_b200 import capi
    b = V.PackedBatch(queries)
    capi.set_path(capi.PATH_SCAN)
    rs = idx.search(b, want)
    assert capi.last_stats()["list_path"] == "scan"
    capi.set_path(capi.PATH_RANK)
    rr = idx.search(b, want)
    st = capi.last_stats()
    assert st["list_path"] == "rank" and st["stages"]["scan_count"]["launches"] == 0
    capi.set_path(capi.PATH_AUTO)
    return rs, rr


@pytest.mark.parametrize("flags", [0, 1])
@pytest.mark.parametrize("cfg,n,nq", [(1, 1_000_000, 1), (2, 4 << 20, 300), (3, 2 << 20, 200)])
def test_rank_lists_equal_scan_lists(V, cfg, n, nq, flags):
    """Every distinct subpattern's list from the rank path equals the scan's: same per-query occurrence totals, same
    tuples for every query; and both equal the oracle."""
    from vlg_matching_b200 import capi, synth
    text = synth.make_text(cfg, n)
    queries = synth.make_queries(cfg, text, nq)
    queries += [([b"zzzzqq"], []), ([b"e", b"zzzzqq"], [(0, 50)]), ([text[-5:].tobytes()], []), ([text[:9].tobytes(), b"a"], [(0, 100)])]
    idx = V.Index.build(text)
    idx.attach_rank(flags)
    rs, rr = both_paths(V, idx, queries, capi.WANT_TUPLES)
    ref = O.search(text, queries, want_tuples=True, threads=8)
    total = 0
    for q in range(len(queries)):
        assert rs.status(q) == rr.status(q) == 0
        assert rs.occurrences(q) == rr.occurrences(q), (q, queries[q])
        assert np.array_equal(rs.tuples(q), rr.tuples(q)), (q, queries[q])
        assert np.array_equal(rr.tuples(q), ref[q][2]), (q, queries[q])
        total += rr.tuples(q).shape[0]
    assert total > 0
    idx.free()


def test_rank_path_dense_and_single_symbol(V):
    """Lists that span many rank blocks at every level (small alphabet) and k = 1 queries incl. the sibling-skip quirk."""
    from vlg_matching_b200 import capi
    rng = np.random.default_rng(5)
    text = rng.integers(97, 99, size=300_000, dtype=np.uint8)
    queries = [([b"a"], []), ([b"b"], []), ([b"ab"], []), ([b"a", b"b"], [(0, 0)]), ([b"a", b"a", b"a"], [(0, 1), (0, 3)]),
               ([b"ab", b"ba", b"ab", b"b"], [(0, 2), (1, 1), (0, 7)]), ([b"aaaaaaaaaaaa"], []), ([b"abababab", b"b"], [(5, 9)])]
    idx = V.Index.build(text)
    idx.attach_rank()
    rs, rr = both_paths(V, idx, queries, capi.WANT_TUPLES)
    ref = O.search(text, queries, want_tuples=True, threads=8)
    for q in range(len(queries)):
        assert np.array_equal(rs.tuples(q), rr.tuples(q)) and np.array_equal(rr.tuples(q), ref[q][2]), q
    idx.free()


def test_rank_path_on_shards(V):
    """A shard with a halo: lists hold halo occurrences too, the owned prefix is what level 0 uses; chained shards equal
    the whole text (same protocol as test_entry_cursor_and_sharded_equivalence, lists from SA ranges)."""
    from vlg_matching_b200 import capi, synth
    text = synth.make_text(2, 1 << 20)
    queries = synth.make_queries(2, text, 80) + [([b"e"], []), ([b"th"], [])]
    span = max(sum(len(s) for s in subs) + sum(g[1] for g in gaps) for subs, gaps in queries)
    whole = V.Index.build(text)
    b = V.PackedBatch(queries)
    capi.set_path(capi.PATH_SCAN)
    rw = whole.search(b, capi.WANT_TUPLES)
    capi.set_path(capi.PATH_RANK)
    G, n = 3, text.size
    cuts = [n * g // G for g in range(G + 1)]
    cursor = np.zeros(len(queries), dtype=np.uint64)
    parts = [[] for _ in queries]
    try:
        for g in range(G):
            lo, hi = cuts[g], cuts[g + 1]
            shard = V.Index.build(text[lo:min(n, hi + span)], global_offset=lo, owned_len=hi - lo, total_len=n)
            shard.attach_rank()
            r = shard.search(b, capi.WANT_TUPLES, entry_cursor=cursor)
            assert capi.last_stats()["list_path"] == "rank"
            cursor = r.exit_cursors()
            for q in range(len(queries)):
                parts[q].append(r.tuples(q))
            shard.free()
    finally:
        capi.set_path(capi.PATH_AUTO)
    for q in range(len(queries)):
        assert np.array_equal(np.concatenate(parts[q], axis=0), rw.tuples(q)), (q, queries[q])


def test_cost_model_and_errors(V, tmp_path):
    """AUTO: a few rare subpatterns go through the rank path, a batch of frequent ones through the scan; forcing the
    rank path without rank structures is an argument error; save/load keeps the structures."""
    from vlg_matching_b200 import capi, synth
    text = synth.make_text(2, 8 << 20)
    idx = V.Index.build(text)
    capi.set_path(capi.PATH_RANK)
    with pytest.raises(capi.VlgbError) as e:
        idx.search(V.PackedBatch([([b"the"], [])]), capi.WANT_COUNT)
    assert e.value.code == capi.ERR_ARG
    capi.set_path(capi.PATH_AUTO)
    info = idx.attach_rank()
    assert info["bytes"] > text.size * 4 and info["build_ms"] > 0
    rare = synth.sample_queries(text, 4, 2, (9, 12), [(0, 2000)], 91, True)
    r = idx.search(V.PackedBatch(rare), capi.WANT_STARTS)
    assert capi.last_stats()["list_path"] == "rank"
    ref = O.search(text, rare, want_tuples=False, threads=4)
    for q in range(len(rare)):
        assert np.array_equal(r.starts(q), ref[q][1])
    frequent = [([b"e"], []), ([b" t", b"he"], [(0, 10)])]
    idx.search(V.PackedBatch(frequent), capi.WANT_COUNT)
    assert capi.last_stats()["list_path"] == "scan"
    p = str(tmp_path / "rank.vlgb")
    idx.save(p)
    idx2 = V.Index.load(p)
    assert idx2.rank_info()["present"] == 1
    r2 = idx2.search(V.PackedBatch(rare), capi.WANT_STARTS)
    assert capi.last_stats()["list_path"] == "rank"
    for q in range(len(rare)):
        assert np.array_equal(r2.starts(q), ref[q][1])
    sym = capi.Index.build_symbols(np.asarray([300, 5, 300, 7], dtype=np.uint64))
    with pytest.raises(capi.VlgbError):
        sym.attach_rank()


@pytest.mark.parametrize("flags", [0, 1])
def test_frequent_phrases_equal_reference_rule(V, flags):
    """vlgb_index_frequent_phrases (device: runs of equal prefixes in the suffix array) against the numpy statement of
    examples/generate-pattern.cpp (synth.frequent_phrases: same heap rule, same output order, same padding)."""
    from vlg_matching_b200 import synth
    for text, cases in ((synth.make_text(2, 1 << 20), [(200, 3), (200, 5), (200, 7), (17, 4)]),
                        (synth.make_text(3, 1 << 18), [(200, 2), (200, 3), (50, 6)]),
                        (np.frombuffer(b"ab ab ab, cd cd x_y x_y x_y x_y !!!!", dtype=np.uint8), [(5, 2), (40, 3), (3, 9)])):
        idx = V.Index.build(text)
        idx.attach_rank(flags)
        for count, length in cases:
            got, occ, nd = idx.frequent_phrases(count, length)
            want = synth.frequent_phrases(text, count, length)
            assert got == want, (count, length, got[:5], want[:5])
            tb = text.tobytes()
            for g, o in list(zip(got, occ))[:nd][-5:]:
                c, i = 0, tb.find(g)
                while i >= 0:
                    c, i = c + 1, tb.find(g, i + 1)
                assert c == int(o), (g, c, o)
        idx.free()
