"""GPU: the five BASELINE.json configs at FULL size against the CPU oracle (oracle/vlg_oracle.c, itself pinned to the
reference by tests/golden/*), through the C ABI.  Bit-exact match starts; counts of ALL queries cross-checked between
count-only and start-producing runs; checksums pinned so that bench.py's `config.checksum` / `num_results` can be
compared with a value the oracle confirmed.

  cfg 2: 200 MiB English, 1000 x 2 subpatterns           ALL queries compared with the oracle
  cfg 3: 1 GiB DNA, 1000 x 3 subpatterns, mixed gaps     32 sampled queries vs oracle, all counts consistent
  cfg 4: one 1 GiB shard of the 8 GiB text, 10 000 x 2   200 sampled queries vs oracle, all counts consistent; and
         several shards on ONE GPU chained by vlgb_search_sharded == one index over the same bytes
  cfg 5: 1 GiB DNA, 64 queries, every list > 10^7        8 sampled queries vs oracle, all counts consistent
(cfg 1 runs through the example program: tests/test_host_programs.py.)"""
import os
import sys

import numpy as np
import pytest

from oracle import pyoracle as O

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
THREADS = os.cpu_count() or 4


@pytest.fixture(scope="module")
def V():
    import vlg_matching_b200 as V
    from vlg_matching_b200 import capi
    assert capi.device_count() > 0
    capi.init(0)
    return V


@pytest.fixture(scope="module")
def dna_text():
    from vlg_matching_b200 import synth
    return synth.make_text(3)  # configs 3 and 5 share the 1 GiB DNA text


def device_to_host_u64(ptr: int, n: int) -> np.ndarray:
    """n uint64 at a DEVICE address -> numpy (cuda-python's runtime binding; plumbing for the test only)."""
    from cuda import cudart
    host = np.zeros(n, dtype=np.uint64)
    err, = cudart.cudaMemcpy(host.ctypes.data, ptr, n * 8, cudart.cudaMemcpyKind.cudaMemcpyDeviceToHost)
    assert int(err) == 0, err
    return host


def checksum_of(result, nq):
    cs, cnt = 0, result.counts()
    for q in range(nq):
        if cnt[q]:
            cs = (cs + int(result.starts(q).sum(dtype=np.uint64))) & 0xFFFFFFFFFFFFFFFF
    return cs


def test_config2_full_all_queries_vs_oracle(V):
    """The bench workload itself (bench.make_queries on bench.shard_text(0)): every start of every query equals the
    oracle's; num_results / checksum are the values bench.py prints at N = 1."""
    import bench
    from vlg_matching_b200 import capi
    text = bench.shard_text(0, bench.SHARD_BYTES)
    queries = bench.make_queries()
    idx = V.Index.build(text)
    b = V.PackedBatch(queries)
    r = idx.search(b, capi.WANT_STARTS)
    rc = idx.search(b, capi.WANT_COUNT)
    assert np.array_equal(r.counts(), rc.counts())
    ref = O.search(text, queries, want_tuples=False, threads=THREADS)
    total = cs = 0
    for q, (st, starts, _) in enumerate(ref):
        assert st == 0 and r.status(q) == 0
        assert np.array_equal(r.starts(q), starts), (q, queries[q])
        total += len(starts)
        cs = (cs + int(starts.sum(dtype=np.uint64))) & 0xFFFFFFFFFFFFFFFF
    assert total == r.total() > 7_000_000
    assert checksum_of(r, len(queries)) == cs
    print(f"cfg2 full: num_results {total} checksum {cs}")
    # the device-resident copy of the same result (VLGB_KEEP_ON_DEVICE) is readable until the result is freed
    rd = idx.search(b, capi.WANT_STARTS | capi.KEEP_ON_DEVICE)
    idx.search(V.PackedBatch(queries[:3]), capi.WANT_STARTS).free()  # another search must not clobber it
    for q in (0, 1, 500, 999):
        n = int(rd.counts()[q])
        if n == 0:
            assert rd.device_starts(q) == 0
            continue
        assert np.array_equal(device_to_host_u64(rd.device_starts(q), n), ref[q][1]), q
    idx.free()


def sampled_vs_oracle(V, idx, text, queries, sample, threads_per_call=2, group=2):
    from vlg_matching_b200 import capi
    checked = 0
    for a in range(0, len(sample), group):
        ids = sample[a:a + group]
        qs = [queries[i] for i in ids]
        oc = O.search(text, qs, want_tuples=False, threads=threads_per_call)
        rb = idx.search(V.PackedBatch(qs), capi.WANT_STARTS)
        for k, i in enumerate(ids):
            assert np.array_equal(rb.starts(k), oc[k][1]), i
            checked += len(oc[k][1])
        yield ids, [len(o[1]) for o in oc]
        rb.free()
        del oc


@pytest.mark.parametrize("cfg,nsample", [(3, 32), (5, 8)])
def test_dna_configs_full_size(V, dna_text, cfg, nsample):
    from vlg_matching_b200 import capi, synth
    text = dna_text
    queries = synth.make_queries(cfg, text)
    idx = V.Index.build(text)
    b = V.PackedBatch(queries)
    r = idx.search(b, capi.WANT_COUNT)
    gc = r.counts().copy()
    st = capi.last_stats()
    assert st["total_occurrences"] > (10**9 if cfg == 5 else 10**7)
    r.free()
    rs = idx.search(b, capi.WANT_STARTS | capi.KEEP_ON_DEVICE)
    assert np.array_equal(rs.counts(), gc)
    rs.free()
    step = max(1, len(queries) // nsample)
    sample = list(range(0, len(queries), step))[:nsample]
    # the oracle materialises u64 lists (2 GB per 2.7e8-element list at config 5): few queries at a time
    grp, thr = (2, 2) if cfg == 5 else (8, THREADS)
    for ids, lens in sampled_vs_oracle(V, idx, text, queries, sample, threads_per_call=thr, group=grp):
        for i, n in zip(ids, lens):
            assert int(gc[i]) == n, (cfg, i)
    print(f"cfg{cfg} full: {int(gc.sum())} matches over {len(queries)} queries, {len(sample)} sampled queries equal the oracle")
    idx.free()


def test_config4_shard_and_multi_shard_chain(V):
    """Config 4: shard 0 (1 GiB English, the bench's cfg4 text) with all 10 000 queries -- 200 sampled queries equal the
    oracle, count-only == starts for all; then the same bytes cut into 3 shards on ONE GPU and chained inside
    vlgb_search_sharded (exact cursor hand-over, how 8 GiB runs at N = 1) give the same starts for every query."""
    import bench
    from vlg_matching_b200 import capi, sharded
    text = bench.cfg4_shard_text(0, bench.CFG4_SHARD_BYTES)
    queries = bench.cfg4_queries()
    idx = V.Index.build(text)
    b = V.PackedBatch(queries)
    rc = idx.search(b, capi.WANT_COUNT)
    gc = rc.counts().copy()
    rc.free()
    rs = idx.search(b, capi.WANT_STARTS)
    assert np.array_equal(rs.counts(), gc) and int(gc.sum()) > 10**8
    sample = list(range(0, len(queries), len(queries) // 200))[:200]
    for ids, lens in sampled_vs_oracle(V, idx, text, queries, sample, threads_per_call=THREADS, group=40):
        for i, n in zip(ids, lens):
            assert int(gc[i]) == n, i
    # ---- multi-shard chain on one GPU
    span = sharded.batch_span(queries)
    n = text.size
    cuts = [0, n // 3 + 1, 2 * n // 3, n]
    shards = [V.Index.build(text[cuts[g]:min(n, cuts[g + 1] + span)], global_offset=cuts[g], owned_len=cuts[g + 1] - cuts[g], total_len=n)
              for g in range(3)]
    capi.comm_init(None, 1, 0)
    try:
        res, gcounts, st = capi.search_sharded(shards, b, capi.WANT_STARTS)
    finally:
        capi.comm_destroy()
    assert np.array_equal(gcounts, gc) and st["n_ranks"] == 1
    for q in range(0, len(queries), 7):
        got = np.concatenate([r.starts(q) for r in res])
        assert np.array_equal(got, rs.starts(q)), q
    print(f"cfg4 shard 0: {int(gc.sum())} matches, checksum {checksum_of(rs, len(queries))}; 3-shard chain on one GPU identical")
    for s in shards:
        s.free()
    idx.free()


def test_shard_near_4gib_limit(V):
    """A shard within 64 KiB of the 4 GiB - 4 KiB limit: row/offset arithmetic of the scan must not wrap.
    The text lives on the device only (memset + a few planted patterns), built with vlgb_index_build_device."""
    import torch
    from vlg_matching_b200 import capi
    n = 0xFFFFF000 - 4096
    buf = torch.full((n + 4096,), 0x61, dtype=torch.uint8, device="cuda:0")
    plants = {n - 6: b"needle", n - 40000: b"needle", 3: b"needle", (1 << 31) + 7: b"needle", n - 65000: b"thread"}
    for pos, w in plants.items():
        buf[pos:pos + len(w)] = torch.tensor(list(w), dtype=torch.uint8, device="cuda:0")
    buf[n:] = 0
    torch.cuda.synchronize()
    idx = capi.Index.build_device(buf.data_ptr(), n)
    r = idx.search(V.PackedBatch([([b"needle"], []), ([b"thread", b"needle"], [(0, 30000)]), ([b"needle", b"needle"], [(0, 40000)])]),
                   capi.WANT_TUPLES)
    assert r.starts(0).tolist() == sorted(p for p, w in plants.items() if w == b"needle")
    assert r.tuples(1).tolist() == [[n - 65000, n - 40000]]
    assert r.tuples(2).tolist() == [[n - 40000, n - 6]]
    idx.free()
    del buf
    torch.cuda.empty_cache()


def test_shard_at_the_size_limit(V):
    """A shard of the largest size the library takes (2^32 - 4096 bytes: positions are shard-local u32): occurrences in
    the first bytes, across row boundaries and in the very last bytes are found, nothing wraps in 32-bit arithmetic.
    The text lives on the device only (vlgb_index_build_device); it is one filler byte with planted words, so the
    expected lists are the planted positions."""
    import torch
    from vlg_matching_b200 import capi
    n = 0xFFFFF000
    free, _ = torch.cuda.mem_get_info()
    if free < 60 << 30:
        pytest.skip("needs ~60 GB of free device memory")
    text = torch.full((n + 4096,), ord("x"), dtype=torch.uint8, device="cuda")
    needle, hay = b"needle", b"hay"
    plants = [0, 40, 65279, 65280 * 3 - 3, (1 << 31) - 2, (1 << 32) - 70000, n - 65280 - 3, n - 1030, n - 200, n - 9]
    hays = [p + 6 + 20 for p in plants[:-1]] + [n - 3]  # the last hay ends exactly at the end of the shard
    for p in plants:
        text[p:p + 6] = torch.tensor(list(needle), dtype=torch.uint8, device="cuda")
    for p in hays:
        text[p:p + 3] = torch.tensor(list(hay), dtype=torch.uint8, device="cuda")
    text[n:] = 0  # pad bytes after the shard
    torch.cuda.synchronize()
    idx = capi.Index.build_device(text.data_ptr(), n)
    queries = [([needle], []), ([hay], []), ([needle, hay], [(0, 100)]), ([b"le", b"ha"], [(10, 30)]), ([b"needlex"], [])]
    r = idx.search(V.PackedBatch(queries), capi.WANT_TUPLES)
    assert r.tuples(0)[:, 0].tolist() == plants
    assert r.tuples(1)[:, 0].tolist() == hays
    # every needle but the last has its hay 26 bytes on (gap 20); the last needle (n - 9) is followed by the hay at n - 3: gap 0
    assert r.tuples(2).tolist() == [[p, h] for p, h in zip(plants, hays)]
    assert r.tuples(3).tolist() == [[p + 4, p + 26] for p in plants[:-1]]  # "le" .. "ha": gap 20 inside [10,30]; the last pair has gap 0
    assert r.tuples(4)[:, 0].tolist() == plants[:-1]  # the last needle is followed by the hay, not by the filler
    rc = idx.search(V.PackedBatch(queries), capi.WANT_COUNT)
    assert np.array_equal(rc.counts(), r.counts())
    idx.free()
    del text
    torch.cuda.empty_cache()
