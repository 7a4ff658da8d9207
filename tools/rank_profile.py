#!/usr/bin/env python
"""Rank path under ncu: 200 MiB index with the suffix-array wavelet tree attached, a batch of 256 rare two-subpattern
queries forced through VLGB_PATH_RANK (SA-range search + in-order enumeration).  usage: ncu ... python tools/rank_profile.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from vlg_matching_b200 import capi
capi.init(0)
S = bench.synth()
text = bench.shard_text(0, bench.SHARD_BYTES)
idx = capi.Index.build(text)
info = idx.attach_rank()
rare = S.sample_queries(text[:bench.QUERY_SOURCE_BYTES], 256, 2, (9, 12), [(100, 1000)], 21, word_aligned=True)
freq = bench.make_queries(64)  # frequent words through the same path: long SA ranges, the enumeration kernel's regime
capi.set_path(capi.PATH_RANK)
for qs in (rare, freq):
    b = capi.PackedBatch(qs)
    for _ in range(3):
        r = idx.search(b, capi.WANT_STARTS)
        m = r.total(); r.free()
    st = capi.last_stats()
    print(len(qs), "queries", m, "matches", {k: round(v["ms"], 4) for k, v in st["stages"].items() if v["ms"] > 0}, "occurrences", st["total_occurrences"])
print("rank build", info)
