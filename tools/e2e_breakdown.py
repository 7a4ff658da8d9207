#!/usr/bin/env python
"""Where the end-to-end step of bench.py spends its wall time (GPU box only)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
from vlg_matching_b200 import capi, synth
capi.init(0)
queries = bench.make_queries()
text = bench.shard_text(0, bench.SHARD_BYTES)
idx = capi.Index.build(text)
strs = [synth.to_benchmark_syntax(q) for q in queries]
def t(fn, n=10):
    fn(); fn()
    t0 = time.perf_counter()
    for _ in range(n): r = fn()
    return (time.perf_counter() - t0) / n * 1e3
pb = capi.ParsedBatch(strs, capi.SYNTAX_BENCHMARK)
print("parse_batch        %.3f ms" % t(lambda: capi.ParsedBatch(strs, capi.SYNTAX_BENCHMARK)))
def s(want):
    r = idx.search(pb, want); r.free()
print("search COUNT       %.3f ms" % t(lambda: s(capi.WANT_COUNT)))
print("search STARTS dev  %.3f ms" % t(lambda: s(capi.WANT_STARTS | capi.KEEP_ON_DEVICE)))
print("search STARTS host %.3f ms" % t(lambda: s(capi.WANT_STARTS)))
st = capi.last_stats()
print("device stages sum  %.3f ms" % sum(v["ms"] for v in st["stages"].values()), "d2h", st["d2h_bytes"], "h2d", st["h2d_bytes"])
