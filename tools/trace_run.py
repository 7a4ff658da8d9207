import os, sys
sys.path.insert(0, "/root/repo")
import bench
from vlg_matching_b200 import capi, synth
capi.init(0)
queries = bench.make_queries()
text = bench.shard_text(0, bench.SHARD_BYTES)
idx = capi.Index.build(text)
strs = [synth.to_benchmark_syntax(q) for q in queries]
pb = capi.ParsedBatch(strs, capi.SYNTAX_BENCHMARK)
for i in range(4):
    sys.stderr.write("--- call %d\n" % i)
    r = idx.search(pb, capi.WANT_STARTS); r.free()
