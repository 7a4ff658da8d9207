#!/usr/bin/env python
"""Full-size BASELINE configs 3 and 5 on one GPU against the C oracle (counts, exit cursors, checksum of the match
starts).  GPU box only; usage: tools/full_config_check.py [cfg ...] [--mib N]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from vlg_matching_b200 import capi, synth
from oracle import pyoracle as O

args = [a for a in sys.argv[1:] if not a.startswith("--")]
mib = int(sys.argv[sys.argv.index("--mib") + 1]) if "--mib" in sys.argv else 1024
capi.init(0)
for cfg in [int(a) for a in args] or [5, 3]:
    t0 = time.time()
    text = synth.make_text(cfg, mib << 20)
    queries = synth.make_queries(cfg, text)
    print(f"cfg {cfg}: text {text.size >> 20} MiB, {len(queries)} queries (made in {time.time() - t0:.1f} s)", flush=True)
    idx = capi.Index.build(text)
    b = capi.PackedBatch(queries)
    t0 = time.time()
    r = idx.search(b, capi.WANT_STARTS | capi.KEEP_ON_DEVICE)
    t1 = time.time()
    st = capi.last_stats()
    gc, gx = r.counts().copy(), r.exit_cursors().copy()
    print(f"  gpu: {t1 - t0:.3f} s wall (first call), {int(gc.sum())} matches, occurrences {st['total_occurrences']}, "
          f"passes {st['scan_passes']}, chunks {st['query_chunks']}, join slots {st['join_slots']}", flush=True)
    r.free()
    t0 = time.time(); r = idx.search(b, capi.WANT_COUNT); t1 = time.time(); r.free()
    print(f"  gpu: {t1 - t0:.3f} s wall (second call)", flush=True)
    # the oracle materialises u64 lists (2 GB per 2.7e8-element list): a sample of the queries, two at a time
    step = max(1, len(queries) // (8 if cfg == 5 else 24))
    sample = list(range(0, len(queries), step))
    t0 = time.time()
    for a in range(0, len(sample), 2):
        ids = sample[a:a + 2]
        qs = [queries[i] for i in ids]
        oc = O.search(text, qs, want_tuples=False, threads=2)
        rb = idx.search(capi.PackedBatch(qs), capi.WANT_STARTS)
        for k, i in enumerate(ids):
            assert int(gc[i]) == len(oc[k][1]) == int(rb.counts()[k]), (cfg, i, int(gc[i]), len(oc[k][1]))
            assert np.array_equal(rb.starts(k), oc[k][1]), (cfg, i)
        rb.free()
        del oc
    print(f"  cfg {cfg}: counts and match starts of {len(sample)} sampled queries equal the oracle ({time.time() - t0:.0f} s); "
          f"largest sampled count {max(int(gc[i]) for i in sample)}", flush=True)
