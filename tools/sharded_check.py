#!/usr/bin/env python
"""Under torchrun (N >= 2): vlgb_search_sharded over N ranks x SHARDS_PER_RANK shards of SHARD_MIB MiB each (halo
overlap, NCCL cursor exchange inside the library, re-entry) must report, for EVERY query, exactly the match starts a
single index over the WHOLE text reports inside each shard's owned range -- checked on every rank for its own shards.
Used by tests/test_gpu_multi.py (2 GPUs) and by hand at 4 / 8."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

import bench
from vlg_matching_b200 import capi, sharded

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
capi.init(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
sharded.comm_init(rank, world, dist)
mib = int(os.environ.get("SHARD_MIB", "200"))
per = int(os.environ.get("SHARDS_PER_RANK", "1"))
nq = int(os.environ.get("QUERIES", "1000"))
n = mib << 20
S = world * per
queries = bench.make_queries(nq)
queries += [([b"e"], []), ([b" "], []), ([b"th"], []), ([b"e", b"t"], [(20, 300)])]  # dense lists: every boundary is crossed
span = sharded.batch_span(queries)
pieces = {g: bench.shard_text(g, n) for g in range(S)} if rank == 0 or True else {}
shards = []
for j in range(per):
    g = rank * per + j
    host = np.concatenate([pieces[g], pieces[g + 1][:span]]) if g + 1 < S else pieces[g]
    shards.append(capi.Index.build(host, global_offset=g * n, owned_len=n, total_len=n * S))
batch = capi.PackedBatch(queries)
res, gcounts, st = capi.search_sharded(shards, batch, capi.WANT_STARTS)
# the same with the result copies on the copy stream (VLGB_ASYNC_COPY): two calls in flight, re-entered queries follow
# the first copy; must equal the blocking call
ra1, gca, _ = capi.search_sharded(shards, batch, capi.WANT_STARTS | capi.ASYNC_COPY)
ra2, gcb, _ = capi.search_sharded(shards, batch, capi.WANT_STARTS | capi.ASYNC_COPY)
assert np.array_equal(gca, gcounts) and np.array_equal(gcb, gcounts)
for j in range(per):
    for q in range(len(queries)):
        s0 = res[j].starts(q)
        assert np.array_equal(ra2[j].starts(q), s0) and np.array_equal(ra1[j].starts(q), s0), (rank, j, q)
for r in ra1 + ra2:
    r.free()
whole = np.concatenate([pieces[g] for g in range(S)])
w = capi.Index.build(whole)
rw = w.search(batch, capi.WANT_STARTS)
assert np.array_equal(rw.counts(), gcounts), np.nonzero(rw.counts() != gcounts)[0][:10]
for j in range(per):
    lo, hi = (rank * per + j) * n, (rank * per + j + 1) * n
    for q in range(len(queries)):
        ws = rw.starts(q)
        assert np.array_equal(ws[(ws >= lo) & (ws < hi)], res[j].starts(q)), (rank, j, q)
ok = torch.ones(1, device=dev)
dist.all_reduce(ok)
if rank == 0:
    print(f"sharded x{world} ranks x{per} shards ({mib} MiB each): {int(gcounts.sum())} matches, starts of all {len(queries)} queries equal "
          f"the single-index search on every rank; rounds {st['rounds']}, shards re-entered {st['reentered_shards']}, "
          f"exchange {st['exchange_ms']:.3f} ms", flush=True)
capi.comm_destroy()
dist.barrier()
dist.destroy_process_group()
