#!/usr/bin/env python
"""Under torchrun (N >= 2): the sharded search of bench.py's workload (N x 200 MiB, halo overlap, cursor exchange,
re-entry) must report exactly the counts a single index over the WHOLE text reports on rank 0's GPU."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
import bench
from vlg_matching_b200 import capi, sharded
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local); capi.init(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
mib = int(os.environ.get("SHARD_MIB", "200"))
n = mib << 20
queries = bench.make_queries(); span = sharded.batch_span(queries)
own = bench.shard_text(rank, n)
host = np.concatenate([own, bench.shard_text(rank + 1, span)]) if rank + 1 < world else own
idx = capi.Index.build(host, global_offset=rank * n, owned_len=n, total_len=n * world)
batch = capi.PackedBatch(queries)
S = sharded.ShardedSearcher(idx, rank, world, dist, dev)
r, gcounts, refixed = S.search(batch, capi.WANT_STARTS)
local_counts = r.counts().copy()
starts0 = [r.starts(q).copy() for q in range(0, len(queries), 97)]
r.free(); idx.free()
dist.barrier()
if rank == 0:
    whole = np.concatenate([bench.shard_text(g, n) for g in range(world)])
    w = capi.Index.build(whole)
    rw = w.search(batch, capi.WANT_STARTS)
    wc = rw.counts().copy()
    assert np.array_equal(wc, gcounts.astype(np.uint64)), np.nonzero(wc != gcounts)[0][:10]
    for k, q in enumerate(range(0, len(queries), 97)):  # rank 0's own matches are the whole-text matches below n
        ws = rw.starts(q)
        assert np.array_equal(ws[ws < n], starts0[k]), q
    print(f"sharded x{world} ({mib} MiB per shard): {int(wc.sum())} matches, per-query counts equal the single-index search; "
          f"{refixed} shard re-entries; rank 0 starts equal", flush=True)
dist.barrier(); dist.destroy_process_group()
