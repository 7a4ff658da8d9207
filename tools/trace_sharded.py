#!/usr/bin/env python
"""Wall-clock phases of ShardedSearcher.search under torchrun (GPU box, N >= 2)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
import bench
from vlg_matching_b200 import capi, sharded
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local); capi.init(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
queries = bench.make_queries(); span = sharded.batch_span(queries)
own = bench.shard_text(rank, bench.SHARD_BYTES)
host = np.concatenate([own, bench.shard_text(rank + 1, span)]) if rank + 1 < world else own
idx = capi.Index.build(host, global_offset=rank * bench.SHARD_BYTES, owned_len=bench.SHARD_BYTES, total_len=bench.SHARD_BYTES * world)
batch = capi.PackedBatch(queries)
S = sharded.ShardedSearcher(idx, rank, world, dist, dev)
marks = []
def mark(name): marks.append((name, time.perf_counter()))
orig_ag, orig_fix = S._all_gather, None
def ag(arr):
    mark("before all_gather"); r = orig_ag(arr); mark("after all_gather"); return r
S._all_gather = ag
for it in range(6):
    dist.barrier(); torch.cuda.synchronize()
    marks.clear(); mark("start")
    r = idx.search(batch, capi.WANT_STARTS | capi.KEEP_ON_DEVICE); mark("search")
    loc = np.stack([r.counts(), r.exit_cursors(), r.first_starts()], axis=1).astype(np.uint64); mark("stack")
    def fixup(entry):
        mark("before fixup"); n = r.fixup(entry); mark("after fixup (%d redone)" % n)
        return np.stack([r.counts(), r.exit_cursors(), r.first_starts()], axis=1).astype(np.uint64)
    table, refixed = sharded.chain_cursors(rank, world, loc, S._all_gather, None, fixup); mark("chain done")
    r.free()
if rank == world - 1:
    t0 = marks[0][1]
    for (n, t), (_, tp) in zip(marks[1:], marks[:-1]): print(f"  {n:32s} +{(t-tp)*1e3:7.3f} ms (at {(t-t0)*1e3:7.3f})")
dist.barrier(); dist.destroy_process_group()
