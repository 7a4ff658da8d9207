#!/usr/bin/env python
"""Under torchrun: a few resident + e2e steps of the bench workload with VLGB_TRACE on rank 0 (host timeline of the
sharded search: where the exchange time goes).  usage: torchrun ... tools/trace_n.py"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ["LOCAL_RANK"])
if rank == 0:
    os.environ["VLGB_TRACE"] = "1"
import numpy as np, torch, torch.distributed as dist
import bench
from vlg_matching_b200 import capi, sharded
bench.bind_to_gpu_numa_node(local)
torch.cuda.set_device(local); capi.init(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
sharded.comm_init(rank, world, dist)
q = bench.make_queries(); span = sharded.batch_span(q)
own = bench.shard_text(rank, bench.SHARD_BYTES)
host = np.concatenate([own, bench.shard_text(rank + 1, span)]) if rank + 1 < world else own
idx = capi.Index.build(host, global_offset=rank * bench.SHARD_BYTES, owned_len=bench.SHARD_BYTES, total_len=bench.SHARD_BYTES * world)
b = capi.PackedBatch(q)
for mode, want in (("resident", capi.WANT_STARTS | capi.KEEP_ON_DEVICE), ("e2e", capi.WANT_STARTS)):
    for i in range(5):
        dist.barrier(); torch.cuda.synchronize()
        if rank == 0 and i == 4: print(f"==== {mode} step", file=sys.stderr, flush=True)
        t0 = time.perf_counter()
        rs, gc, st = capi.search_sharded([idx], b, want)
        dt = (time.perf_counter() - t0) * 1e3
        if i == 4: print(f"[rank {rank}] {mode}: {dt:.3f} ms, exchange {st['exchange_ms']:.3f} ms", file=sys.stderr, flush=True)
        for r in rs: r.free()
capi.comm_destroy(); dist.barrier(); dist.destroy_process_group()
