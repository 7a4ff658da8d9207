#!/usr/bin/env python
"""Host trace (VLGB_TRACE) of the pipelined end-to-end loop of bench.py (VLGB_ASYNC_COPY) under torchrun, printed for
ONE rank (TRACE_RANK, default the last) so that the library's marks are not interleaved.  Measurement only."""
import os, sys, time
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
trace_rank = int(os.environ.get("TRACE_RANK", world - 1))
if rank == trace_rank:
    os.environ["VLGB_TRACE"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
import bench
from vlg_matching_b200 import capi, sharded
torch.cuda.set_device(local); capi.init(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
sharded.comm_init(rank, world, dist if world > 1 else None)
queries = bench.make_queries(); span = sharded.batch_span(queries)
S = bench.synth()
own = bench.shard_text(rank, bench.SHARD_BYTES)
host = np.concatenate([own, bench.shard_text(rank + 1, span)]) if rank + 1 < world else own
idx = capi.Index.build(host, global_offset=rank * bench.SHARD_BYTES, owned_len=bench.SHARD_BYTES, total_len=bench.SHARD_BYTES * world)
pattern_file = b"\n".join(S.to_benchmark_syntax(q) for q in queries) + b"\n"
def say(*a):
    if rank == trace_rank: print(*a, file=sys.stderr, flush=True)
for mode, flags in (("serial", 0), ("pipelined", capi.ASYNC_COPY)):
    prev = None
    for it in range(int(os.environ.get("STEPS", "6"))):
        if world > 1 and it == 0: dist.barrier()
        t0 = time.perf_counter()
        say(f"---- {mode} step {it}")
        b = capi.ParsedBatch(pattern_file, capi.SYNTAX_BENCHMARK); t1 = time.perf_counter()
        rs, gc, st = capi.search_sharded([idx], b, capi.WANT_STARTS | flags); t2 = time.perf_counter()
        if flags and prev is not None:
            prev[0].starts(0); prev[0].starts(prev[1] - 1); t3 = time.perf_counter(); prev[0].free()
        elif not flags:
            rs[0].starts(0); rs[0].starts(len(b) - 1); t3 = time.perf_counter(); rs[0].free()
        else:
            t3 = time.perf_counter()
        t4 = time.perf_counter()
        prev = (rs[0], len(b)) if flags else None
        say(f"     python: parse {1e3*(t1-t0):.3f}  search_sharded {1e3*(t2-t1):.3f}  read {1e3*(t3-t2):.3f}  free {1e3*(t4-t3):.3f}  total {1e3*(t4-t0):.3f} ms")
    if prev: prev[0].starts(0); prev[0].free()
capi.comm_destroy()
if world > 1:
    dist.barrier(); dist.destroy_process_group()
