"""Multi-GPU protocol on CPU: world_size 2 and 3 over gloo.  vlg_matching_b200/sharded.chain_cursors is the
exchange step of SURVEY §8e (all-gather of (count, exit cursor, first start), left-to-right cursor chaining,
re-entry of a shard whose true entry cursor passes its first speculative match).  The per-shard search is
injected: here it is the CPU oracle run on the shard's bytes (a search that enters at cursor c is the search of
text[c:end], because viability only looks right and selection only looks left), on the GPU box it is
vlgb_search_batch / vlgb_result_fixup (tests/test_gpu_parity.py::test_entry_cursor_and_sharded_equivalence)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import pyoracle as O
from vlg_matching_b200 import synth
from vlg_matching_b200.sharded import batch_span, chain_cursors, shard_bounds

NONE = np.uint64(0xFFFFFFFFFFFFFFFF)


def make_inputs():
    text = synth.english_text(96 << 10, 31)
    q = synth.sample_queries(text, 24, 2, (2, 5), [(5, 60), (100, 1000)], 32, word_aligned=True)
    q += synth.sample_queries(text, 8, 3, (2, 3), [(0, 30)], 33, word_aligned=False)
    # dense lists: a match that starts left of a boundary ends right of it, so the next shard must re-enter
    q += [([b"e "], []), ([b"th"], []), ([b"e", b"t"], [(20, 300)]), ([b"a", b"o", b"e"], [(50, 400), (10, 500)]),
          ([b" ", b" "], [(900, 1000)])]
    return text, q


def shard_search(text, queries, lo, hi, end, entry):
    """[Q,3] (count, exit cursor, first start) of the shard [lo,hi) + halo up to `end`, entered at `entry`."""
    out = np.zeros((len(queries), 3), dtype=np.uint64)
    for qi, (subs, gaps) in enumerate(queries):
        c = max(int(entry[qi]), lo)
        out[qi] = (0, entry[qi], NONE)
        if c >= end:
            continue
        (_, starts, tup), = O.search(text[c:end], [(subs, gaps)], want_tuples=True)
        keep = starts + c < hi
        if keep.any():
            t = tup[keep] + c
            out[qi] = (keep.sum(), t[-1, -1] + len(subs[-1]), t[0, 0])
    return out


def worker(rank, world, port, result_q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    text, queries = make_inputs()
    lo, hi, end = shard_bounds(text.size, world, batch_span(queries))[rank]
    local = shard_search(text, queries, lo, hi, end, np.zeros(len(queries), dtype=np.uint64))  # speculative

    def all_gather(arr):
        t = torch.from_numpy(arr.view(np.int64).copy())
        out = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(out, t)
        return torch.stack(out).numpy().view(np.uint64)

    def fixup(entry):
        return shard_search(text, queries, lo, hi, end, entry)

    table, refixed = chain_cursors(rank, world, local, all_gather, fixup)  # decisions: the library's vlgb_chain_plan
    result_q.put((rank, table.tolist(), refixed))
    dist.barrier()
    dist.destroy_process_group()


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("world", [2, 3])
def test_chain_cursors_over_gloo(world):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = [q.get(timeout=240) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    tables = {r: np.asarray(t, dtype=np.uint64) for r, t, _ in got}
    for r in range(1, world):
        assert np.array_equal(tables[0], tables[r])  # every rank derives the same final table
    text, queries = make_inputs()
    whole = O.search(text, queries, want_tuples=True, threads=4)
    counts = tables[0][:, :, 0].sum(axis=0)
    for qi, (_, starts, tup) in enumerate(whole):
        assert int(counts[qi]) == len(starts), (qi, queries[qi])
        if len(starts):
            last_end = int(tup[-1, -1]) + len(queries[qi][0][-1])
            assert int(tables[0][world - 1, qi, 1]) == last_end  # the cursor the last shard leaves = whole-text cursor
    assert sum(rf for _, _, rf in got) > 0  # at least one shard had to re-enter


def test_shard_bounds_cover_text_once():
    b = shard_bounds(1000, 3, 50)
    assert b[0][0] == 0 and b[-1][1] == 1000 and all(b[i][1] == b[i + 1][0] for i in range(2))
    assert all(e == min(1000, h + 50) for _, h, e in b)


def test_chain_plan_equals_numpy_protocol():
    """vlgb_chain_plan (the host logic inside vlgb_search_sharded) against the numpy statement of the same rule."""
    from vlg_matching_b200 import capi
    from vlg_matching_b200.sharded import implied_entries
    rng = np.random.default_rng(7)
    for G, Q in ((1, 5), (2, 1), (3, 40), (8, 257)):
        rows = np.zeros((G, Q, 3), dtype=np.uint64)
        rows[:, :, 0] = rng.integers(0, 3, (G, Q))
        rows[:, :, 1] = rng.integers(0, 1000, (G, Q))
        rows[:, :, 2] = rng.integers(0, 1000, (G, Q))
        used = rng.integers(0, 3, (G, Q)).astype(np.uint64) * np.uint64(300)
        entries, redo = capi.chain_plan(rows, used)
        want = implied_entries(rows)
        cnt, first = rows[:, :, 0], rows[:, :, 2]
        moved = want != used
        harmless = (want > used) & ((cnt == 0) | (want <= first))
        assert np.array_equal(entries, want) and np.array_equal(redo, (moved & ~harmless).any(axis=1))
