"""Multi-GPU search: contiguous text shards that overlap by the maximum query span, one rank per GPU, every rank
answers the whole query batch on its shards.

The exchange lives in the library (vlgb_search_sharded, csrc/engine.cu + csrc/comm.hpp): per round ONE NCCL
all-gather of (match count, exit cursor, first start) per query and rank -- 24 B -- then every rank derives every
rank's true entry cursor (the pull_forward state handed over the boundary, vlg_index.hpp:253-264) and the ranks whose
reported set can change re-enter on the device (vlgb_result_fixup).  A match is owned by the shard that contains
its first subpattern's start, so no match is reported twice and no match data crosses NVLink.

This module is the thin Python caller: shard geometry, communicator bootstrap over torch.distributed (plumbing:
only the 128-byte NCCL id travels through it), and `chain_cursors`, the same round protocol written against
injected search/all-gather functions so that it can run on CPU over gloo -- its decisions come from the library's
own host logic (vlgb_chain_plan).
"""
from __future__ import annotations

import numpy as np

NONE = np.uint64(0xFFFFFFFFFFFFFFFF)


def shard_bounds(total_len: int, world: int, span: int):
    """[(lo, hi, end)] per shard: owned [lo, hi), resident [lo, end) with end = min(total, hi + span)."""
    cuts = [total_len * g // world for g in range(world + 1)]
    return [(cuts[g], cuts[g + 1], min(total_len, cuts[g + 1] + span)) for g in range(world)]


def batch_span(queries) -> int:
    """Largest distance from a match start to one past its last byte over the batch."""
    return max((sum(len(s) for s in subs) + sum(g[1] for g in gaps) for subs, gaps in queries), default=0)


def implied_entries(table: np.ndarray):
    """Entry cursor of every shard implied by the gathered rows (numpy statement of vlgb_chain_plan's `entries`,
    kept for the tests that compare the two)."""
    world, nq = table.shape[0], table.shape[1]
    cursor = np.zeros(nq, dtype=np.uint64)
    entries = np.zeros((world, nq), dtype=np.uint64)
    for g in range(world):
        entries[g] = cursor
        cursor = np.where(table[g, :, 0] > 0, table[g, :, 1], cursor)
    return entries


def chain_cursors(rank: int, world: int, local: np.ndarray, all_gather, fixup, plan=None):
    """The round protocol of vlgb_search_sharded with the search injected (CPU tests over gloo).
    local: uint64 [Q, 3] = (count, exit_cursor, first_start) of this rank's speculative result (entry cursor 0);
    all_gather(arr) -> [world, Q, 3]; fixup(entry[Q]) -> new local [Q, 3]; plan(rows, used) -> (entries, redo),
    default the library's vlgb_chain_plan.
    Returns (table [world, Q, 3] with the final values, number of shard re-entries over all rounds).

    After the first all-gather every shard knows the cursor its left neighbours would hand it.  All shards whose
    reported set can change under that cursor re-enter AT THE SAME TIME, then one more all-gather.  A re-entered
    shard almost always leaves the same exit cursor (orbits of the monotone successor coalesce within a few
    matches), so the second table already is the fixpoint; when an exit cursor did move, the shards to its right go
    again.  Shard g is final after round g at the latest.  Every rank derives the same decisions from the same
    table, so the collective sequence is identical on all ranks."""
    if plan is None:
        from . import capi
        plan = capi.chain_plan
    table = all_gather(local)
    nq = local.shape[0]
    used = np.zeros((world, nq), dtype=np.uint64)  # entry cursor every shard's current row was computed with
    refixed = 0
    for _ in range(world + 1):
        entries, redo = plan(table, used)
        if (entries[rank] != used[rank]).any():
            local = fixup(entries[rank].copy())  # records the cursors; searches again only what can change
        used = entries
        if not redo.any():
            # rows of shards without a match carry the cursor through
            table[:, :, 1] = np.where(table[:, :, 0] > 0, table[:, :, 1], entries)
            break
        refixed += int(redo.sum())
        table = all_gather(local)
    return table, refixed


def comm_init(rank: int, world: int, dist=None):
    """Create the library's NCCL communicator for this rank.  The 128-byte id travels over torch.distributed."""
    from . import capi
    if world == 1:
        capi.comm_init(None, 1, 0)
        return
    obj = [capi.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(obj, src=0)
    capi.comm_init(obj[0], world, rank)


class ShardedSearcher:
    """One instance per rank: `shards` are this rank's capi.Index objects, consecutive pieces of the text, each built
    over [lo, end) with global_offset=lo, owned_len=hi-lo, total_len=total.  comm_init() must have run."""

    def __init__(self, shards):
        self.shards = list(shards) if isinstance(shards, (list, tuple)) else [shards]

    def search(self, batch, want):
        """-> (results per local shard, global per-query counts over all ranks, stats of the exchange)."""
        from . import capi
        return capi.search_sharded(self.shards, batch, want)
