"""Multi-GPU search: contiguous text shards that overlap by the maximum query span, one process per
GPU, every rank answers the whole query batch on its shard.

The only exchange is tiny: per (query, shard) the match count, the exit cursor and the first reported
start (24 B).  One all-gather, then the non-overlap cursor is chained left to right: shard g enters
with the cursor shard g-1 left (vlg_iterator::pull_forward state, vlg_index.hpp:253-264).  A shard
whose true entry cursor lies beyond its first speculative match re-enters on the device
(vlgb_result_fixup) and broadcasts its new counters; every rank derives the same decision from the
gathered data, so the collective sequence is identical on all ranks.  A match is owned by the shard
that contains its first subpattern's start, so no match is reported twice.

torch.distributed is plumbing only (NCCL on the GPU box, gloo in the CPU tests); `search_fn` /
`fixup_fn` are injected so the protocol itself can be tested without a GPU.
"""
from __future__ import annotations

import numpy as np

NONE = np.uint64(0xFFFFFFFFFFFFFFFF)


def shard_bounds(total_len: int, world: int, span: int):
    """[(lo, hi, end)] per rank: owned [lo, hi), resident [lo, end) with end = min(total, hi + span)."""
    cuts = [total_len * g // world for g in range(world + 1)]
    return [(cuts[g], cuts[g + 1], min(total_len, cuts[g + 1] + span)) for g in range(world)]


def batch_span(queries) -> int:
    """Largest distance from a match start to one past its last byte over the batch."""
    return max((sum(len(s) for s in subs) + sum(g[1] for g in gaps) for subs, gaps in queries), default=0)


def chain_cursors(rank: int, world: int, local: np.ndarray, all_gather, broadcast, fixup):
    """local: uint64 [Q, 3] = (count, exit_cursor, first_start) of this rank's speculative result
    (entry cursor 0).  all_gather(arr) -> [world, Q, 3]; broadcast(arr, src) -> arr of rank src;
    fixup(entry[Q]) -> new local [Q, 3] (runs on this rank only when needed).
    Returns (table [world, Q, 3] with the final values, number of shards that re-entered)."""
    table = all_gather(local)
    nq = local.shape[0]
    cursor = np.zeros(nq, dtype=np.uint64)
    refixed = 0
    for g in range(world):
        cnt, ext, first = table[g, :, 0], table[g, :, 1], table[g, :, 2]
        need = (cnt > 0) & (cursor > first)
        if g == rank:
            mine = fixup(cursor.copy())  # records the entry cursors; re-searches only the `need` queries
        if need.any():
            refixed += 1
            buf = mine if g == rank else np.zeros((nq, 3), dtype=np.uint64)
            table[g] = broadcast(buf, g)
            cursor = table[g, :, 1].copy()
        else:
            cursor = np.where(cnt > 0, ext, cursor)
            table[g, :, 1] = cursor
    return table, refixed


class ShardedSearcher:
    """One instance per rank.  `index` is this rank's capi.Index built over [lo, end) with
    global_offset=lo, owned_len=hi-lo, total_len=total."""

    def __init__(self, index, rank: int, world: int, dist=None, device=None):
        self.index, self.rank, self.world, self.dist, self.device = index, rank, world, dist, device

    def _all_gather(self, arr):
        if self.world == 1:
            return arr[None].copy()
        import torch
        t = torch.from_numpy(arr.view(np.int64)).to(self.device)
        out = torch.empty((self.world,) + tuple(t.shape), dtype=t.dtype, device=self.device)
        self.dist.all_gather_into_tensor(out, t)
        return out.cpu().numpy().view(np.uint64)

    def _broadcast(self, arr, src):
        import torch
        t = torch.from_numpy(arr.view(np.int64).copy()).to(self.device)
        self.dist.broadcast(t, src)
        return t.cpu().numpy().view(np.uint64)

    def search(self, batch, want):
        """-> (local capi.Result holding this shard's owned matches, global counts [Q], shards re-entered)."""
        r = self.index.search(batch, want)
        local = np.stack([r.counts(), r.exit_cursors(), r.first_starts()], axis=1).astype(np.uint64)

        def fixup(entry):
            r.fixup(entry)
            return np.stack([r.counts(), r.exit_cursors(), r.first_starts()], axis=1).astype(np.uint64)

        table, refixed = chain_cursors(self.rank, self.world, local, self._all_gather, self._broadcast, fixup)
        return r, table[:, :, 0].sum(axis=0), refixed
