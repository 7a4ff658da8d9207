"""Multi-GPU search: contiguous text shards that overlap by the maximum query span, one process per
GPU, every rank answers the whole query batch on its shard.

The only exchange is tiny: per (query, shard) the match count, the exit cursor and the first reported
start (24 B).  One all-gather, then the non-overlap cursor is chained left to right: shard g enters
with the cursor shard g-1 left (vlg_iterator::pull_forward state, vlg_index.hpp:253-264).  Shards
whose true entry cursor lies beyond their first speculative match re-enter on the device
(vlgb_result_fixup), all at once, followed by another all-gather; this repeats until no entry cursor
moves (chain_cursors).  Every rank derives the same decisions from the gathered data, so the collective
sequence is identical on all ranks.  A match is owned by the shard that contains its first
subpattern's start, so no match is reported twice.

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


def implied_entries(table: np.ndarray):
    """Entry cursor of every shard implied by the gathered rows: shard g enters with the cursor the nearest
    shard to its left that reported a match left behind (0 for shard 0)."""
    world, nq = table.shape[0], table.shape[1]
    cursor = np.zeros(nq, dtype=np.uint64)
    entries = np.zeros((world, nq), dtype=np.uint64)
    for g in range(world):
        entries[g] = cursor
        cursor = np.where(table[g, :, 0] > 0, table[g, :, 1], cursor)
    return entries


def chain_cursors(rank: int, world: int, local: np.ndarray, all_gather, broadcast, fixup):
    """local: uint64 [Q, 3] = (count, exit_cursor, first_start) of this rank's speculative result
    (entry cursor 0).  all_gather(arr) -> [world, Q, 3]; fixup(entry[Q]) -> new local [Q, 3] (re-enters the
    shard on the device for the queries that can change, vlgb_result_fixup).  `broadcast` is unused (kept for
    callers of the earlier sequential protocol).
    Returns (table [world, Q, 3] with the final values, number of shard re-entries over all rounds).

    Rounds instead of a left-to-right walk: after the first all-gather every shard knows the cursor its left
    neighbours would hand it.  All shards whose reported set can change under that cursor re-enter AT THE SAME
    TIME, then one more all-gather.  A re-entered shard almost always leaves the same exit cursor (orbits of the
    monotone successor coalesce within a few matches), so the second table already is the fixpoint; when an
    exit cursor did move, the shards to its right go again.  Shard g is final after round g at the latest, so
    the loop ends after <= world rounds -- typically 1 -- and costs one small collective per round instead of
    one per shard.  Every rank derives the same decisions from the same table, so the collective sequence is
    identical on all ranks."""
    del broadcast
    table = all_gather(local)
    nq = local.shape[0]
    used = np.zeros((world, nq), dtype=np.uint64)  # entry cursor every shard's current row was computed with
    refixed = 0
    for _ in range(world):
        entries = implied_entries(table)
        cnt, first = table[:, :, 0], table[:, :, 2]
        # same rule as vlgb_result_fixup: a query is searched again unless the cursor is unchanged, or moved
        # right without passing the first reported start (no match: nothing to report either way)
        moved = entries != used
        harmless = (entries > used) & ((cnt == 0) | (entries <= first))
        redo = (moved & ~harmless).any(axis=1)
        if moved[rank].any():
            local = fixup(entries[rank].copy())  # records the cursors; searches again only what can change
        used = entries
        if not redo.any():
            # rows of shards without a match carry the cursor through
            table[:, :, 1] = np.where(cnt > 0, table[:, :, 1], entries)
            break
        refixed += int(redo.sum())
        table = all_gather(local)
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
        # persistent pinned host + device buffers: two async copies and one NCCL call, no allocation per step
        key = arr.shape
        if getattr(self, "_ag_key", None) != key:
            n = int(np.prod(arr.shape))
            self._ag_host_in = torch.empty(n, dtype=torch.int64).pin_memory()
            self._ag_host_out = torch.empty(self.world * n, dtype=torch.int64).pin_memory()
            self._ag_dev_in = torch.empty(n, dtype=torch.int64, device=self.device)
            self._ag_dev_out = torch.empty(self.world * n, dtype=torch.int64, device=self.device)
            self._ag_key = key
        self._ag_host_in.numpy()[:] = arr.reshape(-1).view(np.int64)
        self._ag_dev_in.copy_(self._ag_host_in, non_blocking=True)
        self.dist.all_gather_into_tensor(self._ag_dev_out, self._ag_dev_in)
        self._ag_host_out.copy_(self._ag_dev_out, non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        return self._ag_host_out.numpy().view(np.uint64).reshape((self.world,) + tuple(arr.shape)).copy()

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
