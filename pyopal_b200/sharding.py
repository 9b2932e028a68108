"""Database sharding across the GPUs of one box and the score gather.

The reference's only data parallelism is a thread pool over contiguous
chunks of equal sequence COUNT (/root/reference/src/pyopal/_align.py:145-172),
which is unbalanced on skewed length distributions.  Here the database is cut
by cumulative CELL count (= residues, the query being common): targets sorted
by length are dealt to the ranks in snake order, so every rank receives the
same length mix, the longest targets spread out and the per-rank residue
totals differ by at most one target.  Targets are independent, so the only
exchange is the gather of the per-shard score arrays (one int32 per target),
done with ``torch.distributed`` -- NCCL over NVLink on the GPUs, gloo in the
CPU tests.
"""

import numpy as np


def shard_indices(lengths, world, rank):
    """Database indices (ascending) owned by ``rank`` of ``world``."""
    lengths = np.asarray(lengths)
    order = np.argsort(-lengths.astype(np.int64), kind="stable")
    pos = np.arange(len(order))
    rnd, k = pos // world, pos % world
    owner = np.where(rnd % 2 == 0, k, world - 1 - k)
    return np.sort(order[owner == rank])


def pack_shard(residues, offsets, lengths, mine):
    """Packed (residues, offsets, lengths) of the targets ``mine`` of a packed database."""
    residues = np.asarray(residues)
    offsets = np.asarray(offsets, dtype=np.int64)
    sl = np.ascontiguousarray(np.asarray(lengths)[mine]).astype(np.int32)
    so = np.zeros(len(mine), dtype=np.int64)
    if len(mine) > 1:
        np.cumsum(sl[:-1], out=so[1:])
    total = int(sl.sum())
    if total == 0:
        return np.zeros(0, dtype=np.uint8), so, sl
    # source position of every destination byte
    src = np.repeat(offsets[mine] - so, sl) + np.arange(total, dtype=np.int64)
    return np.ascontiguousarray(residues[src]), so, sl


def padded_shard_size(n_all, world):
    """Slots per rank in the gathered buffer (equal for every rank, as all_gather needs)."""
    return (n_all + world - 1) // world + 1


class ScoreGather:
    """All-gather of per-shard scores and their un-permutation to database order.

    ``mine`` are this rank's database indices; ``device`` is where the score
    tensors live ("cuda" with the NCCL backend, "cpu" with gloo).
    """

    def __init__(self, mine, n_all, world, device, group=None):
        import torch
        import torch.distributed as dist

        self.world = world
        self.n_all = n_all
        self.n_pad = padded_shard_size(n_all, world)
        self.group = group
        self.device = device
        self._dist = dist
        self._torch = torch
        if len(mine) > self.n_pad:
            raise ValueError("shard larger than the padded shard size")
        idx = torch.full((self.n_pad,), -1, dtype=torch.int64, device=device)
        idx[: len(mine)] = torch.as_tensor(np.asarray(mine, dtype=np.int64), device=device)
        if world > 1:
            all_idx = torch.empty(world * self.n_pad, dtype=torch.int64, device=device)
            dist.all_gather_into_tensor(all_idx, idx, group=group)
        else:
            all_idx = idx
        self.valid = (all_idx >= 0).nonzero().squeeze(1)
        self.dest = all_idx[self.valid]
        if len(self.dest) != n_all or len(torch.unique(self.dest)) != n_all:
            raise ValueError("shards do not partition the database")

    def new_local(self):
        """A score buffer of the padded shard size for this rank's search to fill."""
        return self._torch.zeros(self.n_pad, dtype=self._torch.int32, device=self.device)

    def gather(self, local, gathered=None, out=None):
        """All-gather ``local`` (n_pad int32) and return the scores in database order."""
        torch = self._torch
        if self.world > 1:
            if gathered is None:
                gathered = torch.empty(self.world * self.n_pad, dtype=torch.int32, device=self.device)
            self._dist.all_gather_into_tensor(gathered, local, group=self.group)
        else:
            gathered = local
        if out is None:
            out = torch.empty(self.n_all, dtype=torch.int32, device=self.device)
        out[self.dest] = gathered[self.valid]
        return out


class BlockGather:
    """Score gather for a database dealt to the ranks in equal BLOCKS (block b -> rank b % world).

    Used for databases too large to length-sort globally on one host (config 5 of
    BASELINE.json, 5e7 targets): every rank draws / loads its own blocks, scores them
    in local order, and database order is recovered from the all-gathered buffer by a
    strided view -- no index arrays (the 5e7-entry permutation of ScoreGather would
    cost 400 MB per rank).  `blocks_per_rank * world` blocks of `block` targets each.
    """

    def __init__(self, block, blocks_per_rank, world, device, group=None):
        import torch
        import torch.distributed as dist

        self.block, self.nb, self.world = int(block), int(blocks_per_rank), int(world)
        self.n_local = self.block * self.nb
        self.n_all = self.n_local * self.world
        self.group = group
        self.device = device
        self._torch, self._dist = torch, dist

    def new_local(self):
        return self._torch.zeros(self.n_local, dtype=self._torch.int32, device=self.device)

    def gather(self, local, gathered=None, out=None):
        """All-gather ``local`` (n_local int32); returns the scores in database order."""
        torch = self._torch
        if self.world == 1:
            return local
        if gathered is None:
            gathered = torch.empty(self.world * self.n_local, dtype=torch.int32, device=self.device)
        self._dist.all_gather_into_tensor(gathered, local, group=self.group)
        if out is None:
            out = torch.empty(self.n_all, dtype=torch.int32, device=self.device)
        # gathered[r, k, i] is target (k * world + r) * block + i
        out.view(self.nb, self.world, self.block).copy_(gathered.view(self.world, self.nb, self.block).permute(1, 0, 2))
        return out
