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


class ShardedDatabase:
    """This rank's shard of a database on its GPU, plus the gather: the multi-GPU form of a score search.

    One process per GPU (``torch.distributed`` initialised by the caller, NCCL backend; a single process
    without a process group is world size 1).  Two ways to shard:

    * ``ShardedDatabase.deal(residues, offsets, lengths)`` -- every rank holds the packed database on the
      host; targets are dealt by cumulative cells in snake order over the length-sorted list (config 2);
    * ``ShardedDatabase.from_blocks(shard, block, blocks_per_rank)`` -- every rank brings its own blocks
      (block b -> rank b mod world) of a database too large to sort in one place (config 5).

    ``search_scores`` returns the scores of ALL targets in database order as a device tensor on every rank;
    ``search_topk`` adds a device top-k per query and copies only the hits to the host.
    """

    def __init__(self, store, gather, n_local, stream=None):
        import torch

        self.store, self.gather, self.n_local = store, gather, n_local
        self.world = gather.world
        self.n_all = gather.n_all
        self.stream = stream or torch.cuda.current_stream()
        self._local = gather.new_local()
        self._gathered = (torch.empty(self.world * len(self._local), dtype=torch.int32, device="cuda")
                          if self.world > 1 else None)
        self._torch = torch

    @staticmethod
    def _world(group=None):
        import torch.distributed as dist

        if dist.is_available() and dist.is_initialized():
            return dist.get_world_size(group), dist.get_rank(group)
        return 1, 0

    @classmethod
    def deal(cls, residues, offsets, lengths, alphabet_length=24, device=None, group=None, stream=None):
        from .platform import cuda

        world, rank = cls._world(group)
        lengths = np.asarray(lengths)
        mine = shard_indices(lengths, world, rank) if world > 1 else np.arange(len(lengths))
        sr, so, sl = pack_shard(residues, offsets, lengths, mine) if world > 1 else (residues, offsets, lengths)
        store = cuda.DeviceStore(sr, so, np.ascontiguousarray(sl), alphabet_length, device=device)
        self = cls(store, ScoreGather(mine, len(lengths), world, "cuda", group), len(mine), stream)
        self.local_residues = int(np.asarray(sl, dtype=np.int64).sum())
        return self

    @classmethod
    def from_blocks(cls, shard, block, blocks_per_rank, alphabet_length=24, device=None, group=None, stream=None):
        from .platform import cuda

        world, _ = cls._world(group)
        residues, offsets, lengths = shard[:3]
        if len(lengths) != block * blocks_per_rank:
            raise ValueError("shard does not hold blocks_per_rank blocks")
        store = cuda.DeviceStore(residues, offsets, np.ascontiguousarray(lengths), alphabet_length, device=device)
        self = cls(store, BlockGather(block, blocks_per_rank, world, "cuda", group), len(lengths), stream)
        self.local_residues = int(np.asarray(lengths, dtype=np.int64).sum())
        return self

    def search_scores(self, query, gap_open, gap_extend, int_matrix, algorithm=3, out=None):
        """Score ``query`` (encoded bytes) against the whole database; int32 device tensor in database order.

        The call is complete when it returns (search, gather and un-permutation synchronised), like
        ``Aligner.align``: an un-synchronised all-gather followed by the next search's persistent kernels
        measured +21 ms per query on two B200s."""
        self.store.search_scores_device(query, gap_open, gap_extend, int_matrix, algorithm, self._local.data_ptr(),
                                        self.stream.cuda_stream)
        self.stream.synchronize()  # the store's per-search workspace is reused by the next query
        if self.world == 1:  # the local order is the database order; the tensor is reused by the next search
            return self._local[: self.n_all]
        res = self.gather.gather(self._local, self._gathered, out)
        self.stream.synchronize()
        return res

    def search_topk(self, queries, gap_open, gap_extend, int_matrix, k, algorithm=3, to_host=True):
        """Batched search: the ``k`` best (score, database index) per query; only the hits leave the device."""
        torch = self._torch
        k = min(k, self.n_all)
        vals = torch.empty((len(queries), k), dtype=torch.int32, device="cuda")
        idx = torch.empty((len(queries), k), dtype=torch.int32, device="cuda")
        for qi, q in enumerate(queries):
            scores = self.search_scores(q, gap_open, gap_extend, int_matrix, algorithm)
            # exact radix select on the device (csrc/topk.cu), unsorted; ties go to the lower database index
            self.store.topk_device(scores.data_ptr(), self.n_all, k, vals[qi].data_ptr(), idx[qi].data_ptr(),
                                   self.stream.cuda_stream)
        if not to_host:
            self.stream.synchronize()
            return vals, idx
        hv = torch.empty(vals.shape, dtype=vals.dtype).pin_memory()
        hi = torch.empty(idx.shape, dtype=idx.dtype).pin_memory()
        hv.copy_(vals, non_blocking=True)
        hi.copy_(idx, non_blocking=True)
        self.stream.synchronize()
        # order the k hits of every query: score descending, ties by ascending index
        order = np.lexsort((hi.numpy(), -hv.numpy().astype(np.int64)), axis=1)
        return torch.from_numpy(np.take_along_axis(hv.numpy(), order, 1)), torch.from_numpy(np.take_along_axis(hi.numpy(), order, 1))
