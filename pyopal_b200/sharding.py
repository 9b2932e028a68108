"""Database sharding across the GPUs of one box, one process per GPU.

The reference's only data parallelism is a thread pool over contiguous
chunks of equal sequence COUNT (/root/reference/src/pyopal/_align.py:145-172),
which is unbalanced on skewed length distributions.  Here the database is cut
by cumulative CELL count (= residues, the query being common): targets sorted
by length are dealt to the ranks in snake order, so every rank receives the
same length mix, the longest targets spread out and the per-rank residue
totals differ by at most one target.  Targets are independent, so the only
exchange is the gather of the per-shard score arrays (one int32 per target):
one ``ncclAllGather`` issued by ``libpyopal_cuda.so`` itself (csrc/multi.cu,
NCCL bound with ``dlopen``), followed by an un-permutation kernel.

Two layouts, same C code underneath:

* ONE process, every visible GPU: nothing to do here -- `pyopal.Database`
  builds a `platform.cuda.MultiDeviceStore` and `Aligner.align` /
  `pyopal.align` use all the GPUs;
* one process PER GPU (``torchrun``, ``mpirun``, ``srun`` ...): `ShardedDatabase`
  below.  The processes only have to agree on 128 bytes, the NCCL unique id:
  `exchange_nccl_id` passes it through a file, or the launcher's own broadcast
  can be used.
"""

import array
import ctypes
import os
import time

from .platform import cuda as _cuda


def shard_indices(lengths, world, rank):
    """Database indices (ascending) owned by ``rank`` of ``world`` (the deal of csrc/multi.cu)."""
    owner = _cuda.deal(lengths, world)
    return array.array("i", [i for i, o in enumerate(owner) if o == rank])


def exchange_nccl_id(rank, key=None, directory=None, timeout=120.0):
    """The communicator's 128-byte NCCL id, created on rank 0 and read by the others from a file.

    ``key`` names the rendezvous; by default ``MASTER_PORT`` plus the parent process id (all the ranks of one
    ``torchrun`` / ``mp.spawn`` launch share their parent), so concurrent or repeated launches do not collide.
    """
    if key is None:
        key = f"{os.environ.get('MASTER_PORT', '0')}_{os.getppid()}"
    directory = directory or os.environ.get("TMPDIR", "/tmp")
    path = os.path.join(directory, f"pyopal_nccl_id_{key}")
    if rank == 0:
        blob = _cuda.nccl_unique_id()
        tmp = f"{path}.{os.getpid()}.tmp"
        with open(tmp, "wb") as handle:
            handle.write(blob)
        os.replace(tmp, path)
        return blob
    deadline = time.monotonic() + timeout
    while time.monotonic() < deadline:
        try:
            with open(path, "rb") as handle:
                blob = handle.read()
            if len(blob) == 128:
                return blob
        except FileNotFoundError:
            pass
        time.sleep(0.01)
    raise TimeoutError(f"rank {rank}: no NCCL id at {path} after {timeout} s")


def discard_nccl_id(key=None, directory=None):
    """Remove the rendezvous file (rank 0, once every rank has created its handle)."""
    if key is None:
        key = f"{os.environ.get('MASTER_PORT', '0')}_{os.getppid()}"
    path = os.path.join(directory or os.environ.get("TMPDIR", "/tmp"), f"pyopal_nccl_id_{key}")
    try:
        os.remove(path)
    except OSError:
        pass


class ShardedDatabase(_cuda.MultiDeviceStore):
    """This process's shard of a database on its GPU plus the NCCL gather (one process per GPU).

    * ``ShardedDatabase.deal(residues, offsets, lengths, rank=, world=, nccl_id=)`` -- every rank holds the packed
      database on the host; the cell-balanced deal is computed inside the library and this rank uploads its
      shard only (config 2 of BASELINE.json);
    * ``ShardedDatabase.from_blocks(shard, block, blocks_per_rank, rank=, world=, nccl_id=)`` -- every rank brings
      its own blocks (block b -> rank b mod world) of a database too large to sort in one place (config 5).

    Searches are collective: every rank makes the same call; all of them receive the scores in database order.
    `search`, `search_batch`, `search_batch_topk`, `stats`, `time_begin` / `time_end` are those of
    `platform.cuda.MultiDeviceStore`.
    """

    @classmethod
    def _create(cls, fn, args, n_all, alphabet_length, device, world=1):
        handle = ctypes.c_void_p(None)
        if world > 1:
            cls._ensure_nccl()
        _cuda.check(fn(*args, ctypes.byref(handle)))
        return cls._wrap(handle, n_all, alphabet_length, [device])

    @classmethod
    def deal(cls, residues, offsets, lengths, alphabet_length=24, *, rank=0, world=1, nccl_id=None, device=None):
        lib = _cuda.load_library()
        device = _cuda.default_device() if device is None else device
        lens = _cuda.as_int32_array(lengths)
        keep = (residues, offsets, lens)
        self = cls._create(
            lib.pyopal_cuda_mdb_create_rank_dealt,
            (device, rank, world, _cuda._addr(nccl_id), _cuda._addr(residues), _cuda._addr(offsets), _cuda._addr(lens),
             len(lens), alphabet_length),
            len(lens), alphabet_length, device, world,
        )
        del keep
        self.rank, self.n_all = rank, len(lens)
        return self

    @classmethod
    def from_blocks(cls, shard, block, blocks_per_rank, alphabet_length=24, *, rank=0, world=1, nccl_id=None, device=None):
        lib = _cuda.load_library()
        device = _cuda.default_device() if device is None else device
        residues, offsets, lengths = shard[:3]
        lens = _cuda.as_int32_array(lengths)
        if len(lens) != block * blocks_per_rank:
            raise ValueError("shard does not hold blocks_per_rank blocks")
        n_all = block * blocks_per_rank * world
        index = None
        if world > 1:
            # block j of this rank is database block j * world + rank
            index = array.array("i", bytes(4 * len(lens)))
            view = memoryview(index)
            for j in range(blocks_per_rank):
                first = (j * world + rank) * block
                view[j * block : (j + 1) * block] = array.array("i", range(first, first + block))
        self = cls._create(
            lib.pyopal_cuda_mdb_create_rank,
            (device, rank, world, _cuda._addr(nccl_id), _cuda._addr(residues), _cuda._addr(offsets), _cuda._addr(lens),
             len(lens), _cuda._addr(index), n_all, alphabet_length),
            n_all, alphabet_length, device, world,
        )
        self.rank, self.n_all = rank, n_all
        return self
