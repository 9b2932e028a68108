"""Multi-rank path on CPU: world_size 2 and 3 over gloo.

Each rank scores its cell-balanced shard (here with the oracle -- tests may),
the per-shard score arrays are all-gathered and un-permuted exactly as
bench.py does over NCCL, and the result must equal the single-rank search.
"""

import os
import socket

import numpy as np
import pytest

import synthetic
from pyopal_b200 import sharding


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, ret):
    import torch
    import torch.distributed as dist

    import oracle
    import pyopal_b200 as pyopal

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        residues, offsets, lengths = synthetic.database(900, 700, 4.5, 0.9, 0, 3000, giants=(9000, 7000, 5000))
        q = synthetic.random_residues(np.random.Generator(np.random.PCG64(901)), 60).tobytes()
        M = pyopal.Aligner("BLOSUM62")._int_matrix
        mine = sharding.shard_indices(lengths, world, rank)
        sr, so, sl = sharding.pack_shard(residues, offsets, lengths, mine)
        import array

        local = oracle.search_flat(q, sr.tobytes() + b"\0", array.array("q", so.tolist()), array.array("i", sl.tolist()),
                                   M, 24, 11, 1, "sw", "score", 1)["score"]
        gather = sharding.ScoreGather(mine, len(lengths), world, "cpu")
        buf = gather.new_local()
        buf[: len(mine)] = torch.tensor(local, dtype=torch.int32)
        got = gather.gather(buf).tolist()
        if rank == 0:
            want = oracle.search_flat(q, residues.tobytes() + b"\0", array.array("q", offsets.tolist()),
                                      array.array("i", lengths.tolist()), M, 24, 11, 1, "sw", "score", 4)["score"]
            ret["ok"] = got == want
            ret["residues"] = [int(lengths[sharding.shard_indices(lengths, world, r)].sum()) for r in range(world)]
            ret["longest"] = int(lengths.max())
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_search_matches_single_rank(oracle, world):
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        ret = mgr.dict()
        port = _free_port()
        procs = [ctx.Process(target=_worker, args=(r, world, port, ret)) for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(180)
        assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
        assert ret["ok"]
        # cell balance: snake dealing keeps every shard within one longest target of the others
        assert max(ret["residues"]) - min(ret["residues"]) <= ret["longest"]


def test_shards_partition_and_balance():
    rng = np.random.Generator(np.random.PCG64(5))
    lengths = synthetic.random_lengths(rng, 10_000, 5.65, 0.68, 2, 35_213)
    for world in (1, 2, 4, 8):
        parts = [sharding.shard_indices(lengths, world, r) for r in range(world)]
        allidx = np.concatenate(parts)
        assert len(allidx) == len(lengths) and len(np.unique(allidx)) == len(lengths)
        tot = [int(lengths[p].sum()) for p in parts]
        assert max(tot) - min(tot) <= int(lengths.max())
        assert all(len(p) <= sharding.padded_shard_size(len(lengths), world) for p in parts)


def test_pack_shard_roundtrip():
    residues, offsets, lengths = synthetic.database(7, 300, 4.0, 0.8, 0, 500)
    mine = sharding.shard_indices(lengths, 3, 1)
    sr, so, sl = sharding.pack_shard(residues, offsets, lengths, mine)
    for k, i in enumerate(mine):
        assert bytes(sr[so[k] : so[k] + sl[k]]) == bytes(residues[offsets[i] : offsets[i] + lengths[i]])
