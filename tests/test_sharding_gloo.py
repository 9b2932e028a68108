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


def _block_worker(rank, world, port, ret):
    """Config-5 shape: blocks dealt round-robin, batched queries, view-based un-permutation."""
    import array

    import torch
    import torch.distributed as dist

    import oracle
    import pyopal_b200 as pyopal

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        block, nb = 150, 3
        M = pyopal.Aligner("BLOSUM62")._int_matrix
        queries = [q.tobytes() for q in synthetic.config5_queries(3)]
        sr, so, sl, mine = synthetic.config5_shard(rank, world, nb * world, block)
        assert mine == list(range(rank, nb * world, world))
        gather = sharding.BlockGather(block, nb, world, "cpu")
        rows = []
        for q in queries:
            local = oracle.search_flat(q, sr.tobytes() + b"\0", array.array("q", so.tolist()),
                                       array.array("i", sl.tolist()), M, 24, 11, 1, "sw", "score", 1)["score"]
            buf = gather.new_local()
            buf[:] = torch.tensor(local, dtype=torch.int32)
            rows.append(gather.gather(buf).tolist())
        if rank == 0:
            wr, wo, wl, _ = synthetic.config5_shard(0, 1, nb * world, block)
            ok = True
            for q, got in zip(queries, rows):
                want = oracle.search_flat(q, wr.tobytes() + b"\0", array.array("q", wo.tolist()),
                                          array.array("i", wl.tolist()), M, 24, 11, 1, "sw", "score", 4)["score"]
                ok = ok and got == want
            ret["ok"] = ok
    finally:
        dist.destroy_process_group()


def test_block_dealt_batch_matches_single_rank(oracle):
    import torch.multiprocessing as mp

    world = 2
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        ret = mgr.dict()
        port = _free_port()
        procs = [ctx.Process(target=_block_worker, args=(r, world, port, ret)) for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(180)
        assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
        assert ret["ok"]


def test_config5_blocks_are_balanced():
    lens = [synthetic.config5_block_lengths(b, 20_000) for b in range(8)]
    tot = [int(l.sum()) for l in lens]
    assert (max(tot) - min(tot)) / (sum(tot) / len(tot)) < 0.03
    r, l = synthetic.config5_block(3, 500)
    assert len(r) == int(l.sum()) and r.max() < 23 and l.min() >= 11
    assert np.array_equal(l, synthetic.config5_block_lengths(3, 500))
