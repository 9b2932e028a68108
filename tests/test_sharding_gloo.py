"""Multi-rank path on CPU: world_size 2 and 3 over gloo.

Each rank takes the shard the LIBRARY's deal gives it (`pyopal_cuda_deal`, the host code of csrc/multi.cu, which
needs no GPU), scores it -- here with the oracle, tests may --, the per-shard score arrays are all-gathered in
padded form and un-permuted with the same destination table the device kernel uses (slot j of rank r = database
index mine_r[j], -1 padding), and the result must equal the single-rank search.  The NCCL form of the same
exchange is exercised on hardware by tests/test_multi_gpu.py.
"""

import array
import os
import socket

import numpy as np
import pytest

import synthetic
from pyopal_b200 import sharding
from pyopal_b200.platform import cuda


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _gather_unpermute(dist, torch, local_scores, mine_all, rank, n_all):
    """What csrc/multi.cu does with ncclAllGather + unpermute_kernel, over gloo."""
    world = len(mine_all)
    n_pad = max(len(m) for m in mine_all)
    buf = torch.zeros(n_pad, dtype=torch.int32)
    buf[: len(local_scores)] = torch.tensor(local_scores, dtype=torch.int32)
    gathered = torch.empty(world * n_pad, dtype=torch.int32)
    dist.all_gather_into_tensor(gathered, buf)
    dest = np.full(world * n_pad, -1, dtype=np.int64)
    for r, m in enumerate(mine_all):
        dest[r * n_pad : r * n_pad + len(m)] = m
    out = np.full(n_all, np.iinfo(np.int32).min, dtype=np.int32)
    valid = dest >= 0
    out[dest[valid]] = gathered.numpy()[valid]
    return out.tolist()


def _worker(rank, world, port, ret):
    import torch
    import torch.distributed as dist

    import oracle
    import pyopal_b200 as pyopal

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        db = synthetic.database(900, 700, 4.5, 0.9, 0, 3000, giants=(9000, 7000, 5000))
        residues, offsets, lengths = db
        q = synthetic.random_residues(np.random.Generator(np.random.PCG64(901)), 60).tobytes()
        M = pyopal.Aligner("BLOSUM62")._int_matrix
        mine_all = [np.asarray(sharding.shard_indices(lengths, world, r)) for r in range(world)]
        sr, so, sl = synthetic.pack_subset(db, mine_all[rank])
        local = oracle.search_flat(q, sr.tobytes() + b"\0", array.array("q", so.tolist()), array.array("i", sl.tolist()),
                                   M, 24, 11, 1, "sw", "score", 1)["score"]
        got = _gather_unpermute(dist, torch, local, mine_all, rank, len(lengths))
        if rank == 0:
            want = oracle.search_flat(q, residues.tobytes() + b"\0", array.array("q", offsets.tolist()),
                                      array.array("i", lengths.tolist()), M, 24, 11, 1, "sw", "score", 4)["score"]
            ret["ok"] = got == want
            ret["residues"] = [int(lengths[m].sum()) for m in mine_all]
            ret["longest"] = int(lengths.max())
    finally:
        dist.destroy_process_group()


def _spawn(target, world, extra=()):
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        ret = mgr.dict()
        port = _free_port()
        procs = [ctx.Process(target=target, args=(r, world, port, ret) + tuple(extra)) for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(180)
        assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
        return dict(ret)


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_search_matches_single_rank(oracle, world):
    ret = _spawn(_worker, world)
    assert ret["ok"]
    # cell balance: snake dealing keeps every shard within one longest target of the others
    assert max(ret["residues"]) - min(ret["residues"]) <= ret["longest"]


def test_shards_partition_and_balance():
    rng = np.random.Generator(np.random.PCG64(5))
    lengths = synthetic.random_lengths(rng, 10_000, 5.65, 0.68, 2, 35_213)
    for world in (1, 2, 4, 8):
        owner = np.asarray(cuda.deal(lengths, world))
        parts = [np.asarray(sharding.shard_indices(lengths, world, r)) for r in range(world)]
        assert all(np.array_equal(np.nonzero(owner == r)[0], parts[r]) for r in range(world))
        allidx = np.concatenate(parts)
        assert len(allidx) == len(lengths) and len(np.unique(allidx)) == len(lengths)
        tot = [int(lengths[p].sum()) for p in parts]
        assert max(tot) - min(tot) <= int(lengths.max())
        assert all(np.all(np.diff(p) > 0) for p in parts if len(p) > 1)  # ascending: slices map to contiguous ranges
    # the deal is the snake over the length-sorted order (descending, ties by index)
    order = np.argsort(-lengths.astype(np.int64), kind="stable")
    owner = np.asarray(cuda.deal(lengths, 3))
    assert owner[order[:7]].tolist() == [0, 1, 2, 2, 1, 0, 0]


def test_pack_subset_roundtrip():
    residues, offsets, lengths = synthetic.database(7, 300, 4.0, 0.8, 0, 500)
    mine = np.asarray(sharding.shard_indices(lengths, 3, 1))
    sr, so, sl = synthetic.pack_subset((residues, offsets, lengths), mine)
    for k, i in enumerate(mine):
        assert bytes(sr[so[k] : so[k] + sl[k]]) == bytes(residues[offsets[i] : offsets[i] + lengths[i]])


def _block_worker(rank, world, port, ret):
    """Config-5 shape: blocks dealt round-robin (block b -> rank b mod world), batched queries."""
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
        # the index table ShardedDatabase.from_blocks hands to the library: block j of rank r is database block j * world + r
        mine_all = [np.concatenate([np.arange((j * world + r) * block, (j * world + r + 1) * block) for j in range(nb)])
                    for r in range(world)]
        rows = []
        for q in queries:
            local = oracle.search_flat(q, sr.tobytes() + b"\0", array.array("q", so.tolist()),
                                       array.array("i", sl.tolist()), M, 24, 11, 1, "sw", "score", 1)["score"]
            rows.append(_gather_unpermute(dist, torch, local, mine_all, rank, nb * world * block))
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
    assert _spawn(_block_worker, 2)["ok"]


def test_config5_blocks_are_balanced():
    lens = [synthetic.config5_block_lengths(b, 20_000) for b in range(8)]
    tot = [int(l.sum()) for l in lens]
    assert (max(tot) - min(tot)) / (sum(tot) / len(tot)) < 0.03
    r, l = synthetic.config5_block(3, 500)
    assert len(r) == int(l.sum()) and r.max() < 23 and l.min() >= 11
    assert np.array_equal(l, synthetic.config5_block_lengths(3, 500))


def test_nccl_id_file_rendezvous(tmp_path):
    # the 128 bytes the ranks agree on travel through a file when no launcher broadcast is at hand; rank 0 needs NCCL
    # (absent or without a GPU it raises RuntimeError, which is the library failing loudly, not a fallback)
    key = "t%d" % os.getpid()
    path = tmp_path / f"pyopal_nccl_id_{key}"
    path.write_bytes(bytes(range(128)))
    assert sharding.exchange_nccl_id(1, key=key, directory=str(tmp_path), timeout=2.0) == bytes(range(128))
    sharding.discard_nccl_id(key=key, directory=str(tmp_path))
    assert not path.exists()
    with pytest.raises(TimeoutError):
        sharding.exchange_nccl_id(1, key=key, directory=str(tmp_path), timeout=0.2)
