"""Several GPUs behind the reference API (csrc/multi.cu): needs a box with at least two B200s, skipped otherwise.

* `pyopal.align` / `Aligner.align` on a Database large enough to be dealt over every visible GPU must return the
  single-GPU result (the reference's own "chunked == whole" check, src/pyopal/tests/test_align.py:9-37);
* the one-process-per-GPU layout: two spawned ranks, ncclCommInitRank through the file rendezvous, gathered
  scores == single-GPU scores == oracle.
"""

import array
import os

import numpy as np
import pytest

import pyopal_b200 as pyopal
import synthetic
from pyopal_b200 import sharding

pytestmark = pytest.mark.gpu
GO, GE = 11, 1


def _need_two(cuda):
    if cuda.device_count() < 2:
        pytest.skip("needs two GPUs")


def test_align_uses_every_gpu_and_matches_one(cuda, oracle, monkeypatch):
    _need_two(cuda)
    db = synthetic.database(4242, 60_000, 5.3, 0.7, 0, 6000, giants=(20_000, 9_000))
    residues, offsets, lengths = db
    aligner = pyopal.Aligner("BLOSUM62", GO, GE)
    q = synthetic.random_residues(np.random.Generator(np.random.PCG64(1)), 240).tobytes()
    qs = aligner.alphabet.decode(q)
    monkeypatch.setenv("PYOPAL_CUDA_DEVICES", "0")
    one = pyopal.Database.from_packed(residues, lengths)
    want = {a: list(aligner.align_scores(qs, one, algorithm=a)) for a in ("sw", "nw", "hw", "ov")}
    assert type(one._device_store()).__name__ == "DeviceStore"
    monkeypatch.setenv("PYOPAL_CUDA_DEVICES", "all")
    many = pyopal.Database.from_packed(residues, lengths)
    store = many._device_store()
    assert type(store).__name__ == "MultiDeviceStore" and store.world == cuda.device_count()
    for a in ("sw", "nw", "hw", "ov"):
        assert list(aligner.align_scores(qs, many, algorithm=a)) == want[a], a
    # the reference-facing generator, result objects, slices
    hits = list(pyopal.align(qs, many, "BLOSUM62", gap_open=GO, gap_extend=GE, threads=4))
    assert [h.score for h in hits] == want["sw"] and [h.target_index for h in hits] == list(range(len(lengths)))
    part = aligner.align(qs, many, start=1234, end=40_321)
    assert [h.score for h in part] == want["sw"][1234:40_321] and part[0].target_index == 1234
    # oracle on a subsample
    idx = np.unique(np.concatenate([np.random.Generator(np.random.PCG64(2)).choice(len(lengths), 1500, replace=False),
                                    np.nonzero(lengths >= 9000)[0]]))
    sr, so, sl = synthetic.pack_subset(db, idx)
    ref = oracle.search_flat(q, sr.tobytes() + b"\0", array.array("q", so.tolist()), array.array("i", sl.tolist()),
                             aligner._int_matrix, 24, GO, GE, "sw", "score", 8)["score"]
    assert [want["sw"][i] for i in idx] == list(ref)
    # end and full mode go shard by shard, results scattered back to database order
    sub_idx = list(range(0, 60_000, 97))
    sub = many.extract(sub_idx)
    monkeypatch.setenv("PYOPAL_CUDA_DEVICES", "0")
    sub_one = one.extract(sub_idx)
    for mode in ("end", "full"):
        monkeypatch.setenv("PYOPAL_CUDA_DEVICES", "all")
        a = aligner.align(qs, sub, mode=mode)
        monkeypatch.setenv("PYOPAL_CUDA_DEVICES", "0")
        b = aligner.align(qs, sub_one, mode=mode)
        assert a == b, mode
    # batched calls
    queries = [aligner.alphabet.decode(x.tobytes()) for x in synthetic.config5_queries(5)]
    monkeypatch.setenv("PYOPAL_CUDA_DEVICES", "all")
    rows_many = aligner.align_batch(queries, many)
    top_many = aligner.align_batch_topk(queries, many, 25)
    rows_one = aligner.align_batch(queries, one)
    assert [list(r) for r in rows_many] == [list(r) for r in rows_one]
    for r, t in zip(rows_one, top_many):
        order = sorted(range(len(r)), key=lambda i: (-r[i], i))[:25]
        assert [(h.target_index, h.score) for h in t] == [(i, r[i]) for i in order]


def _queries():
    # lengths 551, 851, 63, 270 plus two truncated copies (800, 500): sorted by length the batch is two pairs within
    # 0.7 of each other and two single queries (csrc/api.cu batch_units)
    base = [q.tobytes() for q in synthetic.config5_queries(4)]
    return base + [base[1][:800], base[0][:500]]


def _rank(rank, world, port, ret):
    os.environ["MASTER_PORT"] = str(port)
    os.environ["PYOPAL_CUDA_DEVICE"] = str(rank)
    from pyopal_b200 import sharding as sh

    db = synthetic.database(515, 30_000, 5.3, 0.7, 0, 5000, giants=(15_000,))
    residues, offsets, lengths = db
    M = pyopal.Aligner("BLOSUM62", GO, GE)._int_matrix
    nid = sh.exchange_nccl_id(rank)
    sdb = sh.ShardedDatabase.deal(residues, offsets, np.ascontiguousarray(lengths), 24, rank=rank, world=world, nccl_id=nid, device=rank)
    queries = _queries()
    rows = sdb.search_batch(queries, GO, GE, M, 3)
    one = sdb.search(queries[1], GO, GE, M, 0, 1)[0]  # HW, the blocking call
    vals, idx = sdb.search_batch_topk(queries, GO, GE, M, 3, 20)
    sdb.close()
    # the same batch with the two-queries-per-pass kernel forced (this database is small enough to be searched one
    # query at a time by default): both rows of a dual search travel in one all-gather
    os.environ["PYOPAL_CUDA_DUAL"] = "1"
    nid2 = sh.exchange_nccl_id(rank, key=f"{port}_dual")
    sdb2 = sh.ShardedDatabase.deal(residues, offsets, np.ascontiguousarray(lengths), 24, rank=rank, world=world, nccl_id=nid2, device=rank)
    rows2 = sdb2.search_batch(queries, GO, GE, M, 3)
    vals2, idx2 = sdb2.search_batch_topk(queries, GO, GE, M, 3, 20)
    sdb2.search_batch([queries[1], queries[4]], GO, GE, M, 3, to_host=False)   # one unit: a pair
    dual_used = sdb2.stats()["profile_in_smem"] == 2
    sdb2.close()
    ret[rank] = (bytes(rows), bytes(one), bytes(vals), bytes(idx), bytes(rows2) == bytes(rows) and bytes(idx2) == bytes(idx), dual_used)


def test_two_nccl_ranks_match_single_gpu(cuda, oracle):
    _need_two(cuda)
    import multiprocessing as mp
    import socket

    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        ret = mgr.dict()
        procs = [ctx.Process(target=_rank, args=(r, 2, port, ret)) for r in range(2)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(600)
        assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
        got = dict(ret)
    sharding.discard_nccl_id(key=f"{port}_{os.getpid()}")
    assert got[0] == got[1]  # every rank holds the same database-order result
    assert got[0][4] and got[1][4]   # two queries per pass == one per pass, on both ranks
    assert got[0][5] and got[1][5]   # ... and the dual kernel did run
    sharding.discard_nccl_id(key=f"{port}_dual")
    db = synthetic.database(515, 30_000, 5.3, 0.7, 0, 5000, giants=(15_000,))
    residues, offsets, lengths = db
    M = pyopal.Aligner("BLOSUM62", GO, GE)._int_matrix
    store = cuda.DeviceStore(residues, offsets, np.ascontiguousarray(lengths), 24, device=0)
    queries = _queries()
    rows = np.frombuffer(got[0][0], dtype=np.int32).reshape(len(queries), -1)
    for qi, q in enumerate(queries):
        assert np.array_equal(rows[qi], np.frombuffer(store.search(q, GO, GE, M, 0, 3)[0], dtype=np.int32)), qi
    assert np.array_equal(np.frombuffer(got[0][1], dtype=np.int32), np.frombuffer(store.search(queries[1], GO, GE, M, 0, 1)[0], dtype=np.int32))
    vals = np.frombuffer(got[0][2], dtype=np.int32).reshape(len(queries), 20)
    idx = np.frombuffer(got[0][3], dtype=np.int32).reshape(len(queries), 20)
    for qi in range(len(queries)):
        order = sorted(range(rows.shape[1]), key=lambda i: (-int(rows[qi][i]), i))[:20]
        assert idx[qi].tolist() == order and vals[qi].tolist() == [int(rows[qi][i]) for i in order]
    want = oracle.search_flat(queries[0], residues.tobytes() + b"\0", array.array("q", offsets.tolist()),
                              array.array("i", lengths.tolist()), M, 24, GO, GE, "sw", "score", 8)["score"]
    assert rows[0].tolist() == list(want)
