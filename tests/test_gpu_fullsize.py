"""GPU parity at BASELINE.json's full sizes (config 2: 570k Swiss-Prot-shaped proteins, 2.05e8 residues).

The oracle cannot score 1e11+ cells in seconds, so the full-size runs are
checked through size-independent properties: the oracle on a random subsample
plus every planted giant, ordering between the algorithms, planted homologs,
and agreement between the whole-database search and its slices / shards /
batched form.
"""

import array

import numpy as np
import pytest

import pyopal_b200 as pyopal
import synthetic
from pyopal_b200 import sharding
from pyopal_b200.platform import cuda as _cuda_mod

pytestmark = pytest.mark.gpu

GO, GE = 11, 1
ALGO = {"nw": 0, "hw": 1, "ov": 2, "sw": 3}


@pytest.fixture(scope="module")
def full(cuda):
    db = synthetic.config2()
    residues, offsets, lengths = db
    store = cuda.DeviceStore(residues, offsets, np.ascontiguousarray(lengths), 24, device=0)
    aligner = pyopal.Aligner("BLOSUM62", GO, GE)
    queries = synthetic.config2_queries(db)
    return {"db": db, "store": store, "M": aligner._int_matrix, "queries": [q.tobytes() for q in queries]}


def gpu_scores(full, q, algo, **kw):
    return np.frombuffer(full["store"].search(q, GO, GE, full["M"], 0, ALGO[algo], **kw)[0], dtype=np.int32)


def oracle_subset(oracle, full, q, algo, idx):
    residues, offsets, lengths = full["db"]
    sr, so, sl = synthetic.pack_subset(full["db"], np.asarray(idx))
    out = oracle.search_flat(q, sr.tobytes() + b"\0", array.array("q", so.tolist()), array.array("i", sl.tolist()),
                             full["M"], 24, GO, GE, algo, "score", 8)
    return np.asarray(out["score"], dtype=np.int32)


@pytest.mark.parametrize("qi,algo", [(0, "sw"), (1, "sw"), (2, "sw"), (3, "sw"), (4, "sw"), (1, "nw"), (1, "hw"), (1, "ov"),
                                     (4, "nw"), (4, "hw"), (4, "ov")])
def test_subsample_and_giants_match_oracle(full, oracle, qi, algo):
    # a random subsample + all planted giants for EVERY query of the benchmark step (SW: the
    # one-pass tiles of the 100 / 300-aa queries and the 7 / 14 / 35-pass bulk plans of the 1000 / 2000 / 5000-aa
    # ones, whose random targets never touch the giant kernels) and the 300- and 5000-aa queries for NW / HW / OV
    _, _, lengths = full["db"]
    rng = np.random.Generator(np.random.PCG64(99 + qi))
    giants = np.nonzero(lengths >= 10_000)[0]
    idx = np.unique(np.concatenate([rng.choice(len(lengths), 3000 if qi < 3 else 1200, replace=False), giants]))
    q = full["queries"][qi]
    got = gpu_scores(full, q, algo)
    assert len(got) == len(lengths)
    want = oracle_subset(oracle, full, q, algo, idx)
    bad = np.nonzero(got[idx] != want)[0]
    assert len(bad) == 0, f"{algo}: {len(bad)} mismatches, first {[(int(idx[b]), int(got[idx[b]]), int(want[b])) for b in bad[:5]]}"


def test_long_query_giants_overflow_ladder(full, oracle):
    # the giants against a long query (SW, NW, multi-pass tiles) and against the 100-aa query
    # (NW scores below -32768: the static 32-bit rung)
    _, _, lengths = full["db"]
    giants = np.nonzero(lengths >= 10_000)[0]
    for qi, algo in ((4, "sw"), (4, "nw"), (0, "nw"), (0, "hw")):
        q = full["queries"][qi]
        got = gpu_scores(full, q, algo)
        want = oracle_subset(oracle, full, q, algo, giants)
        assert np.array_equal(got[giants], want), (qi, algo)
        if (qi, algo) == (0, "nw"):
            assert got[giants].min() < -32768


def test_algorithm_ordering_full_size(full):
    # free end gaps can only help: SW >= OV >= HW >= NW for every target; SW >= 0
    q = full["queries"][0]
    sw, ov, hw, nw = (gpu_scores(full, q, a) for a in ("sw", "ov", "hw", "nw"))
    assert (sw >= 0).all()
    assert (sw >= ov).all() and (ov >= hw).all() and (hw >= nw).all()


def test_planted_homolog_is_top_hit(full):
    # an exact copy of a database entry scores at least its identity diagonal and is the best hit
    residues, offsets, lengths = full["db"]
    k = int(np.nonzero((lengths > 400) & (lengths < 600))[0][7])
    q = residues[offsets[k]: offsets[k] + lengths[k]].tobytes()
    M = np.asarray(full["M"]).reshape(24, 24)
    diag = int(sum(M[c, c] for c in q))
    sw = gpu_scores(full, q, "sw")
    assert sw[k] >= diag
    assert sw[k] == sw.max()
    assert gpu_scores(full, q, "nw")[k] == diag


def test_slices_reproduce_whole(full):
    q = full["queries"][1]
    whole = gpu_scores(full, q, "sw")
    n = len(whole)
    for s, e in [(0, 1000), (123_456, 234_567), (n - 777, n), (n // 2, n // 2 + 1)]:
        part = gpu_scores(full, q, "sw", start=s, end=e)
        assert np.array_equal(part, whole[s:e]), (s, e)


def test_four_way_shards_reproduce_whole(full, cuda):
    # the N-GPU partition (sharding.shard_indices) run shard by shard on one device
    residues, offsets, lengths = full["db"]
    q = full["queries"][2]
    whole = gpu_scores(full, q, "sw")
    out = np.full(len(lengths), np.iinfo(np.int32).min, dtype=np.int32)
    totals = []
    for rank in range(4):
        mine = np.asarray(sharding.shard_indices(lengths, 4, rank))
        sr, so, sl = synthetic.pack_subset((residues, offsets, lengths), mine)
        store = cuda.DeviceStore(sr, so, sl, 24, device=0)
        out[mine] = np.frombuffer(store.search(q, GO, GE, full["M"], 0, 3)[0], dtype=np.int32)
        totals.append(int(sl.sum()))
        store.close()
    assert np.array_equal(out, whole)
    assert max(totals) - min(totals) <= int(lengths.max())


def test_stream_kernel_equals_classic_on_a_shard(full, cuda, oracle, monkeypatch):
    # rank 0's 1/8 shard of config 2 (what one GPU of an 8-GPU box searches): the planner picks 32-lane tiles there
    # and hands them to the continuous-wavefront kernel (csrc/stream.cuh); its scores must equal the classic
    # kernel's for every query of the step (one and several passes), for NW on the longest query, and the oracle's
    # on a subsample
    db = full["db"]
    mine = np.asarray(sharding.shard_indices(db[2], 8, 0))
    sr, so, sl = synthetic.pack_subset(db, mine)
    rows = {}
    for env in ("PYOPAL_CUDA_NO_STREAM", "PYOPAL_CUDA_STREAM"):
        monkeypatch.delenv("PYOPAL_CUDA_NO_STREAM", raising=False)
        monkeypatch.delenv("PYOPAL_CUDA_STREAM", raising=False)
        monkeypatch.setenv(env, "1")
        store = cuda.DeviceStore(sr, so, np.ascontiguousarray(sl), 24, device=0)
        rows[env] = [np.frombuffer(store.search(q, GO, GE, full["M"], 0, 3)[0], dtype=np.int32).copy() for q in full["queries"]]
        rows[env].append(np.frombuffer(store.search(full["queries"][4], GO, GE, full["M"], 0, 0)[0], dtype=np.int32).copy())
        rows[env].append(np.frombuffer(store.search(full["queries"][2], GO, GE, full["M"], 0, 2)[0], dtype=np.int32).copy())
        store.close()
    for a, b in zip(rows["PYOPAL_CUDA_NO_STREAM"], rows["PYOPAL_CUDA_STREAM"]):
        assert np.array_equal(a, b)
    monkeypatch.delenv("PYOPAL_CUDA_STREAM", raising=False)
    store = cuda.DeviceStore(sr, so, np.ascontiguousarray(sl), 24, device=0)   # the default: the planner's choice
    rng = np.random.Generator(np.random.PCG64(31))
    idx = np.unique(np.concatenate([rng.choice(len(sl), 1500, replace=False), np.nonzero(sl >= 10_000)[0]]))
    sub = synthetic.pack_subset((sr, so, sl), idx)
    for qi in (2, 4):
        q = full["queries"][qi]
        got = np.frombuffer(store.search(q, GO, GE, full["M"], 0, 3)[0], dtype=np.int32)
        assert np.array_equal(got, rows["PYOPAL_CUDA_STREAM"][qi])
        want = oracle.search_flat(q, sub[0].tobytes() + b"\0", array.array("q", sub[1].tolist()), array.array("i", sub[2].tolist()),
                                  full["M"], 24, GO, GE, "sw", "score", 8)["score"]
        assert got[idx].tolist() == list(want), qi
    store.close()


def test_batch_equals_single_full_size(full):
    qs = full["queries"][:3]
    flat = np.frombuffer(full["store"].search_batch(qs, GO, GE, full["M"], 3), dtype=np.int32)
    n = full["store"].size
    for i, q in enumerate(qs):
        assert np.array_equal(flat[i * n: (i + 1) * n], gpu_scores(full, q, "sw")), i


def test_config5_block_batch(cuda, oracle):
    # config 5's shape on one rank: two blocks dealt to "rank 1 of 2", the batched search, the block gather's
    # view-based order (world 1 = identity) and the oracle on a subsample incl. the longest targets
    block = 60_000
    sr, so, sl, mine = synthetic.config5_shard(1, 2, 4, block)
    assert mine == [1, 3] and len(sl) == 2 * block
    store = cuda.DeviceStore(sr, so, np.ascontiguousarray(sl), 24, device=0)
    M = pyopal.Aligner("BLOSUM62", GO, GE)._int_matrix
    queries = [q.tobytes() for q in synthetic.config5_queries(6)]
    rows = np.frombuffer(store.search_batch(queries, GO, GE, M, 3), dtype=np.int32).reshape(len(queries), -1)
    rng = np.random.Generator(np.random.PCG64(55))
    idx = np.unique(np.concatenate([rng.choice(len(sl), 1500, replace=False), np.argsort(sl)[-20:], np.argsort(sl)[:20]]))
    sub_r, sub_o, sub_l = synthetic.pack_subset((sr, so, sl), idx)
    for qi, q in enumerate(queries):
        want = oracle.search_flat(q, sub_r.tobytes() + b"\0", array.array("q", sub_o.tolist()),
                                  array.array("i", sub_l.tolist()), M, 24, GO, GE, "sw", "score", 8)["score"]
        assert rows[qi][idx].tolist() == list(want), qi
        single = np.frombuffer(store.search(q, GO, GE, M, 0, 3)[0], dtype=np.int32)
        assert np.array_equal(single, rows[qi])


def test_dual_query_batch_equals_single_searches(cuda, oracle, monkeypatch):
    # config 5's shape again, the batch searched two queries at a time by dual_kernel (forced: this 120k-protein
    # stand-in is small enough to have "giants"): every score row equals the single-query search's, a subsample the
    # oracle's; through the one-GPU handle and through the multi-GPU handle's batch path (world 1)
    block = 60_000
    sr, so, sl, _ = synthetic.config5_shard(0, 2, 4, block)
    M = pyopal.Aligner("BLOSUM62", GO, GE)._int_matrix
    queries = [q.tobytes() for q in synthetic.config5_queries(7)]
    monkeypatch.setenv("PYOPAL_CUDA_NO_DUAL", "1")
    store = cuda.DeviceStore(sr, so, np.ascontiguousarray(sl), 24, device=0)
    want = np.frombuffer(store.search_batch(queries, GO, GE, M, 3), dtype=np.int32).reshape(len(queries), -1).copy()
    store.close()
    monkeypatch.delenv("PYOPAL_CUDA_NO_DUAL")
    monkeypatch.setenv("PYOPAL_CUDA_DUAL", "1")
    store = cuda.DeviceStore(sr, so, np.ascontiguousarray(sl), 24, device=0)
    got = np.frombuffer(store.search_batch(queries, GO, GE, M, 3), dtype=np.int32).reshape(len(queries), -1)
    assert np.array_equal(got, want)
    vals, idx = store.search_batch_topk(queries, GO, GE, M, 3, 50)
    for qi in range(len(queries)):
        order = np.lexsort((np.arange(want.shape[1]), -want[qi].astype(np.int64)))[:50]
        assert list(idx[qi * 50: qi * 50 + 50]) == order.tolist()
        assert list(vals[qi * 50: qi * 50 + 50]) == want[qi][order].tolist()
    store.close()
    sdb = sharding.ShardedDatabase.from_blocks((sr, so, np.ascontiguousarray(sl)), block, 2, device=0)
    flat = np.frombuffer(sdb.search_batch(queries, GO, GE, M, 3), dtype=np.int32).reshape(len(queries), -1)
    assert np.array_equal(flat, want)
    sdb.close()
    rng = np.random.Generator(np.random.PCG64(56))
    pick = np.unique(np.concatenate([rng.choice(len(sl), 1200, replace=False), np.argsort(sl)[-10:]]))
    sub = synthetic.pack_subset((sr, so, sl), pick)
    for qi in (0, 3, 6):
        ref = oracle.search_flat(queries[qi], sub[0].tobytes() + b"\0", array.array("q", sub[1].tolist()),
                                 array.array("i", sub[2].tolist()), M, 24, GO, GE, "sw", "score", 8)["score"]
        assert want[qi][pick].tolist() == list(ref)


def test_sharded_database_api_single_rank(cuda, oracle):
    # sharding.ShardedDatabase at world size 1 (no communicator): both ways of sharding, scores in database order,
    # the batched call and the batched device top-k; N > 1 is covered by test_multi_gpu.py (NCCL, when the box has
    # two GPUs) and tests/test_sharding_gloo.py (the deal and the gather's index arithmetic, on CPU)
    residues, offsets, lengths = synthetic.database(77, 3000, 5.0, 0.7, 1, 2500, giants=(6000, 4500))
    M = pyopal.Aligner("BLOSUM62", GO, GE)._int_matrix
    queries = [q.tobytes() for q in synthetic.config5_queries(3)]
    sdb = sharding.ShardedDatabase.deal(residues, offsets, lengths, device=0)
    assert sdb.world == 1 and sdb.n_all == len(lengths)
    want = [oracle.search_flat(q, residues.tobytes() + b"\0", array.array("q", offsets.tolist()),
                               array.array("i", lengths.tolist()), M, 24, GO, GE, "sw", "score", 8)["score"] for q in queries]
    for q, w in zip(queries, want):
        assert list(sdb.search(q, GO, GE, M, 0, 3)[0]) == list(w)
    flat = sdb.search_batch(queries, GO, GE, M, 3)
    n = len(lengths)
    for qi, w in enumerate(want):
        assert list(flat[qi * n: (qi + 1) * n]) == list(w)
    # the caller's own result buffer (a high-rate caller reuses one instead of allocating rows per call)
    mine = np.full(len(queries) * n, -1, dtype=np.int32)
    assert sdb.search_batch(queries, GO, GE, M, 3, out=mine) is mine
    assert mine.tolist() == list(flat)
    with pytest.raises(ValueError):
        sdb.search_batch(queries, GO, GE, M, 3, out=np.zeros(5, dtype=np.int32))
    vals, idx = sdb.search_batch_topk(queries, GO, GE, M, 3, 10)
    for qi, w in enumerate(want):
        order = sorted(range(n), key=lambda i: (-w[i], i))[:10]
        assert list(vals[qi * 10: qi * 10 + 10]) == [w[i] for i in order]
        assert list(idx[qi * 10: qi * 10 + 10]) == order
    part = sdb.search(queries[0], GO, GE, M, 0, 3, start=100, end=900)[0]
    assert list(part) == list(want[0][100:900])
    blk = sharding.ShardedDatabase.from_blocks((residues, offsets, lengths), 1500, 2, device=0)
    assert list(blk.search(queries[0], GO, GE, M, 0, 3)[0]) == list(want[0])
    with pytest.raises(ValueError):
        sharding.ShardedDatabase.from_blocks((residues, offsets, lengths), 1000, 2, device=0)


@pytest.mark.parametrize("qi,algo", [(2, "sw"), (3, "sw"), (3, "hw"), (2, "ov"), (1, "nw")])
def test_config4_top_hits_end_and_full(full, oracle, cuda, qi, algo):
    # BASELINE.json config 4 at full size: score pass over the 570k-protein database, the top hits (which include
    # the planted 10 000 - 35 213-aa giants: SW scores grow with length) re-aligned in end and full mode, every
    # score, end cell, start cell and operation string against the oracle.  1000- and 2000-aa queries make
    # pairdp.cu walk the giants as concurrent bands, four columns a step.
    residues, offsets, lengths = full["db"]
    q = full["queries"][qi]
    sc = gpu_scores(full, q, algo)
    k = 120
    top = np.sort(np.argsort(-sc.astype(np.int64), kind="stable")[:k])
    giants = np.nonzero(lengths >= 30_000)[0][:3]
    top = np.unique(np.concatenate([top, giants]))
    sr, so, sl = synthetic.pack_subset((residues, offsets, lengths), top)
    sub = cuda.DeviceStore(sr, so, sl, 24, device=0)
    flat = (sr.tobytes() + b"\0", array.array("q", so.tolist()), array.array("i", sl.tolist()))
    for mode, st in (("end", 1), ("full", 2)):
        got = sub.search(q, GO, GE, full["M"], st, ALGO[algo], lengths=sl)
        want = oracle.search_flat(q, *flat, full["M"], 24, GO, GE, algo, mode, 8)
        assert list(got[0]) == want["score"], (algo, mode)
        assert list(got[0]) == [int(x) for x in sc[top]]
        assert list(got[1]) == want["query_end"] and list(got[2]) == want["target_end"], (algo, mode)
        if st == 2:
            assert list(got[3]) == want["query_start"] and list(got[4]) == want["target_start"], (algo, mode)
            ops, oo, ol = got[5], got[6], got[7]
            for j in range(len(top)):
                assert bytes(ops[oo[j] : oo[j] + ol[j]]) == want["alignment"][j], (algo, mode, j)
    stats = sub.stats()
    if qi >= 2:  # queries of 1000+ aa against a handful of very long hits: the latency-bound shape was chosen
        assert stats["K"] in (4, 8) and stats["passes"] > 1


@pytest.mark.parametrize("qi,algo", [(1, "sw"), (3, "sw"), (3, "hw")])
def test_whole_database_end_mode(full, oracle, qi, algo):
    # end mode over all 570k proteins (the throughput shape of pairdp.cu: one pass of K = 10 rows per lane for the
    # 300-aa query, two concurrent bands of K = 32 for the 2000-aa one): scores equal the score-mode search
    # everywhere, end cells equal the oracle's on a random subsample plus every planted giant.
    residues, offsets, lengths = full["db"]
    q = full["queries"][qi]
    out = full["store"].search(q, GO, GE, full["M"], 1, ALGO[algo], lengths=lengths)
    got_s, got_q, got_t = (np.frombuffer(x, dtype=np.int32) for x in out[:3])
    assert np.array_equal(got_s, gpu_scores(full, q, algo))
    rng = np.random.Generator(np.random.PCG64(1234 + qi))
    giants = np.nonzero(lengths >= 10_000)[0]
    idx = np.unique(np.concatenate([rng.choice(len(lengths), 1500, replace=False), giants]))
    sr, so, sl = synthetic.pack_subset((residues, offsets, lengths), idx)
    want = oracle.search_flat(q, sr.tobytes() + b"\0", array.array("q", so.tolist()), array.array("i", sl.tolist()),
                              full["M"], 24, GO, GE, algo, "end", 8)
    assert got_s[idx].tolist() == want["score"]
    assert got_q[idx].tolist() == want["query_end"]
    assert got_t[idx].tolist() == want["target_end"]
