"""GPU parity: score mode through the C-ABI vs the oracle (bit-exact)."""

import array

import numpy as np
import pytest

import pyopal_b200 as pyopal
import synthetic

pytestmark = pytest.mark.gpu

ALGOS = ("sw", "nw", "hw", "ov")


def make_db(seqs, alphabet=None):
    db = pyopal.Database.__new__(pyopal.Database)
    db.alphabet = alphabet or pyopal.Alphabet()
    db._sequences = list(seqs)
    db._lengths = [len(s) for s in seqs]
    return db


def rand_seqs(seed, n, mu, sigma, lo, hi):
    return synthetic.as_sequences(synthetic.database(seed, n, mu, sigma, lo, hi))


def rand_query(seed, n):
    return synthetic.random_residues(np.random.Generator(np.random.PCG64(seed)), n).tobytes()


def check(oracle, aligner, query, seqs, algo, db=None, **kw):
    db = db or make_db(seqs, aligner.alphabet)
    got = list(aligner.align_scores(query_str(aligner, query), db, algorithm=algo, **kw))
    s, e = kw.get("start", 0), min(kw.get("end", len(seqs)), len(seqs))
    want = oracle.search(query, seqs[s:e], aligner._int_matrix, len(aligner.alphabet), aligner.gap_open,
                         aligner.gap_extend, algo, "score", threads=8)["score"]
    assert len(got) == len(want)
    bad = [i for i, (a, b) in enumerate(zip(got, want)) if a != b]
    assert not bad, f"{algo}: {len(bad)} mismatches, first {[(i, got[i], want[i], len(seqs[s + i])) for i in bad[:5]]}"
    return db


def query_str(aligner, encoded):
    return aligner.alphabet.decode(encoded)


def test_golden_vectors_through_api(cuda):
    # /root/reference/src/pyopal/tests/test_aligner.py:42-60,93-111 and _align.py:106-111
    aligner = pyopal.Aligner()
    db = pyopal.Database(["AACCGCTG"])
    assert aligner.align("ACCTCG", db, algorithm="nw")[0].score == 44
    assert aligner.align("ACCTCG", db, algorithm="sw")[0].score == 47
    assert isinstance(aligner.align("ACCTCG", db)[0], pyopal.ScoreResult)
    targets = ["AACCGCTG", "ATGCGCT", "TTATTACG"]
    res = list(pyopal.align("ACCTG", targets, gap_open=2, ordered=True))
    assert [r.score for r in res] == [41, 31, 23]
    assert [r.target_index for r in res] == [0, 1, 2]


@pytest.mark.parametrize("algo", ALGOS)
@pytest.mark.parametrize("lq", [1, 7, 33, 100, 300, 517])
def test_random_db(cuda, oracle, algo, lq):
    seqs = rand_seqs(100 + lq, 1500, 5.0, 0.7, 1, 4000)
    aligner = pyopal.Aligner("BLOSUM62", 11, 1)
    check(oracle, aligner, rand_query(lq, lq), seqs, algo)


@pytest.mark.parametrize("algo", ALGOS)
def test_default_scoring_low_gap_costs(cuda, oracle, algo):
    # reference defaults BLOSUM50 / 3 / 1 (lib.pyx:1153-1155): higher scores, stresses the ladder
    seqs = rand_seqs(21, 800, 5.3, 0.6, 1, 3000)
    check(oracle, pyopal.Aligner(), rand_query(22, 250), seqs, algo)


@pytest.mark.parametrize("algo", ALGOS)
def test_multi_pass_long_query(cuda, oracle, algo):
    # 1500 and 2900 rows need 2 and 3 passes of the widest tile (32 threads x 40 rows)
    seqs = rand_seqs(31, 300, 5.5, 0.8, 1, 6000)
    aligner = pyopal.Aligner("BLOSUM62", 11, 1)
    for lq in (1500, 2900):
        check(oracle, aligner, rand_query(lq, lq), seqs, algo)


@pytest.mark.parametrize("algo", ALGOS)
def test_empty_and_tiny_targets(cuda, oracle, algo):
    seqs = [b"", b"\x00", b"\x01\x02", b"", rand_query(1, 50), b"\x05" * 3, b""] + rand_seqs(41, 40, 2.0, 0.8, 1, 30)
    aligner = pyopal.Aligner("BLOSUM62", 11, 1)
    for lq in (1, 5, 64):
        check(oracle, aligner, rand_query(40 + lq, lq), seqs, algo)


def test_sw_overflow_escalation(cuda, oracle):
    # query = long mutated copies of targets => SW scores above what the biased u16 rung holds
    # (~65000): 16 -> 32 bit rescore; base[1] scores between 32768 and 65535 and stays at 16 bit
    rng = np.random.Generator(np.random.PCG64(51))
    base = [synthetic.random_residues(rng, n) for n in (24000, 7000, 8000)]
    seqs = [b.tobytes() for b in base] + [synthetic.mutate(rng, base[0]).tobytes()] + rand_seqs(52, 200, 5.5, 0.6, 10, 2000)
    aligner = pyopal.Aligner("BLOSUM62", 11, 1)
    query = np.concatenate([base[0][:22000], base[1]]).tobytes()
    db = check(oracle, aligner, query, seqs, "sw")
    st = db._device_store().stats()
    assert st["escalated"] >= 2, st
    got = list(aligner.align_scores(query_str(aligner, query), db))
    assert max(got) > 65535
    assert 32767 < got[1] < 65535


@pytest.mark.parametrize("algo", ALGOS)
@pytest.mark.parametrize("lq", [50, 700, 2500, 4500])
def test_giant_teams(cuda, oracle, algo, lq):
    # a few long targets in a small database go to giant.cu: one warp (Lq 50), a team of 6 warps with
    # 4 rows per lane (700), 8 rows (2500), 16 rows (4500); lengths not multiples of the 4-column block
    seqs = synthetic.as_sequences(synthetic.database(71, 400, 5.0, 0.6, 1, 1500, giants=(5003, 4001, 3002, 2999)))
    aligner = pyopal.Aligner("BLOSUM62", 11, 1)
    db = check(oracle, aligner, rand_query(70 + lq, lq), seqs, algo)
    assert db._device_store().stats()["targets32"] >= 4


@pytest.mark.parametrize("lq", [50, 700, 2500, 4500])
def test_giant_teams_one_target_per_team(cuda, oracle, monkeypatch, lq):
    # Smith-Waterman giants are walked two to a team at 16 bit by default (the two targets of a pair);
    # PYOPAL_CUDA_NO_GIANT16 keeps the 32-bit, one-target-per-team kernel that NW / HW / OV always use
    monkeypatch.setenv("PYOPAL_CUDA_NO_GIANT16", "1")
    seqs = synthetic.as_sequences(synthetic.database(71, 400, 5.0, 0.6, 1, 1500, giants=(5003, 4001, 3002, 2999)))
    aligner = pyopal.Aligner("BLOSUM62", 11, 1)
    db = check(oracle, aligner, rand_query(70 + lq, lq), seqs, "sw")
    assert db._device_store().stats()["targets32"] >= 4


def test_packed_giants_overflow_is_rescored(cuda, oracle):
    # a pair of long targets one of which is a near copy of the query under a matrix with a 100-point match: its
    # packed 16-bit score overflows inside giant.cu, both targets of the pair go to the 32-bit rescue list, scores
    # stay exact
    rng = np.random.Generator(np.random.PCG64(4321))
    base = pyopal.ScoringMatrix.from_name("BLOSUM62")
    rows = base.matrix
    w = base.alphabet.index("W")
    rows[w][w] = 100
    aligner = pyopal.Aligner(pyopal.ScoringMatrix(rows, base.alphabet), 11, 1)
    seqs = synthetic.as_sequences(synthetic.database(72, 300, 5.0, 0.6, 1, 1200))
    twin = np.full(3000, w, dtype=np.uint8)
    twin[::97] = 0
    seqs += [twin.tobytes(), synthetic.random_residues(rng, 2990).tobytes()]
    query = bytes([w]) * 900
    db = check(oracle, aligner, query, seqs, "sw")
    got = list(aligner.align_scores(query_str(aligner, query), db))
    assert max(got) > 65535
    st = db._device_store().stats()
    assert st["targets32"] >= 2 and st["escalated"] >= 1, st


@pytest.mark.parametrize("algo", ALGOS)
@pytest.mark.parametrize("shape", ["1,16,0", "2,16,0", "3,32,0", "4,32,0", "8,16,0", "4,32"])
@pytest.mark.parametrize("lq", [1, 37, 300, 1100])
def test_transposed_giants(cuda, oracle, monkeypatch, algo, shape, lq):
    # tgiant.cu: the target over the lanes of a team of 1..8 warps (16 / 32 positions per lane), the
    # query streamed; lengths around the slab sizes (512 .. 4096) so that one-slab, multi-slab
    # (boundary FIFO), partly idle teams and the warp-to-warp ring are all taken
    monkeypatch.setenv("PYOPAL_CUDA_TGIANT", shape)
    giants = (9001, 4097, 4096, 4095, 3073, 2049, 2048, 1537, 1100, 1025)
    seqs = synthetic.as_sequences(synthetic.database(81, 300, 5.0, 0.6, 1, 1000, giants=giants))
    aligner = pyopal.Aligner("BLOSUM62", 11, 1)
    db = check(oracle, aligner, rand_query(80 + lq, lq), seqs, algo)
    assert db._device_store().stats()["targets32"] >= len(giants)


def test_transposed_giants_match_query_teams(cuda, oracle, monkeypatch):
    # the two giant kernels (query over the lanes / target over the lanes) agree on a high-scoring homolog
    rng = np.random.Generator(np.random.PCG64(83))
    big = synthetic.random_residues(rng, 12000)
    seqs = [big.tobytes(), synthetic.mutate(rng, big).tobytes()] + rand_seqs(84, 200, 5.0, 0.6, 1, 900)
    aligner = pyopal.Aligner("BLOSUM50", 10, 2)
    q = big[3000:5200].tobytes()
    for algo in ALGOS:
        monkeypatch.delenv("PYOPAL_CUDA_NO_TGIANT", raising=False)
        a = list(aligner.align_scores(query_str(aligner, q), make_db(seqs, aligner.alphabet), algorithm=algo))
        monkeypatch.setenv("PYOPAL_CUDA_NO_TGIANT", "1")
        b = list(aligner.align_scores(query_str(aligner, q), make_db(seqs, aligner.alphabet), algorithm=algo))
        assert a == b, algo
    monkeypatch.delenv("PYOPAL_CUDA_NO_TGIANT", raising=False)
    check(oracle, aligner, q, seqs, "sw")


@pytest.mark.parametrize("algo", ("nw", "hw", "ov"))
def test_static_32bit_rung(cuda, oracle, algo):
    # NW of a short query against 30k+ targets goes below -32768; long identical pairs go above +32767
    rng = np.random.Generator(np.random.PCG64(61))
    big = synthetic.random_residues(rng, 33000)
    seqs = [big.tobytes(), synthetic.random_residues(rng, 31000).tobytes(), big[:6000].tobytes()] + rand_seqs(62, 100, 5.5, 0.6, 10, 2000)
    aligner = pyopal.Aligner("BLOSUM62", 11, 1)
    db = check(oracle, aligner, rand_query(63, 120), seqs, algo)
    assert db._device_store().stats()["targets32"] >= 2
    check(oracle, aligner, big[:5000].tobytes(), seqs, algo)


def test_slices(cuda, oracle):
    seqs = rand_seqs(71, 501, 5.0, 0.6, 1, 1500)
    aligner = pyopal.Aligner("BLOSUM62", 11, 1)
    q = rand_query(72, 90)
    db = make_db(seqs, aligner.alphabet)
    for s, e in ((0, 10), (3, 4), (250, 501), (100, 10_000), (7, 7)):
        for algo in ("sw", "nw"):
            check(oracle, aligner, q, seqs, algo, db=db, start=s, end=e)
    res = aligner.align(query_str(aligner, q), db, start=5, end=9)
    assert [r.target_index for r in res] == [5, 6, 7, 8]
    with pytest.raises(IndexError):
        aligner.align(query_str(aligner, q), db, start=9, end=5)


def test_wide_scores_and_gaps(cuda, oracle):
    # scores / gap costs outside the int8 profile: everything runs on the 32-bit wide-profile kernel
    rows = [[(37 * i + 11 * j) % 400 - 150 if i != j else 300 for j in range(24)] for i in range(24)]
    for i in range(24):
        for j in range(i):
            rows[i][j] = rows[j][i]
    m = pyopal.ScoringMatrix(rows)
    aligner = pyopal.Aligner(m, 200, 30)
    seqs = rand_seqs(81, 200, 4.5, 0.6, 1, 800)
    for algo in ALGOS:
        check(oracle, aligner, rand_query(82, 77), seqs, algo)
        check(oracle, aligner, rand_query(83, 1300), seqs, algo)


def test_small_alphabet(cuda, oracle):
    alphabet = pyopal.Alphabet("ACGT")
    m = pyopal.ScoringMatrix([[5 if i == j else -4 for j in range(4)] for i in range(4)], "ACGT")
    aligner = pyopal.Aligner(m, 10, 1)
    rng = np.random.Generator(np.random.PCG64(91))
    seqs = [rng.integers(0, 4, size=int(n), dtype=np.uint8).tobytes() for n in rng.integers(1, 900, size=300)]
    q = rng.integers(0, 4, size=200, dtype=np.uint8).tobytes()
    for algo in ALGOS:
        check(oracle, aligner, q, seqs, algo, db=make_db(seqs, alphabet))


def test_generic_entry_point_any_basedatabase(cuda, oracle):
    """A foreign BaseDatabase subclass goes through opalSearchDatabase's own signature (opal.pxd:40-54)."""

    class ListDatabase(pyopal.BaseDatabase):
        def __init__(self, seqs, alphabet=None):
            super().__init__(alphabet=alphabet)
            self._seqs = list(seqs)

        def get_sequences(self):
            return self._seqs

        def get_lengths(self):
            return [len(s) for s in self._seqs]

        def get_size(self):
            return len(self._seqs)

    seqs = rand_seqs(95, 64, 4.5, 0.6, 1, 600)
    aligner = pyopal.Aligner("BLOSUM62", 11, 1)
    q = rand_query(96, 80)
    db = ListDatabase(seqs, aligner.alphabet)
    for algo in ALGOS:
        res = aligner.align(query_str(aligner, q), db, algorithm=algo)
        want = oracle.search(q, seqs, aligner._int_matrix, 24, 11, 1, algo, "score")["score"]
        assert [r.score for r in res] == want
        assert [r.target_index for r in res] == list(range(len(seqs)))


def test_batch_matches_single(cuda, oracle):
    seqs = rand_seqs(97, 700, 5.0, 0.6, 1, 1500)
    aligner = pyopal.Aligner("BLOSUM62", 11, 1)
    db = make_db(seqs, aligner.alphabet)
    queries = [rand_query(200 + i, n) for i, n in enumerate((50, 300, 1400, 9))]
    rows = aligner.align_batch([query_str(aligner, q) for q in queries], db)
    for q, row in zip(queries, rows):
        assert list(row) == oracle.search(q, seqs, aligner._int_matrix, 24, 11, 1, "sw", "score", threads=8)["score"]


def test_mutation_invalidates_device_store(cuda):
    aligner = pyopal.Aligner()
    db = pyopal.Database(["AACCGCTG"])
    assert aligner.align("ACCTCG", db)[0].score == 47
    db.append("ACCTCG")
    res = aligner.align("ACCTCG", db)
    assert len(res) == 2 and res[1].score == 5 + 13 + 13 + 5 + 13 + 8
    db[0] = "TTTT"
    assert aligner.align("ACCTCG", db)[0].score == 5


def test_concurrent_aligners_share_a_database(cuda, oracle):
    # the reference's documented pattern (README.md:116-144): a ThreadPool of queries against ONE Database;
    # every call leases its own workspace (streams, parameter block, buffers) from the handle's pool, results must
    # equal the sequential ones
    from multiprocessing.pool import ThreadPool

    seqs = rand_seqs(301, 900, 5.0, 0.6, 1, 1500) + [rand_query(302, 5000)]
    aligner = pyopal.Aligner("BLOSUM62", 11, 1)
    db = make_db(seqs, aligner.alphabet)
    queries = [rand_query(310 + i, n) for i, n in enumerate((40, 90, 150, 260, 400, 700, 33, 1200))]
    algos = ["sw", "nw", "hw", "ov", "sw", "sw", "ov", "nw"]

    def one(args):
        q, algo = args
        return list(aligner.align_scores(query_str(aligner, q), db, algorithm=algo))

    with ThreadPool(4) as pool:
        got = pool.map(one, list(zip(queries, algos)) * 3)
    want = [oracle.search(q, seqs, aligner._int_matrix, 24, 11, 1, a, "score", threads=8)["score"] for q, a in zip(queries, algos)]
    for k, row in enumerate(got):
        assert row == want[k % len(queries)], (k, algos[k % len(queries)])
    # result objects from threads as well (pyopal.align(threads=...) maps onto the same call)
    res = list(pyopal.align(query_str(aligner, queries[2]), db, "BLOSUM62", gap_open=11, gap_extend=1, algorithm=algos[2], threads=3, ordered=True))
    assert [r.score for r in res] == want[2]


def test_concurrent_searches_overlap(cuda, oracle):
    # searches from several host threads on ONE handle overlap on the device (a pool of per-search workspaces, no
    # mutex around the search: SURVEY.md 8(b) "Threading").  The workload is latency-bound on purpose - one
    # 30 000-aa target walked by a single team of warps while the other SMs idle - so four threads must finish
    # their 24 searches in well under the time 24 solo searches take.
    import time
    from multiprocessing.pool import ThreadPool

    seqs = rand_seqs(331, 40, 5.0, 0.5, 1, 600) + [rand_query(332, 30_000)]
    aligner = pyopal.Aligner("BLOSUM62", 11, 1)
    db = make_db(seqs, aligner.alphabet)
    queries = [rand_query(340 + i, 1000) for i in range(4)]
    want = [oracle.search(q, seqs, aligner._int_matrix, 24, 11, 1, "sw", "score", threads=8)["score"] for q in queries]

    def one(i):
        return list(aligner.align_scores(query_str(aligner, queries[i % 4]), db))

    for i in range(4):
        assert one(i) == want[i]          # builds the store, warms the workspaces
    with ThreadPool(4) as pool:
        pool.map(one, range(8))           # ... of every thread
        t0 = time.perf_counter()
        solo = [one(i) for i in range(24)]
        t_solo = time.perf_counter() - t0
        t0 = time.perf_counter()
        both = pool.map(one, range(24))
        t_pool = time.perf_counter() - t0
    assert all(r == want[i % 4] for i, r in enumerate(solo)) and all(r == want[i % 4] for i, r in enumerate(both))
    assert t_pool < 0.85 * t_solo, (t_pool, t_solo)


def _expected_topk(scores, k):
    order = sorted(range(len(scores)), key=lambda i: (-scores[i], i))[:k]
    return [scores[i] for i in order], order


@pytest.mark.parametrize("algo", ALGOS)
def test_topk_matches_sorted_oracle(cuda, oracle, algo):
    # csrc/topk.cu: exact radix select on the device; ties go to the lower database index.  Short queries give
    # many equal scores, NW/HW give negative ones (the key is order-preserving over the whole int32 range)
    seqs = rand_seqs(401, 5000, 4.5, 0.7, 0, 900) + [b""]
    aligner = pyopal.Aligner("BLOSUM62", 11, 1)
    db = make_db(seqs, aligner.alphabet)
    for lq in (7, 120):
        q = rand_query(410 + lq, lq)
        want = oracle.search(q, seqs, aligner._int_matrix, 24, 11, 1, algo, "score", threads=8)["score"]
        for k in (1, 10, 257, 5000, len(seqs), len(seqs) + 5):
            hits = aligner.align_topk(query_str(aligner, q), db, k, algorithm=algo)
            ws, wi = _expected_topk(want, k)
            assert [h.score for h in hits] == ws, (algo, lq, k)
            assert [h.target_index for h in hits] == wi, (algo, lq, k)
    # slices report absolute indices
    q = rand_query(499, 60)
    want = oracle.search(q, seqs, aligner._int_matrix, 24, 11, 1, algo, "score", threads=8)["score"]
    hits = aligner.align_topk(query_str(aligner, q), db, 25, algorithm=algo, start=1000, end=3000)
    ws, wi = _expected_topk(want[1000:3000], 25)
    assert [h.score for h in hits] == ws and [h.target_index for h in hits] == [i + 1000 for i in wi]
    assert aligner.align_topk(query_str(aligner, q), db, 0) == []


def test_topk_all_equal_and_large(cuda):
    # every target identical: the k lowest indices; and a 2e6-entry device array through the device entry
    aligner = pyopal.Aligner("BLOSUM62", 11, 1)
    db = make_db([rand_query(5, 50)] * 3000, aligner.alphabet)
    hits = aligner.align_topk(query_str(aligner, rand_query(5, 50)), db, 40)
    assert [h.target_index for h in hits] == list(range(40)) and len({h.score for h in hits}) == 1
    import torch

    g = torch.Generator().manual_seed(3)
    scores = torch.randint(-70000, 70000, (2_000_003,), generator=g, dtype=torch.int32)
    scores[::7] = 123  # a fat tie class
    d = scores.cuda()
    store = db._device_store()
    for k in (1, 1000, 300_000):
        vals = torch.empty(k, dtype=torch.int32, device="cuda")
        idx = torch.empty(k, dtype=torch.int32, device="cuda")
        store.topk_device(d.data_ptr(), d.numel(), k, vals.data_ptr(), idx.data_ptr(), torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        v, i = vals.cpu().numpy(), idx.cpu().numpy()
        order = np.lexsort((i, -v.astype(np.int64)))
        ref = np.lexsort((np.arange(scores.numel()), -scores.numpy().astype(np.int64)))[:k]
        assert np.array_equal(i[order], ref) and np.array_equal(v[order], scores.numpy()[ref])
