"""Synthetic protein databases and queries for the benchmark and the parity tests.

Shapes follow BASELINE.json `configs` / SURVEY.md 8(d): residues i.i.d. from the
Swiss-Prot background composition (plus B, Z, X at 1e-4 each), lengths
round(lognormal(mu, sigma)) clipped, `numpy.random.Generator(PCG64(seed))`.
Sequences are produced directly in encoded form (alphabet indices of
``ARNDCQEGHILKMFPSTWYVBZX*``) as one packed uint8 buffer + offsets + lengths.
"""

import numpy as np

ALPHABET = "ARNDCQEGHILKMFPSTWYVBZX*"
_BG = {
    "L": 9.65, "A": 8.25, "G": 7.07, "V": 6.86, "E": 6.72, "S": 6.65, "I": 5.91, "K": 5.80, "R": 5.53,
    "D": 5.46, "T": 5.36, "P": 4.74, "N": 4.06, "Q": 3.93, "F": 3.86, "Y": 2.92, "M": 2.41, "H": 2.27,
    "C": 1.38, "W": 1.10,
}


def _probabilities():
    p = np.zeros(len(ALPHABET))
    for c, v in _BG.items():
        p[ALPHABET.index(c)] = v
    p = p / p.sum() * (1 - 3e-4)
    for c in "BZX":
        p[ALPHABET.index(c)] = 1e-4
    return p / p.sum()


def random_residues(rng, n):
    return rng.choice(len(ALPHABET), size=int(n), p=_probabilities()).astype(np.uint8)


def random_lengths(rng, n, mu, sigma, lo, hi):
    return np.clip(np.rint(rng.lognormal(mu, sigma, size=int(n))), lo, hi).astype(np.int32)


def pack(lengths, residues):
    offsets = np.zeros(len(lengths), dtype=np.int64)
    if len(lengths) > 1:
        np.cumsum(lengths[:-1], out=offsets[1:])
    return residues, offsets, lengths


def mutate(rng, seq, sub=0.10, indel=0.02):
    """A mutated copy: `sub` substitutions, `indel` insertions+deletions per residue."""
    out = []
    p = _probabilities()
    for r in seq:
        u = rng.random()
        if u < indel / 2:
            continue
        if u < indel:
            out.append(rng.choice(len(ALPHABET), p=p))
        out.append(rng.choice(len(ALPHABET), p=p) if rng.random() < sub else r)
    return np.array(out, dtype=np.uint8)


def database(seed, n, mu, sigma, lo, hi, giants=()):
    """(residues uint8[sum], offsets int64[n], lengths int32[n]); `giants` = extra planted lengths."""
    rng = np.random.Generator(np.random.PCG64(seed))
    lengths = random_lengths(rng, n, mu, sigma, lo, hi)
    if len(giants):
        pos = rng.choice(n, size=len(giants), replace=False)
        lengths[pos] = np.asarray(giants, dtype=np.int32)
    residues = random_residues(rng, int(lengths.sum()))
    return pack(lengths, residues)


def config1(n=100_000):
    """C1: 300-aa query x 100k proteins (mean ~350)."""
    db = database(1, n, 5.68, 0.60, 30, 5000)
    q = random_residues(np.random.Generator(np.random.PCG64(2)), 300)
    return q, db


def config2(n=570_000, giants=32):
    """C2: Swiss-Prot-shaped db (~570k, mean ~358) with planted giants of 10000..35213 aa."""
    g = np.linspace(10_000, 35_213, giants).astype(np.int32) if giants else ()
    return database(3, n, 5.65, 0.68, 2, 35_213, g)


def config2_queries(db, lengths=(100, 300, 1000, 2000, 5000)):
    """Half random, half mutated copies of database entries (so true high scores exist)."""
    residues, offsets, lens = db
    rng = np.random.Generator(np.random.PCG64(4))
    out = []
    for i, lq in enumerate(lengths):
        if i % 2 == 0:
            out.append(random_residues(rng, lq))
        else:
            cand = np.nonzero(lens >= lq)[0]
            k = int(cand[rng.integers(len(cand))])
            src = residues[offsets[k] : offsets[k] + lq]
            m = mutate(rng, src)
            out.append(m[:lq] if len(m) >= lq else np.concatenate([m, random_residues(rng, lq - len(m))]))
    return out


def pack_subset(db, idx):
    """Packed (residues, offsets, lengths) of the targets ``idx`` of a packed database."""
    residues, offsets, lengths = db[:3]
    idx = np.asarray(idx, dtype=np.int64)
    sl = np.ascontiguousarray(np.asarray(lengths)[idx]).astype(np.int32)
    so = np.zeros(len(idx), dtype=np.int64)
    if len(idx) > 1:
        np.cumsum(sl[:-1], out=so[1:])
    total = int(sl.sum())
    if total == 0:
        return np.zeros(0, dtype=np.uint8), so, sl
    src = np.repeat(np.asarray(offsets, dtype=np.int64)[idx] - so, sl) + np.arange(total, dtype=np.int64)
    return np.ascontiguousarray(np.asarray(residues)[src]), so, sl


def as_sequences(db):
    """List of encoded `bytes`, one per target."""
    residues, offsets, lengths = db
    buf = residues.tobytes()
    return [buf[int(o) : int(o) + int(n)] for o, n in zip(offsets, lengths)]


# ---------------------------------------------------------------------------
# C5: UniRef50-shaped database (~50 M proteins, ~1.4e10 residues).  Too large to
# draw in one piece, so it is defined block by block: block b holds proteins
# [b * C5_BLOCK, (b + 1) * C5_BLOCK), drawn from PCG64([5, b]); the blocks are
# dealt round-robin to the ranks (lengths are i.i.d., so every rank gets the
# same length mix and the same cell count to ~0.1 %).  Residues come from a
# 65536-entry inverse-CDF table indexed by 16 random bits (composition exact to
# 2^-16, ~10x faster than Generator.choice on 1.8e9 residues per rank).

C5_TARGETS = 50_000_000
C5_BLOCK = 781_250  # 64 blocks


def _residue_table():
    cdf = np.cumsum(_probabilities())
    u = (np.arange(65536) + 0.5) / 65536.0
    return np.minimum(np.searchsorted(cdf, u), len(ALPHABET) - 2).astype(np.uint8)


def config5_block_lengths(block, n=C5_BLOCK):
    rng = np.random.Generator(np.random.PCG64([5, int(block)]))
    return random_lengths(rng, n, 5.40, 0.70, 11, 40_000)


def config5_block(block, n=C5_BLOCK):
    """(residues, lengths) of block `block`."""
    rng = np.random.Generator(np.random.PCG64([5, int(block)]))
    lengths = random_lengths(rng, n, 5.40, 0.70, 11, 40_000)
    table = _residue_table()
    total = int(lengths.sum())
    residues = np.empty(total, dtype=np.uint8)
    step = 1 << 26
    for lo in range(0, total, step):
        hi = min(total, lo + step)
        np.take(table, rng.integers(0, 65536, size=hi - lo, dtype=np.uint16), out=residues[lo:hi])
    return residues, lengths


def config5_shard(rank, world, n_blocks, block=C5_BLOCK, threads=8):
    """This rank's blocks (b = rank, rank + world, ...) of an `n_blocks`-block database, packed:
    (residues, offsets, lengths, block ids)."""
    from concurrent.futures import ThreadPoolExecutor

    mine = list(range(rank, n_blocks, world))
    with ThreadPoolExecutor(max(1, min(threads, len(mine)))) as pool:
        parts = list(pool.map(lambda b: config5_block(b, block), mine))
    lengths = np.concatenate([p[1] for p in parts]) if parts else np.zeros(0, np.int32)
    residues = np.concatenate([p[0] for p in parts]) if parts else np.zeros(0, np.uint8)
    return pack(lengths, residues) + (mine,)


def config5_queries(n=64):
    """64 random queries, lengths lognormal as C1 clipped to [50, 2000] (seed 6)."""
    rng = np.random.Generator(np.random.PCG64(6))
    lens = random_lengths(rng, n, 5.68, 0.60, 50, 2000)
    return [random_residues(rng, int(l)) for l in lens]
