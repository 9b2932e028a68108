"""``pyopal.align`` for the B200 backend (reference: /root/reference/src/pyopal/_align.py)."""

import contextlib
import functools
import multiprocessing.pool

from .lib import Aligner, BaseDatabase, Database
from .matrices import ScoringMatrix, as_scoring_matrix


def align(
    query,
    database,
    scoring_matrix=None,
    *,
    gap_open=3,
    gap_extend=1,
    mode="score",
    overflow="buckets",
    algorithm="sw",
    threads=0,
    pool=None,
    ordered=False,
):
    """Align the query sequence to every database sequence in parallel.

    Same signature and results as the reference (_align.py:28-172).  The
    parallelism is the GPU's: ``threads=0`` (the default) issues ONE search
    over the whole database, which the device spreads over its SMs; a
    positive ``threads`` keeps the reference behaviour of cutting the
    database in ``threads`` contiguous chunks dispatched from a thread pool
    (_align.py:145-172) -- the chunks then serialise on the device, so it only
    exists for drop-in compatibility.

    Example:
        >>> targets = ["AACCGCTG", "ATGCGCT", "TTATTACG"]
        >>> for res in pyopal.align("ACCTG", targets, gap_open=2, ordered=True):
        ...     print(res.score, targets[res.target_index])
        41 AACCGCTG
        31 ATGCGCT
        23 TTATTACG

    """
    if threads == 0:
        threads = 1
    if scoring_matrix is None:
        scoring_matrix = Aligner._DEFAULT_SCORING_MATRIX
    elif isinstance(scoring_matrix, str):
        scoring_matrix = ScoringMatrix.from_name(scoring_matrix)
    elif as_scoring_matrix(scoring_matrix) is None:
        ty = type(scoring_matrix).__name__
        raise TypeError(f"expected str or ScoringMatrix, got {ty}")
    if not isinstance(database, BaseDatabase):
        database = Database(database, scoring_matrix.alphabet)

    # avoid using more threads than necessary
    if threads > len(database):
        threads = len(database) or 1

    aligner = Aligner(scoring_matrix, gap_open=gap_open, gap_extend=gap_extend)
    if threads == 1:
        yield from aligner.align(query, database, mode=mode, overflow=overflow, algorithm=algorithm)
    else:
        if pool is None:
            pool_context = multiprocessing.pool.ThreadPool(threads)
        else:
            pool_context = contextlib.nullcontext(pool)
        # cut the database in chunks of similar length
        chunk_length = len(database) // threads
        with pool_context as pool:
            _align = functools.partial(
                aligner.align, query, database, mode=mode, overflow=overflow, algorithm=algorithm
            )
            mapper = pool.imap if ordered else pool.imap_unordered
            chunk_hits = mapper(
                lambda x: _align(start=x, end=x + chunk_length),
                range(0, len(database), chunk_length),
            )
            for hits in chunk_hits:
                yield from hits
