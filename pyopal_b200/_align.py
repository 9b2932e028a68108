"""``pyopal.align`` for the B200 backend (reference: /root/reference/src/pyopal/_align.py)."""

from .lib import Aligner, BaseDatabase, Database
from .matrices import ScoringMatrix, as_scoring_matrix


def _resolve_matrix(scoring_matrix):
    if scoring_matrix is None:
        return Aligner._DEFAULT_SCORING_MATRIX
    if isinstance(scoring_matrix, str):
        return ScoringMatrix.from_name(scoring_matrix)
    if as_scoring_matrix(scoring_matrix) is None:
        ty = type(scoring_matrix).__name__
        raise TypeError(f"expected str or ScoringMatrix, got {ty}")
    return as_scoring_matrix(scoring_matrix)


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

    Same signature and results as the reference (_align.py:28-172), whose
    parallelism is a thread pool over ``threads`` contiguous chunks of equal
    sequence count.  Here the fan-out happens below the C-ABI: the database
    store is dealt to every visible GPU by cumulative cell count
    (csrc/multi.cu; ``PYOPAL_CUDA_DEVICES`` restricts the devices), each GPU
    scores its shard with one host thread and the shards' results come back in
    database order.  ``threads`` and ``pool`` are therefore accepted for
    compatibility and do not change how the search is dispatched, and the
    hits always arrive in database order (what ``ordered=True`` asks for).

    Example:
        >>> targets = ["AACCGCTG", "ATGCGCT", "TTATTACG"]
        >>> for res in pyopal.align("ACCTG", targets, gap_open=2, ordered=True):
        ...     print(res.score, targets[res.target_index])
        41 AACCGCTG
        31 ATGCGCT
        23 TTATTACG

    """
    if threads < 0:
        raise ValueError(f"invalid thread count: {threads!r}")
    scoring_matrix = _resolve_matrix(scoring_matrix)
    if not isinstance(database, BaseDatabase):
        database = Database(database, scoring_matrix.alphabet)
    aligner = Aligner(scoring_matrix, gap_open=gap_open, gap_extend=gap_extend)
    yield from aligner.align(query, database, mode=mode, overflow=overflow, algorithm=algorithm)


def align_batch(queries, database, scoring_matrix=None, *, gap_open=3, gap_extend=1, algorithm="sw", k=None):
    """Score several queries against the database in one device-side pipeline (additive).

    The reference's pattern for many queries is a ``ThreadPool`` over `Aligner.align` calls
    (/root/reference/README.md:116-144); here the queries go down in ONE call, the searches are
    stream-ordered on every GPU with nothing waiting on the host in between, and Smith-Waterman queries of similar
    length are searched two per pass of the score kernel (csrc/dual.cuh).  Yields, per query and in order, an ``array('i')``
    of scores in database order -- or, with ``k``, the list of the ``k`` best `ScoreResult` (selected on the
    device; score descending, ties by target index).
    """
    scoring_matrix = _resolve_matrix(scoring_matrix)
    if not isinstance(database, BaseDatabase):
        database = Database(database, scoring_matrix.alphabet)
    aligner = Aligner(scoring_matrix, gap_open=gap_open, gap_extend=gap_extend)
    if k is None:
        yield from aligner.align_batch(list(queries), database, algorithm=algorithm)
    else:
        yield from aligner.align_batch_topk(list(queries), database, k, algorithm=algorithm)
