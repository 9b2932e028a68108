"""B200-native drop-in for PyOpal's database-search path.

Public surface = /root/reference/src/pyopal/__init__.py:4-16.
"""

from . import lib
from .lib import (
    Aligner,
    Alphabet,
    BaseDatabase,
    Database,
    ScoreResult,
    EndResult,
    FullResult,
)
from .matrices import ScoringMatrix
from ._align import align, align_batch

__version__ = lib.__version__
__all__ = [
    "Aligner",
    "Alphabet",
    "BaseDatabase",
    "Database",
    "ScoreResult",
    "EndResult",
    "FullResult",
    "ScoringMatrix",
    "align",
    "align_batch",
]
