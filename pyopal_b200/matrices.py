"""A minimal ``ScoringMatrix`` compatible with what PyOpal uses from the
external ``scoring_matrices`` package.

The reference imports ``scoring_matrices.ScoringMatrix``
(/root/reference/src/pyopal/lib.pyx:39, _align.py:6) and touches only:
``from_name``, ``alphabet``, ``is_integer()``, ``size()``/``len``, the float
data (``data_ptr``), ``==`` and pickling (lib.pyx:1153,1199-1209,1230-1251).
That package is not installable in this environment, so this module provides
the same surface; when the real package is importable it is preferred so that
every named matrix is available.
"""

from ._matrix_data import ALPHABET as _DEFAULT_ALPHABET, MATRICES as _MATRICES

try:  # pragma: no cover - not installed in the build image
    from scoring_matrices import ScoringMatrix as _RealScoringMatrix
except ImportError:  # the normal case here
    _RealScoringMatrix = None


class ScoringMatrix:
    """A square substitution matrix over an alphabet of uppercase letters."""

    DEFAULT_ALPHABET = _DEFAULT_ALPHABET

    def __init__(self, matrix, alphabet=_DEFAULT_ALPHABET, name=None):
        rows = [list(map(float, row)) for row in matrix]
        alphabet = str(alphabet)
        if len(rows) != len(alphabet):
            raise ValueError("matrix and alphabet have different sizes")
        if any(len(row) != len(rows) for row in rows):
            raise ValueError("`matrix` must be a square matrix")
        if len(set(alphabet)) != len(alphabet):
            raise ValueError("duplicate letters in alphabet")
        self._rows = rows
        self._alphabet = alphabet
        self._name = name

    @classmethod
    def from_name(cls, name="BLOSUM62"):
        """Load a built-in matrix by name (lib.pyx:1202, _align.py:122)."""
        if name in _MATRICES:
            return cls(_MATRICES[name], _DEFAULT_ALPHABET, name=name)
        if _RealScoringMatrix is not None:  # pragma: no cover
            real = _RealScoringMatrix.from_name(name)
            return cls([list(row) for row in real], real.alphabet, name=name)
        raise ValueError(
            f"unknown scoring matrix: {name!r} (built in: {sorted(_MATRICES)}; "
            "install `scoring_matrices` for the full set)"
        )

    # --- properties ---------------------------------------------------------

    @property
    def alphabet(self):
        return self._alphabet

    @property
    def name(self):
        return self._name

    @property
    def matrix(self):
        return [list(row) for row in self._rows]

    def size(self):
        return len(self._rows)

    def __len__(self):
        return len(self._rows)

    def __getitem__(self, item):
        if isinstance(item, tuple):
            i, j = item
            return self._rows[i][j]
        return tuple(self._rows[item])

    def __iter__(self):
        return (tuple(row) for row in self._rows)

    def is_integer(self):
        return all(x == int(x) for row in self._rows for x in row)

    def is_symmetric(self):
        n = len(self._rows)
        return all(self._rows[i][j] == self._rows[j][i] for i in range(n) for j in range(i))

    def min(self):
        return min(min(row) for row in self._rows)

    def max(self):
        return max(max(row) for row in self._rows)

    def shuffle(self, alphabet):
        """Return the matrix re-ordered for another alphabet order."""
        index = {c: i for i, c in enumerate(self._alphabet)}
        try:
            idx = [index[c] for c in alphabet]
        except KeyError as err:
            raise KeyError(f"letter not in matrix alphabet: {err.args[0]!r}") from None
        return type(self)([[self._rows[i][j] for j in idx] for i in idx], alphabet, name=self._name)

    # --- protocol -----------------------------------------------------------

    def __eq__(self, other):
        if _RealScoringMatrix is not None and isinstance(other, _RealScoringMatrix):  # pragma: no cover
            return self._alphabet == other.alphabet and self._rows == [list(r) for r in other]
        if not isinstance(other, ScoringMatrix):
            return NotImplemented
        return self._alphabet == other._alphabet and self._rows == other._rows

    def __ne__(self, other):
        eq = self.__eq__(other)
        return eq if eq is NotImplemented else not eq

    __hash__ = None

    def __reduce__(self):
        return type(self), (self._rows, self._alphabet, self._name)

    def __repr__(self):
        if self._name is not None and self._name in _MATRICES and self == ScoringMatrix.from_name(self._name):
            return f"{type(self).__name__}.from_name({self._name!r})"
        return f"{type(self).__name__}({self._rows!r}, alphabet={self._alphabet!r})"


def as_scoring_matrix(obj):
    """Accept our class or the real ``scoring_matrices.ScoringMatrix``."""
    if isinstance(obj, ScoringMatrix):
        return obj
    if _RealScoringMatrix is not None and isinstance(obj, _RealScoringMatrix):  # pragma: no cover
        return ScoringMatrix([list(row) for row in obj], obj.alphabet, name=getattr(obj, "name", None))
    return None
