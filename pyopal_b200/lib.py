"""Host-side mirror of ``pyopal.lib`` for the B200 backend.

Same classes, signatures, reprs and exceptions as
/root/reference/src/pyopal/lib.pyx: ``Alphabet`` (:186-332), ``BaseDatabase``
(:337-466), ``Database`` (:469-778), ``ScoreResult`` / ``EndResult`` /
``FullResult`` (:783-1119) and ``Aligner`` (:1122-1383).  What changes is
underneath: ``Aligner`` always selects the CUDA platform module
(`pyopal_b200.platform.cuda.searchCUDA`) and `Database` keeps a
device-resident mirror of its sequences that is rebuilt after any mutation.
"""

import array
import operator
import threading

from .matrices import ScoringMatrix, as_scoring_matrix
from .platform import cuda as _cuda

try:
    from . import _results
except ImportError:  # not built yet: compile the host extension in-tree (gcc, one second), then import
    import os as _os
    import subprocess as _subprocess

    _subprocess.check_call(["bash", _os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "csrc", "build_host.sh")])
    from . import _results

__version__ = "0.7.3+b200"

MAX_ALPHABET_SIZE = 32  # lib.pxd:28-32
UINT32_MAX = 0xFFFFFFFF

_OPAL_SEARCH_MODES = {
    "score": _cuda.OPAL_SEARCH_SCORE,
    "end": _cuda.OPAL_SEARCH_SCORE_END,
    "full": _cuda.OPAL_SEARCH_ALIGNMENT,
}
_OPAL_OVERFLOW_MODES = {"simple": _cuda.OPAL_OVERFLOW_SIMPLE, "buckets": _cuda.OPAL_OVERFLOW_BUCKETS}
_OPAL_ALGORITHMS = {
    "nw": _cuda.OPAL_MODE_NW,
    "hw": _cuda.OPAL_MODE_HW,
    "ov": _cuda.OPAL_MODE_OV,
    "sw": _cuda.OPAL_MODE_SW,
}
_OPAL_ALIGNMENT_OPERATION = {
    "M": _cuda.OPAL_ALIGN_MATCH,
    "D": _cuda.OPAL_ALIGN_DEL,
    "I": _cuda.OPAL_ALIGN_INS,
    "X": _cuda.OPAL_ALIGN_MISMATCH,
}
_ALIGNMENT_SYMBOLS = b"MDIX"  # lib.pyx:991


def _gpu_info():
    """Build / runtime support of the CUDA backend (analogue of ``_cpu_info``, lib.pyx:133-148)."""
    import os

    built = os.path.exists(_cuda.library_path())
    count = _cuda.device_count() if built else 0
    return {
        "cuda": {"build": built, "runtime": count > 0, "devices": count},
        "device": _cuda.device_info(_cuda.default_device()) if count > 0 else None,
    }


# --- Exclusive read/write lock (lib.pyx:153-181) -----------------------------


class SharedMutex:
    """A readers/writer lock; ``std::shared_mutex`` in the reference (shared_mutex.pxd)."""

    def __init__(self):
        self._cond = threading.Condition(threading.Lock())
        self._readers = 0
        self._writer = False
        self.read = ReadLock(self)
        self.write = WriteLock(self)


class ReadLock:
    def __init__(self, owner):
        self.owner = owner

    def __enter__(self):
        o = self.owner
        with o._cond:
            while o._writer:
                o._cond.wait()
            o._readers += 1

    def __exit__(self, exc_type, exc_value, traceback):
        o = self.owner
        with o._cond:
            o._readers -= 1
            if o._readers == 0:
                o._cond.notify_all()


class WriteLock:
    def __init__(self, owner):
        self.owner = owner

    def __enter__(self):
        o = self.owner
        with o._cond:
            while o._writer or o._readers:
                o._cond.wait()
            o._writer = True

    def __exit__(self, exc_type, exc_value, traceback):
        o = self.owner
        with o._cond:
            o._writer = False
            o._cond.notify_all()


# --- Parameters (lib.pyx:186-332) --------------------------------------------


class Alphabet:
    """A class for ordinal encoding of sequences."""

    _DEFAULT_LETTERS = "ARNDCQEGHILKMFPSTWYVBZX*"

    def __init__(self, letters=_DEFAULT_LETTERS):
        if not isinstance(letters, str):
            raise TypeError(f"expected str, found {type(letters).__name__}")
        if len(letters) != len(set(letters)):
            raise ValueError("duplicate symbols in alphabet letters")
        if any(x != "*" and not x.isupper() for x in letters):
            raise ValueError("alphabet must only contain uppercase characters or wildcard")
        if len(letters) > MAX_ALPHABET_SIZE:
            raise ValueError("Cannot use alphabet of more than 32 symbols")
        self.letters = letters
        self.length = len(letters)
        self._unknown = letters.find("*")
        self._letters = letters.encode("ascii")
        # 256-entry translation table (the reference's 255-entry `_ahash`
        # reads out of bounds for byte 255):
        #   0xFF = non-alphabetic byte, 0xFE = alphabetic but not encodable
        table = bytearray([0xFF]) * 256
        for b in range(256):
            if (65 <= b <= 90) or (97 <= b <= 122):
                table[b] = self._unknown if self._unknown >= 0 else 0xFE
        for i, x in enumerate(self._letters):
            table[x] = i
        # a literal '*' is not alphabetic: rejected like the reference (lib.pyx:267-268)
        if self._unknown >= 0:
            table[ord("*")] = 0xFF
        self._table = bytes(table)

    def __len__(self):
        return self.length

    def __contains__(self, item):
        return item in self.letters

    def __getitem__(self, index):
        index_ = operator.index(index)
        if index_ < 0:
            index_ += self.length
        if index_ < 0 or index_ >= self.length:
            raise IndexError(index)
        return self.letters[index_]

    def __reduce__(self):
        return type(self), (self.letters,)

    def __repr__(self):
        if self.letters == self._DEFAULT_LETTERS:
            return f"{type(self).__name__}()"
        return f"{type(self).__name__}({self.letters!r})"

    def __str__(self):
        return self.letters

    def __eq__(self, item):
        if isinstance(item, str):
            return self.letters == item
        elif isinstance(item, Alphabet):
            return self.letters == item.letters
        return False

    __hash__ = None

    def _translate(self, sequence):
        data = bytes(sequence).translate(self._table)
        if b"\xff" in data or b"\xfe" in data:
            raw = bytes(sequence)
            for i, code in enumerate(data):
                if code == 0xFF:
                    raise ValueError(f"character outside ASCII range: {raw[i]!r}")
                if code == 0xFE:
                    raise ValueError(f"non-alphabet character in sequence: {chr(raw[i])!r}")
        return data

    def encode_into(self, sequence, encoded):
        """Encode a sequence to ordinal-encoding into the given buffer."""
        src = memoryview(sequence)
        dst = memoryview(encoded)
        if src.nbytes != dst.nbytes:
            raise ValueError("Buffers do not have the same dimensions")
        dst[:] = self._translate(src)

    def decode_into(self, encoded, sequence):
        """Decode a sequence from ordinal-encoding into the given buffer."""
        src = memoryview(encoded)
        dst = memoryview(sequence)
        if src.nbytes != dst.nbytes:
            raise ValueError("Buffers do not have the same dimensions")
        dst[:] = self._decode_bytes(src)

    def _decode_bytes(self, encoded):
        data = bytes(encoded)
        if data and max(data) >= self.length:
            bad = next(c for c in data if c >= self.length)
            raise ValueError(f"invalid index in encoded sequence: {bad!r}")
        return data.translate(self._letters + bytes(256 - self.length))

    def encode(self, sequence):
        r"""Encode a sequence to an ordinal-encoded sequence using the alphabet.

        Example:
            >>> alphabet = Alphabet("ACGT")
            >>> alphabet.encode("GATACA")
            b'\x02\x00\x03\x00\x01\x00'

        """
        if isinstance(sequence, str):
            try:
                sequence = sequence.encode("ascii")
            except UnicodeEncodeError as err:
                raise ValueError(f"character outside ASCII range: {sequence[err.start]!r}") from None
        elif isinstance(sequence, int):
            raise TypeError(f"expected str or a bytes-like object, found {type(sequence).__name__}")
        return self._translate(sequence)

    def decode(self, encoded):
        r"""Decode an ordinal-encoded sequence using the alphabet.

        Example:
            >>> alphabet = Alphabet("ACGT")
            >>> alphabet.decode(bytearray([2, 0, 3, 0, 1, 0]))
            'GATACA'

        """
        return self._decode_bytes(encoded).decode("ascii")


# --- Sequence storage (lib.pyx:337-778) --------------------------------------


class BaseDatabase:
    """The base class for views of database sequences.

    Subclasses implement `get_size`, `get_lengths` (a sequence of `int`) and
    `get_sequences` (a sequence of encoded `bytes`-like objects); these are
    the Python spellings of the reference's ``cdef`` interface
    (lib.pxd:86-92).  Such a database is searched through the generic entry
    point (`opalSearchDatabase` signature), staged to the device per call.
    """

    _DEFAULT_ALPHABET = Alphabet()

    def __new__(cls, *args, **kwargs):
        self = super().__new__(cls)
        self.lock = SharedMutex()
        self.alphabet = cls._DEFAULT_ALPHABET
        return self

    def __init__(self, sequences=(), alphabet=None):
        if alphabet is None:
            self.alphabet = self._DEFAULT_ALPHABET
        elif isinstance(alphabet, Alphabet):
            self.alphabet = alphabet
        else:
            self.alphabet = Alphabet(alphabet)
        if sequences:
            raise TypeError("cannot create a `BaseDatabase` with sequences")

    @property
    def lengths(self):
        """`list` of `int`: The length of each sequence in the database."""
        with self.lock.read:
            return list(self.get_lengths()[: self.get_size()])

    @property
    def total_length(self):
        """`int`: The total length of the database (64-bit; the reference sums in C ``int``)."""
        with self.lock.read:
            return sum(self.get_lengths()[: self.get_size()])

    # --- Database interface ---

    def get_sequences(self):
        raise NotImplementedError("BaseDatabase.get_sequences")

    def get_lengths(self):
        raise NotImplementedError("BaseDatabase.get_lengths")

    def get_size(self):
        return 0

    # --- Sequence interface ---

    def __contains__(self, query):
        encoded = self.alphabet.encode(query)
        with self.lock.read:
            size = self.get_size()
            sequences = self.get_sequences()
            return any(bytes(sequences[i]) == encoded for i in range(size))

    def __len__(self):
        with self.lock.read:
            return self.get_size()

    def __getitem__(self, index):
        index_ = operator.index(index)
        with self.lock.read:
            size = self.get_size()
            if index_ < 0:
                index_ += size
            if index_ < 0 or index_ >= size:
                raise IndexError(index)
            return self.alphabet.decode(self.get_sequences()[index_])

    def __iter__(self):
        # the reference relies on the legacy __getitem__/IndexError protocol
        i = 0
        while True:
            try:
                yield self[i]
            except IndexError:
                return
            i += 1


class _PackedSequences:
    """Read-only sequence view over a packed residue buffer (what `get_sequences` returns for a packed database)."""

    __slots__ = ("buffer", "offsets", "lengths")

    def __init__(self, buffer, offsets, lengths):
        self.buffer, self.offsets, self.lengths = buffer, offsets, lengths

    def __len__(self):
        return len(self.lengths)

    def __getitem__(self, index):
        if isinstance(index, slice):
            return [self[i] for i in range(*index.indices(len(self)))]
        i = operator.index(index)
        if i < 0:
            i += len(self.lengths)
        o = self.offsets[i]
        return bytes(self.buffer[o : o + self.lengths[i]])

    def __iter__(self):
        return (self[i] for i in range(len(self.lengths)))


class Database(BaseDatabase):
    """A database of target sequences, stored encoded (lib.pyx:469-778).

    Host store, two forms.  A database built one sequence at a time
    (`append` / `extend` / the constructor, lib.pyx:596-636) keeps one encoded
    `bytes` per sequence, shared between the views made by `mask` / `extract`
    like the reference's ``shared_ptr`` buffers.  A database built in bulk
    (`from_packed`, `from_fasta`, `load`) keeps ONE packed residue buffer plus
    ``int64`` offsets and ``int32`` lengths -- zero-copy over the caller's buffer
    or an ``mmap`` -- and is handed to the device as it lies; the first
    per-sequence mutation turns it into the list form.

    The first search uploads the residues once and builds the length-sorted,
    pair-interleaved store on the GPU(s); every mutator invalidates it under
    the write lock.
    """

    def __new__(cls, *args, **kwargs):
        self = super().__new__(cls)
        self._sequences = []
        self._lengths = []
        self._packed = None      # (buffer, offsets array('q'), lengths array('i')) or None
        self._keepalive = None   # e.g. the mmap behind a loaded database
        self._version = 0
        self._store = None
        self._store_version = -1
        self._store_lock = threading.Lock()
        return self

    def __init__(self, sequences=(), alphabet=None):
        super().__init__(alphabet=alphabet)
        self.clear()
        self.extend(sequences)

    def __reduce__(self):
        return (type(self), ((), self.alphabet), None, iter(self))

    # --- Database interface ---

    def get_sequences(self):
        if self._packed is not None:
            return _PackedSequences(*self._packed)
        return self._sequences

    def get_lengths(self):
        return self._packed[2] if self._packed is not None else self._lengths

    def get_size(self):
        return len(self._packed[2]) if self._packed is not None else len(self._sequences)

    @property
    def total_length(self):
        with self.lock.read:
            if self._packed is not None:
                return _cuda.offsets_from_lengths(self._packed[2])[1]
            return sum(self._lengths)

    def _unpack(self):
        """Packed form -> list form (call under the write lock, before a per-sequence mutation)."""
        if self._packed is not None:
            seqs = _PackedSequences(*self._packed)
            self._sequences = [seqs[i] for i in range(len(seqs))]
            self._lengths = list(self._packed[2])
            self._packed = None
            self._keepalive = None

    # --- Device mirror ---

    def _touch(self):
        self._version += 1

    def _device_store(self):
        """The device-resident store for the current contents (call under the read lock).

        Every visible GPU is used (``PYOPAL_CUDA_DEVICES`` restricts them; under a one-process-per-GPU launcher
        the process keeps to its own device): the thread fan-out of the reference's ``pyopal.align``
        (_align.py:145-172) becomes a cell-balanced deal over the devices inside the library.
        """
        with self._store_lock:
            if self._store is None or self._store_version != self._version:
                if self._store is not None:
                    self._store.close()
                devices = _cuda.default_devices()
                if self._packed is not None:
                    buf, offsets, lengths = self._packed
                    total = (offsets[-1] + lengths[-1]) if len(lengths) else 0
                else:
                    buf, offsets, lengths = None, None, self._lengths
                    total = sum(lengths)
                if len(devices) > 1 and total < _cuda.MULTI_GPU_MIN_RESIDUES and not _cuda.devices_forced():
                    devices = devices[:1]  # a small database: one GPU answers before eight have been fed
                if self._packed is not None:
                    self._store = _cuda.make_store(buf, offsets, lengths, self.alphabet.length, devices)
                else:
                    self._store = _cuda.make_store_from_sequences(self._sequences, self.alphabet.length, devices)
                self._store_version = self._version
            return self._store

    # --- Encoding utility ---

    def _encode(self, sequence):
        return self.alphabet.encode(sequence)

    # --- Bulk ingest and persistence (additive) ---
    #
    # The reference appends one sequence at a time: a Python call, an encode
    # and a heap allocation each (lib.pyx:596-636), and pickling re-appends the
    # decoded strings one by one (lib.pyx:496-497).  These constructors take
    # whole buffers instead and never loop over the sequences in Python.

    _MAGIC = b"PYOPALDB2\n"
    _ALIGN = 64

    @classmethod
    def from_packed(cls, residues, lengths, alphabet=None, encoded=True, validate=True):
        """Build a database from one packed buffer and the sequence lengths, without copying it.

        ``residues`` holds the sequences back to back (any C-contiguous
        bytes-like object: `bytes`, ``numpy.uint8``, an ``mmap`` ...): alphabet
        indices when ``encoded`` is true, ASCII letters otherwise (translated in
        one call).  ``lengths``: any buffer or iterable of `int`.
        """
        self = cls((), alphabet)
        data = memoryview(residues).cast("B") if encoded else memoryview(self.alphabet.encode(bytes(residues)))
        lens = _cuda.as_int32_array(lengths)
        offsets, total = _cuda.offsets_from_lengths(lens)
        if (len(lens) and min(lens) < 0) or total != data.nbytes:
            raise ValueError("lengths do not add up to the size of the residue buffer")
        if encoded and validate and data.nbytes and _cuda.has_byte_ge(data, self.alphabet.length):
            raise ValueError("residue code outside the alphabet")
        with self.lock.write:
            self._packed = (data, offsets, lens)
            self._keepalive = residues
            self._sequences, self._lengths = [], []
            self._touch()
        return self

    @classmethod
    def from_fasta(cls, file, alphabet=None):
        """Read every record of a FASTA file (path or binary file object); returns ``(database, names)``."""
        if hasattr(file, "read"):
            blob = file.read()
        else:
            with open(file, "rb") as handle:
                blob = handle.read()
        if isinstance(blob, str):
            blob = blob.encode("ascii")
        names, chunks, lengths = [], [], []
        strip = bytes.maketrans(b"", b"")
        for record in blob.split(b">")[1:]:
            header, _, body = record.partition(b"\n")
            seq = body.translate(strip, b"\r\n \t")
            names.append(header.strip().decode("ascii", "replace"))
            chunks.append(seq)
            lengths.append(len(seq))
        return cls.from_packed(b"".join(chunks), lengths, alphabet, encoded=False), names

    def _packed_view(self):
        """(buffer, offsets, lengths) of the current contents (packing the list form if needed)."""
        if self._packed is not None:
            return self._packed
        lens = array.array("i", self._lengths)
        offsets, _ = _cuda.offsets_from_lengths(lens)
        return memoryview(b"".join(self._sequences)), offsets, lens

    def dump(self, file):
        """Write the encoded database in a packed binary form that `load` maps back without copying.

        Layout (little endian, sections aligned to 64 bytes so that an ``mmap`` of the file can be handed to
        the device as it lies): magic, alphabet, sequence count, residue count, ``int32`` lengths, residues.
        The length sort / pairing / interleaving of the device store is not persisted: it is rebuilt ON the
        GPU from these residues in milliseconds, and the raw residues are half the bytes of the interleaved
        stream to push over PCIe.
        """
        with self.lock.read:
            buf, _, lens = self._packed_view()
            letters = self.alphabet.letters.encode("ascii")
            head = self._MAGIC + len(letters).to_bytes(4, "little") + letters
            head += b"\0" * (-len(head) % 8)
            head += len(lens).to_bytes(8, "little") + buf.nbytes.to_bytes(8, "little")
            head += b"\0" * (-len(head) % self._ALIGN)
            lens_b = memoryview(lens).cast("B")
            payload = [head, lens_b, b"\0" * (-lens_b.nbytes % self._ALIGN), buf]
            if hasattr(file, "write"):
                for part in payload:
                    file.write(part)
            else:
                with open(file, "wb") as handle:
                    for part in payload:
                        handle.write(part)

    @classmethod
    def load(cls, file):
        """Read a database written by `dump`: a path is memory-mapped (no copy, no per-sequence work)."""
        keep = None
        if hasattr(file, "read"):
            view = memoryview(file.read())
        else:
            import mmap

            with open(file, "rb") as handle:
                keep = mmap.mmap(handle.fileno(), 0, access=mmap.ACCESS_READ)
            view = memoryview(keep)
        m = len(cls._MAGIC)
        if bytes(view[:m]) != cls._MAGIC:
            raise ValueError("not a packed pyopal database")
        nl = int.from_bytes(view[m : m + 4], "little")
        letters = bytes(view[m + 4 : m + 4 + nl]).decode("ascii")
        pos = m + 4 + nl
        pos += -pos % 8
        n = int.from_bytes(view[pos : pos + 8], "little")
        total = int.from_bytes(view[pos + 8 : pos + 16], "little")
        pos += 16
        pos += -pos % cls._ALIGN
        lens = view[pos : pos + 4 * n].cast("i")
        pos += 4 * n
        pos += -pos % cls._ALIGN
        self = cls.from_packed(view[pos : pos + total], lens, Alphabet(letters), encoded=True, validate=False)
        self._keepalive = (keep, view)
        return self

    # --- Sequence interface ---

    def __getitem__(self, index):
        if isinstance(index, slice):
            indices = range(*index.indices(len(self)))
            return self.extract(indices)
        return super().__getitem__(index)

    def __setitem__(self, index, sequence):
        index_ = operator.index(index)
        encoded = self._encode(sequence)
        with self.lock.write:
            self._unpack()
            size = len(self._sequences)
            if index_ < 0:
                index_ += size
            if index_ < 0 or index_ >= size:
                raise IndexError(index)
            self._sequences[index_] = encoded
            self._lengths[index_] = len(encoded)
            self._touch()

    def __delitem__(self, index):
        index_ = operator.index(index)
        with self.lock.write:
            self._unpack()
            size = len(self._sequences)
            if index_ < 0:
                index_ += size
            if index_ < 0 or index_ >= size:
                raise IndexError(index)
            del self._sequences[index_]
            del self._lengths[index_]
            self._touch()

    def clear(self):
        """Remove all sequences from the database."""
        with self.lock.write:
            self._packed = None
            self._keepalive = None
            self._sequences = []
            self._lengths = []
            self._touch()

    def extend(self, sequences):
        """Extend the database by adding sequences from an iterable.

        Example:
            >>> db = pyopal.Database(["ATGC"])
            >>> db.extend(["TTCA", "AAAA", "GGTG"])
            >>> list(db)
            ['ATGC', 'TTCA', 'AAAA', 'GGTG']

        """
        for sequence in sequences:
            self.append(sequence)

    def append(self, sequence):
        """Append a single sequence at the end of the database.

        Example:
            >>> db = pyopal.Database(["ATGC", "TTCA"])
            >>> db.append("AAAA")
            >>> list(db)
            ['ATGC', 'TTCA', 'AAAA']

        """
        encoded = self._encode(sequence)
        with self.lock.write:
            self._unpack()
            self._sequences.append(encoded)
            self._lengths.append(len(encoded))
            self._touch()

    def reverse(self):
        """Reverse the database, in place.

        Example:
            >>> db = pyopal.Database(['ATGC', 'TTGC', 'CTGC'])
            >>> db.reverse()
            >>> list(db)
            ['CTGC', 'TTGC', 'ATGC']

        """
        with self.lock.write:
            self._unpack()
            self._sequences.reverse()
            self._lengths.reverse()
            self._touch()

    def insert(self, index, sequence):
        """Insert a sequence in the database at a given position.

        Note:
            If the insertion index is out of bounds, the insertion will
            happen at either end of the database::

                >>> db = pyopal.Database(["ATGC", "TTGC", "CTGC"])
                >>> db.insert(-100, "TTTT")
                >>> db.insert(100, "AAAA")
                >>> list(db)
                ['TTTT', 'ATGC', 'TTGC', 'CTGC', 'AAAA']

        """
        index_ = operator.index(index)
        encoded = self._encode(sequence)
        with self.lock.write:
            self._unpack()
            size = len(self._sequences)
            if index_ < 0:
                index_ += size
            if index_ < 0:
                index_ = 0
            elif index_ >= size:
                index_ = size
            self._sequences.insert(index_, encoded)
            self._lengths.insert(index_, len(encoded))
            self._touch()

    # --- Subset ---

    def _subset(self, positions):
        subdb = Database.__new__(Database)
        subdb.alphabet = self.alphabet
        sequences, lengths = self.get_sequences(), self.get_lengths()
        subdb._sequences = [sequences[i] for i in positions]
        subdb._lengths = [lengths[i] for i in positions]
        return subdb

    def mask(self, bitmask):
        """Extract a subset of the database where the bitmask is `True`.

        Example:
            >>> db = pyopal.Database(['AAAA', 'CCCC', 'KKKK', 'FFFF'])
            >>> list(db.mask([True, False, False, True]))
            ['AAAA', 'FFFF']

        """
        with self.lock.read:
            size = self.get_size()
            positions = []
            i = 0
            for b in bitmask:
                if i >= size:
                    raise IndexError(bitmask)
                if b:
                    positions.append(i)
                i += 1
            if i < size:
                raise IndexError(bitmask)
            return self._subset(positions)

    def extract(self, indices):
        """Extract a subset of the database using the given indices.

        Example:
            >>> db = pyopal.Database(['AAAA', 'CCCC', 'KKKK', 'FFFF'])
            >>> list(db.extract([2, 0]))
            ['KKKK', 'AAAA']

        Caution:
            Negative indexing is not supported.

        """
        with self.lock.read:
            size = self.get_size()
            positions = []
            for index in indices:
                index = operator.index(index)
                if index < 0 or index >= size:
                    raise IndexError(index)
                positions.append(index)
            return self._subset(positions)


# --- Results (lib.pyx:783-1119) -----------------------------------------------


class ScoreResult(_results.ResultBase):
    """The results of a search in ``score`` mode.

    The fields (the ``OpalSearchResult`` of opal.pxd:26-34 with the defaults of
    ``__cinit__``, lib.pyx:787-795) live in the C base class
    ``_results.ResultBase`` (csrc/_results.c), as they live in the Cython
    extension type of the reference (lib.pxd:113-126).
    """

    __slots__ = ()

    def __init__(self, target_index, score):
        self._target_index = operator.index(target_index)
        self._score = operator.index(score)

    def __repr__(self):
        ty = type(self).__name__
        return f"{ty}({self.target_index}, score={self.score!r})"

    def __reduce__(self):
        return type(self), (self.target_index, self.score)

    def __eq__(self, other):
        if not isinstance(other, ScoreResult):
            return NotImplemented
        return self.__reduce__()[1] == other.__reduce__()[1]

    __hash__ = None

    @property
    def target_index(self):
        """`int`: The index of the target in the database."""
        assert self._target_index >= 0
        return self._target_index

    @property
    def score(self):
        """`int`: The score of the alignment."""
        assert self._score is not None
        return self._score


class EndResult(ScoreResult):
    """The results of a search in ``end`` mode."""

    __slots__ = ()

    def __init__(self, target_index, score, query_end, target_end):
        super().__init__(target_index, score)
        self._query_end = operator.index(query_end)
        self._target_end = operator.index(target_end)

    def __repr__(self):
        ty = type(self).__name__
        return (
            f"{ty}({self.target_index}, "
            f"score={self.score!r}, "
            f"query_end={self.query_end!r}, "
            f"target_end={self.target_end!r})"
        )

    def __reduce__(self):
        return type(self), (self.target_index, self.score, self.query_end, self.target_end)

    @property
    def query_end(self):
        """`int`: The coordinate where the alignment ends in the query."""
        return self._query_end

    @property
    def target_end(self):
        """`int`: The coordinate where the alignment ends in the target."""
        return self._target_end


class FullResult(EndResult):
    """The results of a search in ``full`` mode."""

    __slots__ = ()

    def __init__(
        self,
        target_index,
        score,
        query_end,
        target_end,
        query_start,
        target_start,
        query_length,
        target_length,
        alignment,
    ):
        super().__init__(target_index, score, query_end, target_end)
        if not isinstance(alignment, str):
            raise TypeError(f"expected str, found {type(alignment).__name__}")
        self._query_length = operator.index(query_length)
        self._target_length = operator.index(target_length)
        self._query_start = operator.index(query_start)
        self._target_start = operator.index(target_start)
        self._alignment = bytes(_OPAL_ALIGNMENT_OPERATION[x] for x in alignment)

    def __repr__(self):
        ty = type(self).__name__
        return (
            f"{ty}({self.target_index}, "
            f"score={self.score!r}, "
            f"query_end={self.query_end!r}, "
            f"target_end={self.target_end!r}, "
            f"query_start={self.query_start!r}, "
            f"target_start={self.target_start!r}, "
            f"query_length={self.query_length!r}, "
            f"target_length={self.target_length!r}, "
            f"alignment={self.alignment!r})"
        )

    def __reduce__(self):
        return (
            type(self),
            (
                self.target_index,
                self.score,
                self.query_end,
                self.target_end,
                self.query_start,
                self.target_start,
                self.query_length,
                self.target_length,
                self.alignment,
            ),
        )

    @property
    def query_start(self):
        """`int`: The coordinate where the alignment starts in the query."""
        return self._query_start

    @property
    def target_start(self):
        """`int`: The coordinate where the alignment starts in the target."""
        return self._target_start

    @property
    def query_length(self):
        """`int`: The complete length of the query sequence."""
        return self._query_length

    @property
    def target_length(self):
        """`int`: The complete length of the target sequence."""
        return self._target_length

    @property
    def alignment(self):
        """`str`: A string used by Opal to encode alignments."""
        return (self._alignment or b"").translate(_ALIGNMENT_SYMBOLS + bytes(252)).decode("ascii")

    def cigar(self):
        """Create a CIGAR string representing the alignment.

        Example:
            >>> aligner = Aligner()
            >>> db = Database(["AACCGCTG"])
            >>> hit = aligner.align("ACCTCG", db, mode="full", algorithm="nw")[0]
            >>> hit.cigar()
            '1D5M1D1M'

        """
        ops = self._alignment
        if not ops:
            return None  # as the reference (lib.pyx:1020-1021)
        symbols = "MID"  # op % 3: M,X -> M; Opal DEL -> I; Opal INS -> D (lib.pyx:1017-1026)
        chunks = []
        count = 0
        current = ops[0] % 3
        for op in ops:
            symbol = op % 3
            if symbol == current:
                count += 1
            else:
                chunks.append(str(count))
                chunks.append(symbols[current])
                current = symbol
                count = 1
        chunks.append(str(count))
        chunks.append(symbols[current])
        return "".join(chunks)

    def identity(self):
        """Compute the identity of the alignment (matches over aligned columns)."""
        ops = self._alignment or b""
        matches = ops.count(_cuda.OPAL_ALIGN_MATCH)
        mismatches = ops.count(_cuda.OPAL_ALIGN_MISMATCH)
        if matches + mismatches == 0:
            return float("nan")  # the reference divides by zero here
        return matches / (matches + mismatches)

    def coverage(self, reference="query"):
        """Compute the coverage of the alignment.

        Example:
            >>> aligner = Aligner()
            >>> db = Database(["AACCGCTG"])
            >>> hit = aligner.align("ACCTCG", db, mode="full", algorithm="nw")[0]
            >>> hit.coverage("query")
            1.0
            >>> hit.coverage("target")
            0.875

        """
        if reference == "query":
            reflength = self._query_length
            length = self._query_end + 1 - self._query_start
            operation = _cuda.OPAL_ALIGN_DEL
        elif reference == "target":
            reflength = self._target_length
            length = self._target_end + 1 - self._target_start
            operation = _cuda.OPAL_ALIGN_INS
        else:
            raise ValueError(f"Invalid coverage reference: {reference!r}")
        ops = self._alignment or b""
        # trim alignment sides if they correspond to a gap in the reference (lib.pyx:1106-1116)
        for op in ops:
            if op == operation:
                length -= 1
            else:
                break
        for op in reversed(ops):
            if op == operation:
                length -= 1
            else:
                break
        return 0.0 if length < 0 else length / reflength


# --- Aligner (lib.pyx:1122-1383) ----------------------------------------------


class Aligner:
    """The Opal aligner, running on the GPU.

    Attributes:
        scoring_matrix (`ScoringMatrix`): The scoring matrix.
        alphabet (`Alphabet`): The alphabet for encoding sequences.
        gap_open (`int`): The gap opening penalty.
        gap_extend (`int`): The gap extension penalty.

    """

    _DEFAULT_SCORING_MATRIX = ScoringMatrix.from_name("BLOSUM50")
    _DEFAULT_GAP_OPEN = 3
    _DEFAULT_GAP_EXTEND = 1

    def __init__(self, scoring_matrix=None, gap_open=_DEFAULT_GAP_OPEN, gap_extend=_DEFAULT_GAP_EXTEND):
        if scoring_matrix is None:
            self.scoring_matrix = self._DEFAULT_SCORING_MATRIX
        elif isinstance(scoring_matrix, str):
            self.scoring_matrix = ScoringMatrix.from_name(scoring_matrix)
        elif as_scoring_matrix(scoring_matrix) is not None:
            self.scoring_matrix = as_scoring_matrix(scoring_matrix)
        else:
            ty = type(scoring_matrix).__name__
            raise TypeError(f"expected str or ScoringMatrix, found {ty}")

        self.alphabet = Alphabet(self.scoring_matrix.alphabet)
        self.gap_open = operator.index(gap_open)
        self.gap_extend = operator.index(gap_extend)

        # One backend, selected unconditionally (cf. the SSE2 < SSE4 < AVX2 <
        # NEON ladder at lib.pyx:1213-1227).  The import only binds the
        # function; a missing library or GPU raises RuntimeError at search
        # time with the reference's exception class.
        from .platform.cuda import searchCUDA

        self._search = searchCUDA

        if not self.scoring_matrix.is_integer():
            raise ValueError("Integer scoring matrix is expected")
        # row-major float -> int, as lib.pyx:1234-1238
        self._int_matrix = array.array("i", [int(x) for row in self.scoring_matrix for x in row])

    def __repr__(self):
        args = []
        if self.scoring_matrix != self._DEFAULT_SCORING_MATRIX:
            args.append(f"{self.scoring_matrix!r}")
        if self.gap_open != self._DEFAULT_GAP_OPEN:
            args.append(f"gap_open={self.gap_open!r}")
        if self.gap_extend != self._DEFAULT_GAP_EXTEND:
            args.append(f"gap_extend={self.gap_extend!r}")
        return f"{type(self).__name__}({', '.join(args)})"

    def __reduce__(self):
        return type(self), (self.scoring_matrix, self.gap_open, self.gap_extend)

    def __eq__(self, other):
        if not isinstance(other, Aligner):
            return NotImplemented
        return self.__reduce__()[1] == other.__reduce__()[1]

    __hash__ = None

    def _validate(self, mode, overflow, algorithm):
        if mode not in _OPAL_SEARCH_MODES:
            raise ValueError(f"invalid search mode: {mode!r}")
        if overflow not in _OPAL_OVERFLOW_MODES:
            raise ValueError(f"invalid overflow mode: {overflow!r}")
        if algorithm not in _OPAL_ALGORITHMS:
            raise ValueError(f"invalid algorithm: {algorithm!r}")
        return _OPAL_SEARCH_MODES[mode], _OPAL_OVERFLOW_MODES[overflow], _OPAL_ALGORITHMS[algorithm]

    def align(self, query, database, *, mode="score", overflow="buckets", algorithm="sw", start=0, end=UINT32_MAX):
        """Align the query sequence to all targets of the database.

        Same arguments, return types and exceptions as the reference
        (lib.pyx:1258-1383): a `list` with one `ScoreResult` / `EndResult` /
        `FullResult` per target of ``database[start:end]``.
        """
        if query is None:
            raise TypeError("Argument 'query' must not be None")
        if not isinstance(database, BaseDatabase):
            raise TypeError(f"Argument 'database' has incorrect type (expected BaseDatabase, got {type(database).__name__})")
        _mode, _overflow, _algo = self._validate(mode, overflow, algorithm)
        start = operator.index(start)
        end = operator.index(end)
        if start < 0 or end < 0 or start > UINT32_MAX or end > UINT32_MAX:
            raise OverflowError("can't convert value to uint32_t")

        if database.alphabet != self.alphabet:
            raise ValueError("database and score matrix have different alphabets")
        encoded = database.alphabet.encode(query)

        with database.lock.read:
            size = database.get_size()
            if end < start:
                raise IndexError("database slice end is lower than start")
            if end > size:
                end = size
            return self._search(
                encoded, database, _mode, _overflow, _algo, self.gap_open, self.gap_extend, self._int_matrix,
                start, end,
            )

    @staticmethod
    def _slice(database, start, end):
        """``[start, end)`` checked exactly as `align` does (lib.pyx:1365-1370; call under the read lock)."""
        start = operator.index(start)
        end = operator.index(end)
        if start < 0 or end < 0 or start > UINT32_MAX or end > UINT32_MAX:
            raise OverflowError("can't convert value to uint32_t")
        size = database.get_size()
        if end < start:
            raise IndexError("database slice end is lower than start")
        end = min(end, size)
        if start > end:
            raise IndexError("database slice start is past the end")
        return start, end

    def _bulk_args(self, database, algorithm):
        if not isinstance(database, Database):
            raise TypeError("this call needs a pyopal Database (device-resident store)")
        _, _, _algo = self._validate("score", "buckets", algorithm)
        if database.alphabet != self.alphabet:
            raise ValueError("database and score matrix have different alphabets")
        return _algo

    def align_scores(self, query, database, *, algorithm="sw", start=0, end=UINT32_MAX):
        """Score-mode search returning a flat ``array('i')`` instead of result objects.

        Additive to the reference API: for a large
        database the per-target Python objects built at pyx.in:65-72 cost far
        more than the search itself once the DP runs on the GPU.
        """
        _algo = self._bulk_args(database, algorithm)
        encoded = database.alphabet.encode(query)
        with database.lock.read:
            start, end = self._slice(database, start, end)
            if end == start:
                return array.array("i")
            store = database._device_store()
            return store.search(encoded, self.gap_open, self.gap_extend, self._int_matrix,
                                _cuda.OPAL_SEARCH_SCORE, _algo, start=start, end=end)[0]

    def align_topk(self, query, database, k, *, algorithm="sw", start=0, end=UINT32_MAX):
        """The ``k`` best targets for ``query``: a list of `ScoreResult` sorted by score (descending), ties by
        target index.

        Additive to the reference API.  The reference returns one object per target
        (pyx.in:65-72) and the caller sorts them to pick the hits worth re-aligning in ``"full"`` mode
        (docs/examples/abc.ipynb); here the selection runs on the device (csrc/topk.cu) and only ``k``
        (score, index) pairs come back.
        """
        _algo = self._bulk_args(database, algorithm)
        if k < 0:
            raise ValueError("k must be non-negative")
        encoded = database.alphabet.encode(query)
        with database.lock.read:
            start, end = self._slice(database, start, end)
            if end == start or k == 0:
                return []
            store = database._device_store()
            scores, indices = store.search_topk(encoded, self.gap_open, self.gap_extend, self._int_matrix, _algo, k,
                                                start=start, end=end)
        return [ScoreResult(int(i), int(s)) for s, i in zip(scores, indices)]

    def align_batch(self, queries, database, *, algorithm="sw"):
        """Score several queries against the whole database in one device-side pipeline.

        Returns a list of ``array('i')``, one per query.
        """
        _algo = self._bulk_args(database, algorithm)
        encoded = [database.alphabet.encode(q) for q in queries]
        with database.lock.read:
            size = database.get_size()
            if size == 0 or not encoded:
                return [array.array("i") for _ in encoded]
            store = database._device_store()
            flat = store.search_batch(encoded, self.gap_open, self.gap_extend, self._int_matrix, _algo)
            return [flat[i * size : (i + 1) * size] for i in range(len(encoded))]

    def align_batch_topk(self, queries, database, k, *, algorithm="sw"):
        """The ``k`` best `ScoreResult` of every query (device-side selection, only the hits leave the GPU)."""
        _algo = self._bulk_args(database, algorithm)
        if k < 0:
            raise ValueError("k must be non-negative")
        encoded = [database.alphabet.encode(q) for q in queries]
        with database.lock.read:
            size = database.get_size()
            k = min(int(k), size)
            if size == 0 or not encoded or k == 0:
                return [[] for _ in encoded]
            store = database._device_store()
            vals, idx = store.search_batch_topk(encoded, self.gap_open, self.gap_extend, self._int_matrix, _algo, k)
        return [[ScoreResult(int(idx[q * k + j]), int(vals[q * k + j])) for j in range(k)] for q in range(len(encoded))]
