"""Host-side mirror of kmerdb/minimizer.py for the `minimizers` subcommand.

  select_lexicographical_minimizers : kmerdb/minimizer.py:31-85, reference-compatible rows
      (seq_id, i, canonical kmer_id, i % window_size == 0) for i < 4^k - k + 1, including the
      ValueError("WHOOPS") the reference raises for every sequence shorter than 4^k.
  windowed_minimizers               : the standard definition (extension) on the GPU:
      minimum canonical id among w consecutive k-mers, leftmost on ties, by warp shuffles.
Ids come from the CUDA kernels (kdb_minimizers_compat / kdb_minimizers); nothing is computed on the CPU."""
from . import _native


def select_lexicographical_minimizers(seq, seq_id, k, window_size, device=0):
    if type(seq) is not str:
        raise TypeError("kmerdb.minimizer.select_lexicographical_minimizers() expects a str as its first positional argument")
    elif type(seq_id) is not str:
        raise TypeError("kmerdb.minimizer.select_lexicographical_minimizers() expects a str as its second positional argument")
    elif type(k) is not int:
        raise TypeError("kmerdb.minimizer.select_lexicographical_minimizers() expects an int as its third positional argument")
    elif type(window_size) is not int:
        raise TypeError("kmerdb.minimizer.select_lexicographical_minimizers() expects an int as its fourth positional argument")
    if len(seq) < k:
        raise ValueError("Sequence length was less than k")                     # minimizer.py:59-60
    status, ids, flags = _native.minimizers_compat(seq, k, window_size, device=device)
    if status == _native.KDB_E_FORMAT:
        raise ValueError("WHOOPS")                                              # minimizer.py:76-78
    if status == _native.KDB_E_ALPHABET:
        raise NameError("name 'permitted_AA_characters_extended' is not defined")
    if status == _native.KDB_E_IUPAC:
        raise KeyError("non-ACGT IUPAC letter reached the 2-bit encoder")
    none = 2 ** 64 - 1
    return [(seq_id, i, None if int(v) == none else int(v), bool(f)) for i, (v, f) in enumerate(zip(ids, flags))]


def windowed_minimizers(seq, k, w, device=0):
    """[(position, canonical kmer_id)] of the true (w, k) minimizers of one sequence."""
    status, pos, ids = _native.minimizers(seq, k, w, device=device)
    if status == _native.KDB_E_ALPHABET:
        raise NameError("name 'permitted_AA_characters_extended' is not defined")
    if status == _native.KDB_E_IUPAC:
        raise KeyError("ambiguous letter inside a window")
    return [(int(p), int(i)) for p, i in zip(pos, ids)]


def windowed_minimizers_file(fasta_file, k, w, device=0):
    """True (w, k) minimizers of every record of a FASTA/FASTQ file in one device call:
    list of (seq_id, position, canonical kmer_id), record-major."""
    import numpy as np
    from . import parse
    recs = list(parse.parse_sequence_file(fasta_file))
    lens = np.array([len(s) for _, s in recs], dtype=np.uint64)
    off = np.concatenate([np.zeros(1, dtype=np.uint64), np.cumsum(lens, dtype=np.uint64)])
    status, rows = _native.minimizers_batch("".join(s for _, s in recs).encode("ascii", "replace"), off, k, w, device=device)
    if status == _native.KDB_E_ALPHABET:
        raise NameError("name 'permitted_AA_characters_extended' is not defined")
    if status == _native.KDB_E_IUPAC:
        raise KeyError("ambiguous letter inside a window")
    return [(recs[int(r)][0], int(p), int(i)) for r, p, i in rows]
