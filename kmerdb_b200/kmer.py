"""Host-side mirror of the scalar helpers of kmerdb/kmer.py used on the `profile` path.

  kmer_to_id / id_to_kmer : single-k-mer helpers (kmerdb/kmer.py:234-317, :320-363).  One k-mer is
      not device work; these are plain integer arithmetic with the reference's argument checks,
      return values and exception classes.
  kmer_to_id_batch        : the batched form, on the GPU (kdb_kmer_to_id_batch).
  shred                   : kmerdb/kmer.py:489-577 for one record; the window ids come from the
      GPU batch kernel (there is no CPU id loop here).

Only the nucleic-acid branch exists: the reference's amino-acid shred is `RuntimeError
("Unimplemented")` (kmer.py:567-568) and is out of scope."""
from itertools import product

import numpy as np

from . import _native

# kmerdb/kmer.py:44-49 (A0 C1 G2 T3, first letter most significant) and :51-56
letterToBinaryNA = {65: 0, 67: 1, 71: 2, 84: 3}
binaryToLetterNA = ["A", "C", "G", "T"]
standard_lettersNA = set("ACTG")                       # kmer.py:58
IUPAC_NA_DOUBLETS = set("RYSWKM")                      # kmer.py:60-68
IUPAC_NA_TRIPLETS = set("BDHV")                        # kmer.py:70-76
permitted_NA_characters = standard_lettersNA | IUPAC_NA_DOUBLETS | IUPAC_NA_TRIPLETS | {"N"}   # kmer.py:78-80

_COMPLEMENT = {"A": "T", "C": "G", "G": "C", "T": "A"}
_NONE_ID = np.uint64(2 ** 64 - 1)
_KEYERR_ID = np.uint64(2 ** 64 - 2)


def _check_alphabet(s):
    """kmer.py:149-170: a letter outside ACGTRYSWKMBDHVN dies with NameError (the reference hits an
    undefined name on that path; the class is what callers and the CLI's exit-code map observe)."""
    if not permitted_NA_characters.issuperset(s):
        raise NameError("name 'permitted_AA_characters_extended' is not defined")


def kmer_to_id(s, is_aa=False, canonicalize=True):
    """kmerdb/kmer.py:234-317.  None for a k-mer holding N; KeyError for RYSWKMBDHV."""
    if hasattr(s, "seq"):
        s = s.seq
    if type(s) is not str:
        s = str(s) if s.__class__.__name__ == "Seq" else s
        if type(s) is not str:
            raise TypeError("kmerdb.kmer.kmer_to_id expects a str or Bio.SeqRecord.SeqRecord as its positional argument")
    if is_aa:
        raise ValueError("kmerdb_b200.kmer.kmer_to_id: amino-acid ids are out of scope of the profile path")
    if "N" in s:                                           # kmer.py:287-289
        return None
    _check_alphabet(s)                                     # kmer.py:290
    fwd = rc = 0
    for j, ch in enumerate(s):
        c = ord(ch)
        if c not in letterToBinaryNA:
            raise KeyError(c)                              # kmer.py:309
        code = letterToBinaryNA[c]
        fwd = (fwd << 2) | code                            # kmer.py:307-309
        rc |= (3 - code) << (2 * j)                        # kmer.py:310-312 (id of the reverse complement)
    return min(fwd, rc) if canonicalize else fwd           # kmer.py:314-317


def id_to_kmer(id, k, is_aa=False):
    """kmerdb/kmer.py:320-363."""
    if type(id) is not int:
        raise TypeError("kmerdb.id_to_kmer expects an int as its first positional argument")
    elif type(k) is not int:
        raise TypeError("kmerdb.id_to_kmer expects an int as its second positional argument")
    if is_aa:
        raise ValueError("kmerdb_b200.kmer.id_to_kmer: amino-acid ids are out of scope of the profile path")
    if k == 0:
        raise ValueError("kmer.id_to_kmer returned an empty k-mer. The kmer-id was '{0}', the choice of k is {1}".format(id, k))
    letters = []
    for _ in range(k):
        letters.append(binaryToLetterNA[id & 0x03])
        id >>= 2
    return "".join(reversed(letters))


def kmer_to_id_batch(kmers, k, canonicalize=True, device=0):
    """ids of n k-mers packed back to back (bytes of length n*k), computed on the GPU.
    Returns a list of int / None (N) and raises KeyError for RYSWKMBDHV like the scalar call."""
    ids = _native.kmer_to_id_batch(kmers, k, canonical=canonicalize, device=device)
    if (ids == _KEYERR_ID).any():
        i = int(np.nonzero(ids == _KEYERR_ID)[0][0])
        bad = [c for c in kmers[i * k:(i + 1) * k] if c not in letterToBinaryNA and c != ord("N")]
        raise KeyError(bad[0] if bad else None)
    return [None if x == _NONE_ID else int(x) for x in ids]


def reverse_complement(s):
    return "".join(_COMPLEMENT[c] for c in reversed(s))


def substitute_residue_with_chars(kmer, char, substitutions):
    """kmerdb/kmer.py:586-621 (only the 'N' branch is reachable without a crash):
    every occurrence of `char` replaced combinatorially, first occurrence varying slowest."""
    if type(kmer) is not str:
        raise TypeError("kmerdb.kmer.substitute_residue_with_chars() expects a str as its first argument")
    elif type(char) is not str or len(char) > 1:
        raise TypeError("kmerdb.kmer.substitute_residue_with_chars() expects a single character as its second argument")
    elif type(substitutions) is not list or not all(type(c) is str and len(c) == 1 for c in substitutions):
        raise TypeError("kmerdb.kmer.substitute_residue_with_chars() expects a list of single characters as its third argument")
    n = kmer.count(char)
    if n == 0:
        return [kmer]
    parts = kmer.split(char)
    out = []
    for combo in product(substitutions, repeat=n):
        out.append("".join(p + c for p, c in zip(parts, combo + ("",))))
    return out


def shred(seqRecord, k, replace_with_none=False, canonicalize=True, quiet_iupac_warning=True, device=0):
    """kmerdb/kmer.py:489-577 for one record: (kmer_ids, seq_ids, positions).

    Clean windows give one id each; a window with N is dropped (`replace_with_none`) or expanded
    into its 4^n ACGT substitutions (kmer.py:550-565).  The reference iterates a Python set for
    the substitution letters, so the order of the expansions inside one position is not defined
    there; here it is A,C,G,T with the first N varying slowest."""
    if type(k) is not int:
        raise TypeError("kmerdb.kmer.shred() expects an int as its second positional argument")
    seq = str(seqRecord.seq) if hasattr(seqRecord, "seq") else str(seqRecord)
    seq_id = getattr(seqRecord, "id", None)
    _check_alphabet(seq)                                   # kmer.py:519 is_sequence_na over the record
    if len(seq) < k:                                       # kmer.py:461-463; a Seq has no .id -> AttributeError
        raise AttributeError("'Seq' object has no attribute 'id'")
    windows, pos = [], []
    for i in range(len(seq) - k + 1):
        w = seq[i:i + k]
        if "N" in w:
            if replace_with_none:
                continue                                   # kmer.py:542-544
            subs = substitute_residue_with_chars(w, "N", ["A", "C", "G", "T"])
            windows.extend(subs)
            pos.extend([i] * len(subs))
        else:
            windows.append(w)
            pos.append(i)
    ids = kmer_to_id_batch("".join(windows).encode("ascii"), k, canonicalize, device) if windows else []
    return ids, [seq_id] * len(ids), pos
