"""Minimal Bio.Seq.Seq: an immutable str wrapper with slicing, iteration, find/count, equality
and reverse_complement (IUPAC complement table as in Bio.Data.IUPACData.ambiguous_dna_complement)."""
_COMP = str.maketrans("ACGTRYSWKMBDHVNacgtryswkmbdhvn", "TGCAYRSWMKVHDBNtgcayrswmkvhdbn")


class Seq:
    def __init__(self, data):
        self._data = str(data)

    def __str__(self):
        return self._data

    def __repr__(self):
        return "Seq(%r)" % self._data

    def __len__(self):
        return len(self._data)

    def __iter__(self):
        return iter(self._data)

    def __hash__(self):
        return hash(self._data)

    def __eq__(self, other):
        return self._data == str(other)

    def __getitem__(self, idx):
        r = self._data[idx]
        return Seq(r) if isinstance(idx, slice) else r

    def __add__(self, other):
        return Seq(self._data + str(other))

    def __radd__(self, other):
        return Seq(str(other) + self._data)

    def __contains__(self, item):
        return str(item) in self._data

    def find(self, sub, *a):
        return self._data.find(str(sub), *a)

    def count(self, sub, *a):
        return self._data.count(str(sub), *a)

    def upper(self):
        return Seq(self._data.upper())

    def reverse_complement(self):
        return Seq(self._data.translate(_COMP)[::-1])
