"""Host scalar helpers of kmerdb_b200.kmer against the golden vectors of the unmodified reference."""
import pytest

from kmerdb_b200 import kmer


def test_kmer_to_id_golden(golden):
    s = golden["scalars"]
    two = [a + b for a in "ACGT" for b in "ACGT"]
    assert [kmer.kmer_to_id(x, canonicalize=False) for x in two] == s["kmer_to_id_2mers_noncanonical"]
    assert [kmer.kmer_to_id(x) for x in two] == s["kmer_to_id_2mers_canonical"]
    assert kmer.kmer_to_id("ATCNATC") is None
    for km, (canon, fwd) in s["kmer_to_id_misc"].items():
        assert kmer.kmer_to_id(km) == canon and kmer.kmer_to_id(km, canonicalize=False) == fwd


def test_id_to_kmer_golden(golden):
    for key, exp in golden["scalars"]["id_to_kmer"].items():
        i, k = key.split(",")
        assert kmer.id_to_kmer(int(i), int(k)) == exp
        assert kmer.kmer_to_id(exp, canonicalize=False) == int(i)


def test_exception_classes():
    with pytest.raises(TypeError):
        kmer.kmer_to_id(7)
    with pytest.raises(TypeError):
        kmer.id_to_kmer("7", 3)
    with pytest.raises(KeyError):
        kmer.kmer_to_id("ACGRACG")          # kmer.py:309
    with pytest.raises(NameError):
        kmer.kmer_to_id("acgt")             # kmer.py:170


def test_substitute_residue():
    assert kmer.substitute_residue_with_chars("ANC", "N", list("ACGT")) == ["AAC", "ACC", "AGC", "ATC"]
    assert len(kmer.substitute_residue_with_chars("NNA", "N", list("ACGT"))) == 16
    assert kmer.substitute_residue_with_chars("ACG", "N", list("ACGT")) == ["ACG"]
