"""Placeholders so `from Bio import bgzf` imports; the .kdb container is outside the hot path."""


class BgzfWriter:  # pragma: no cover
    def __init__(self, *a, **k):
        raise NotImplementedError("Bio shim: bgzf is not provided")


class BgzfReader(BgzfWriter):
    pass


def BgzfBlocks(handle):  # pragma: no cover
    raise NotImplementedError("Bio shim: bgzf is not provided")
