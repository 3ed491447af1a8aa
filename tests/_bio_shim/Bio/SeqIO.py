"""Minimal Bio.SeqIO.parse for "fasta" and "fastq", following Biopython 1.83's
SimpleFastaParser / FastqGeneralIterator behaviour as documented (restated, not copied):
FASTA: text before the first '>' line is ignored; id = first whitespace token of the title;
sequence = following lines right-stripped and joined, with ' ' and '\\r' removed.
FASTQ: '@' title, sequence line(s) up to the '+' line, quality line(s) until as many
quality characters as sequence letters were read; ValueError on the documented violations."""
from .Seq import Seq
from .SeqRecord import SeqRecord


def _fasta(handle):
    title = None
    lines = []
    for line in handle:
        if line[0] == ">":
            if title is not None:
                yield title, "".join(lines).replace(" ", "").replace("\r", "")
            title = line[1:].rstrip()
            lines = []
        elif title is not None:
            lines.append(line.rstrip())
    if title is not None:
        yield title, "".join(lines).replace(" ", "").replace("\r", "")


def _fastq(handle):
    line = handle.readline()
    while line:
        if line[0] != "@":
            raise ValueError("Records in Fastq files should start with '@' character")
        title = line[1:].rstrip()
        seq = handle.readline().rstrip()
        while True:
            line = handle.readline()
            if not line:
                raise ValueError("End of file without quality information.")
            if line[0] == "+":
                second = line[1:].rstrip()
                if second and second != title:
                    raise ValueError("Sequence and quality captions differ.")
                break
            seq += line.rstrip()
        if " " in seq or "\t" in seq:
            raise ValueError("Whitespace is not allowed in the sequence.")
        qual = handle.readline().rstrip()
        while True:
            line = handle.readline()
            if not line:
                break
            if line[0] == "@" and len(qual) >= len(seq):
                break
            qual += line.rstrip()
        if len(seq) != len(qual):
            raise ValueError("Lengths of sequence and quality values differs for %s (%i and %i)." % (title, len(seq), len(qual)))
        for ch in qual:
            if not 33 <= ord(ch) <= 126:
                raise ValueError("Invalid character in quality string")
        yield title, seq


def parse(handle, fmt):
    if fmt == "fasta":
        gen = _fasta(handle)
    elif fmt == "fastq":
        gen = _fastq(handle)
    else:
        raise ValueError("Bio shim supports only fasta/fastq, got %r" % (fmt,))
    for title, seq in gen:
        rid = title.split(None, 1)[0] if title.split() else ""
        yield SeqRecord(Seq(seq), id=rid, name=rid, description=title)
