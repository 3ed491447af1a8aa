"""Host-side mirror of the edge builder of kmerdb/graph.py for the `graph` subcommand.

  make_edges_from_fasta : kmerdb/graph.py:219-332 — (edge rows, file_metadata, counts).  The host tokenises
      the file (parse.parse_sequence_file, unchanged host code as in graph.py:261); ids, edge rows and the
      k-mer counts come from the CUDA kernels (kdb_edge_rows, the profile kernels).
  edge_weights          : the aggregated (n1, n2, weight) list (legacy .kdbg columns) computed at
      scale from the dense forward (k+1)-mer histogram (kdb_edge_weights) — no per-edge rows are materialised.
Any N raises ValueError in both modes, like the reference."""
import os

import numpy as np

from . import _native, parse


def _raise_graph_status(status, rec):
    if status == _native.KDB_E_ALPHABET:
        raise NameError("name 'permitted_AA_characters_extended' is not defined")        # kmer.py:170
    if status == _native.KDB_E_SHORT:
        raise AttributeError("'Seq' object has no attribute 'id'")                        # kmer.py:461-463
    if status == _native.KDB_E_FORMAT:
        raise ValueError("kmerdb.graph: record %d holds 'N'; ambiguous residues cannot form edges" % rec)
    if status == _native.KDB_E_IUPAC:
        raise KeyError("non-ACGT IUPAC letter in record %d reached the 2-bit encoder" % rec)   # kmer.py:309
    if status == _native.KDB_E_EMPTY:
        raise ValueError("max() arg is an empty sequence")                                # graph.py:314
    if status:
        raise RuntimeError("kmerdb_b200.graph: unexpected device status %d" % status)


def make_edges_from_fasta(filename, k, quiet=True, canonicalize=True, replace_with_none=False, device=0):
    """kmerdb/graph.py:219-332.  Returns (data, file_metadata, counts): data = list of 5-tuples
    (seq_id, pos1, kmer_id1, pos2, kmer_id2) in read order."""
    if type(filename) is not str:
        raise TypeError("kmerdb.graph.make_coo_graph() expects a fasta/fastq sequence filepath as a str as its first positional_regument")
    elif type(k) is not int:
        raise TypeError("kmerdb.graph.make_coo_graph() expects an int for k as the second positional argument")
    elif type(quiet) is not bool:
        raise TypeError("kmerdb.graph.make_coo_graph() expects the keyword argument 'slurp' to be a bool")
    if os.path.exists(filename) is False or os.access(filename, os.R_OK) is False:
        raise ValueError("kmerdb.graph.make_coo_graph() expects the filepath to be be readable on the filesystem")
    recs = list(parse.parse_sequence_file(filename, return_tuple=True))
    ids = [r[0] for r in recs]
    lens = np.array([len(r[1]) for r in recs], dtype=np.uint64)
    rec_off = np.concatenate([np.zeros(1, dtype=np.uint64), np.cumsum(lens, dtype=np.uint64)])
    letters = "".join(r[1] for r in recs).encode("utf-8", "replace")
    if len(letters) != int(rec_off[-1]):                   # non-ASCII letters: one '?' per character keeps the offsets
        letters = "".join(ch if ord(ch) < 128 else "?" for r in recs for ch in r[1]).encode("ascii")
    status, err_rec, rows = _native.edge_rows(letters, rec_off, k, canonical=canonicalize, device=device)
    _raise_graph_status(status, err_rec)
    # counts[kmer_id] += 1 (graph.py:281-283): the same records through the profile kernels
    plain = b"".join(b">\n" + r[1].encode("ascii") + b"\n" for r in recs)
    with _native.Profiler(k, canonical=canonicalize, skip_n=True, fastq=False, device=device) as p:
        p.feed(plain)
        st_code = p.finish()
        st = p.stats()
        parse._raise_for_status(st_code, st)
        counts = p.counts()
    N = 4 ** k
    unique = st["unique_kmers"]
    md5, sha256 = parse.checksum(filename)
    file_metadata = {
        "filename": filename, "md5": md5, "sha256": sha256,
        "total_reads": len(recs), "total_kmers": st["total_kmers"], "unique_kmers": unique,
        "nullomers": N - unique if canonicalize is False else int((N / 2) - unique),      # graph.py:312
        "num_reads": len(recs), "min_read_length": int(lens.min()), "max_read_length": int(lens.max()),
        "avg_read_length": int(np.mean(lens)),
    }
    data = [(ids[int(r)], int(p1), int(i1), int(p2), int(i2)) for r, p1, i1, p2, i2 in rows]
    return data, file_metadata, counts


def edge_weights_bytes(data, k, fmt="fasta", canonicalize=True, device=0):
    """Aggregated edge list of uncompressed FASTA/FASTQ text: (n1, n2, weight) uint64 arrays.
    weight = number of times id n2 directly follows id n1 inside a record."""
    with _native.Profiler(k + 1, canonical=False, skip_n=True, fastq=(fmt == "fastq"), device=device, edge_kplus1=True) as p:
        p.feed(data)
        status = p.finish()
        st = p.stats()
        if status == _native.KDB_E_LAYOUT:
            plain, pending = parse._plain_layout(bytes(data), fmt)
            if pending is not None:
                raise pending
            return edge_weights_bytes(plain, k, "fasta", canonicalize, device)
        parse._raise_for_status(status, st)
        if st["sum_len"] != sum(st["mono"]):               # a letter that is not A/C/G/T survived the raises: N
            raise ValueError("kmerdb.graph: the input holds 'N'; ambiguous residues cannot form edges")
        return p.edge_weights(canonical=canonicalize)


def edge_weights(filename, k, canonicalize=True, device=0):
    """Aggregated (n1, n2, weight) edge list of a sequence file (legacy .kdbg body)."""
    data = parse._read_text_bytes(filename)
    return edge_weights_bytes(data, k, parse._file_format(filename), canonicalize, device)
