"""kmerdb_b200 — B200-native engine for the `profile` hot path of kmerdb v0.9.6.

Mirrors the reference's interface for that path only:
  kmerdb_b200.parse.parsefile / parse_sequence_file   (kmerdb/parse.py:50-163)
  kmerdb_b200.kmer.kmer_to_id / id_to_kmer / shred    (kmerdb/kmer.py:234-577)
Counting runs on sm_100a CUDA kernels behind a C-ABI (include/kmerdb_b200.h); there is no CPU
fallback."""
VERSION = "0.1.0"
REFERENCE_VERSION = "0.9.6"   # kmerdb/config.py:20
