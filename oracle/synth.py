"""TEST INFRASTRUCTURE: NumPy restatement of the synthetic benchmark stream (the product's generator is
kmerdb_b200/csrc/kdb_synth.cuh): reads of i.i.d. uniform ACGT letters from Philox-4x32-10, key = seed,
counter = (read index lo, read index hi, 64-letter block, 0), letter j of a block = 2 bits of word (j & 63) >> 4 at
bit 2 * (j & 15).  FASTA framing '>%09d\\n' + letters + '\\n'.  Used by bench.py's CPU legs so that the reference arm
never loads libkmerdb_b200.so; tests/test_oracle.py checks it byte for byte against kdb_synth_host."""
import numpy as np

_M0, _M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
_W0, _W1 = 0x9E3779B9, 0xBB67AE85
_MASK = np.uint64(0xFFFFFFFF)


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    c0, c1, c2, c3 = (np.asarray(x, dtype=np.uint64) for x in (c0, c1, c2, c3))
    for r in range(10):
        p0, p1 = _M0 * c0, _M1 * c2
        kk0, kk1 = np.uint64((k0 + r * _W0) & 0xFFFFFFFF), np.uint64((k1 + r * _W1) & 0xFFFFFFFF)
        c0, c1, c2, c3 = (p1 >> np.uint64(32)) ^ c1 ^ kk0, p1 & _MASK, (p0 >> np.uint64(32)) ^ c3 ^ kk1, p0 & _MASK
    return c0, c1, c2, c3


def reads(first_read, n_reads, read_len=150, seed=20261104):
    """uint8[n_reads, read_len] of ASCII letters."""
    idx = np.arange(first_read, first_read + n_reads, dtype=np.uint64)
    n_blk = (read_len + 63) // 64
    out = np.empty((n_reads, n_blk * 64), dtype=np.uint8)
    lut = np.frombuffer(b"ACGT", dtype=np.uint8)
    sh = (2 * np.arange(16, dtype=np.uint64))[None, :]
    for b in range(n_blk):
        words = philox4x32_10(idx & _MASK, idx >> np.uint64(32), np.full(n_reads, b, dtype=np.uint64), np.zeros(n_reads, dtype=np.uint64),
                              seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
        for w in range(4):
            codes = ((words[w][:, None] >> sh) & np.uint64(3)).astype(np.intp)
            out[:, 64 * b + 16 * w: 64 * b + 16 * w + 16] = lut[codes]
    return out[:, :read_len]


def fasta(first_read, n_reads, read_len=150, seed=20261104):
    """The FASTA text of reads [first_read, first_read + n_reads) as bytes: 11 + read_len + 1 bytes per record."""
    rb = 11 + read_len + 1
    out = np.empty((n_reads, rb), dtype=np.uint8)
    out[:, 0] = ord(">")
    v = (np.arange(first_read, first_read + n_reads, dtype=np.uint64) % np.uint64(1000000000)).astype(np.int64)
    for d in range(9):
        out[:, 1 + d] = ord("0") + (v // 10 ** (8 - d)) % 10
    out[:, 10] = 10
    out[:, 11:11 + read_len] = reads(first_read, n_reads, read_len, seed)
    out[:, 11 + read_len] = 10
    return out.tobytes()
