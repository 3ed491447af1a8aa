"""Small end-to-end case for compute-sanitizer (one tool per gpurun call): FASTA + FASTQ + N expansion + sweep + K4."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from kmerdb_b200 import _native, parse, graph, minimizer
rng = np.random.default_rng(1)
def fasta(n, alphabet="ACGT", p=None, width=70):
    out = []
    for i in range(n):
        L = int(rng.integers(20, 400)); s = "".join(np.array(list(alphabet))[rng.choice(len(alphabet), size=L, p=p)])
        out.append(">r%d\n" % i + "\n".join(s[j:j + width] for j in range(0, L, width)) + "\n")
    return "".join(out).encode()
t = fasta(300)
for k in (5, 9, 12, 13):
    pr = parse.profile_bytes(t, k)
    assert pr.total_kmers > 0
tn = fasta(200, "ACGTN", [0.245, 0.245, 0.245, 0.245, 0.02])
parse.profile_bytes(tn, 8, replace_with_none=False)
fq = "".join("@q%d\n%s\n+\n%s\n" % (i, "".join(np.array(list("ACGT"))[rng.integers(0, 4, size=100)]), "I" * 100) for i in range(300)).encode()
parse.profile_bytes(fq, 12, fmt="fastq")
parse.profile_multik(_native.synth_host(3000), 4, 13)
graph.edge_weights_bytes(t, 6)
minimizer.windowed_minimizers("ACGTTGCAACGGTACGTTGCAACGGTTTGACCATGACAGTAGGGATACC" * 5, 7, 10)
print("sanitize case done")
