"""Host mirror of the Pearson / Spearman coefficients of `kmerdb distance`.

  correlation(a, b)        : kmerdb/distance.pyx:108-152 (`correlation(a, b, total_kmers)`), on the GPU
  pearson_matrix(profiles) : the pair loop of kmerdb/__init__.py:718-771 for metric == "pearson" in one call
  spearman(x, y)           : kmerdb/python_distances.py:96-116 — scipy.stats.spearmanr is the Pearson coefficient of the
                             average ranks; ranks stay host work (scipy.stats.rankdata), the contraction runs on the GPU
                             on 2 * rank (integers; the coefficient is scale invariant)

`profiles` are uint64 count vectors, one per sample."""
import numpy as np

from . import _native


def correlation(a, b, total_kmers=None, device=0, exact=False):
    """Pearson r of two count vectors.  Default: the reference's arithmetic (float32 mean and residuals, distance.pyx:118-126);
    exact=True: float64 throughout."""
    a = np.asarray(a)
    b = np.asarray(b)
    if a.shape != b.shape or a.ndim != 1:
        raise TypeError("kmerdb.distance.correlation expects two equal size NumPy arrays/vectors as its first and second positional arguments.")
    if total_kmers is not None and (total_kmers != len(a) or total_kmers != len(b)):
        raise ValueError("NumPy kmer count array total does not match length of arrays")        # distance.pyx:120-121
    return float(_native.correlation_matrix(np.stack([a, b]).astype(np.uint64, copy=False), mode=1 if exact else 0, device=device)[0, 1])


def pearson_matrix(profiles, device=0, exact=False):
    """n x n matrix of `correlation` over the rows of `profiles`; the diagonal is the metric's identity value 1.0
    (kmerdb/__init__.py:727-728)."""
    m = _native.correlation_matrix(np.asarray(profiles, dtype=np.uint64), mode=1 if exact else 0, device=device)
    np.fill_diagonal(m, 1.0)
    return m


def spearman(x, y, device=0):
    """Spearman rank correlation (the `cor` of scipy.stats.spearmanr(x, b=y), python_distances.py:113)."""
    from scipy.stats import rankdata
    if type(x) is not np.ndarray:
        raise TypeError("kmerdb.distance.spearman expects a Numpy array as its first positional argument")
    elif type(y) is not np.ndarray:
        raise TypeError("kmerdb.distance.spearman expects a Numpy array as its second positional argument")
    rx = np.rint(2 * rankdata(x)).astype(np.uint64)
    ry = np.rint(2 * rankdata(y)).astype(np.uint64)
    return float(_native.correlation_matrix(np.stack([rx, ry]), mode=1, device=device)[0, 1])
