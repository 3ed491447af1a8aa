"""Pearson / Spearman coefficients on the GPU (kdb_correlation_matrix, kdb_corr.cuh) against golden values of the
UNMODIFIED reference `distance.pyx` (tests/golden/make_golden_distance.py compiles it out of tree) and against numpy / scipy.
Floating point: the reference-arithmetic mode must agree with the reference to 1e-12 relative (both sum float32 products in
extended precision; the last bits depend on the order of the partial sums), the float64 mode with numpy to 1e-10."""
import json
import math
import os

import numpy as np
import pytest

from kmerdb_b200 import distance

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = json.load(open(os.path.join(HERE, "golden", "golden_distance.json")))["cases"]


def _profile(seed, n, kind):
    rng = np.random.default_rng(seed)
    if kind == "poisson":
        return rng.poisson(83.0, size=n).astype(np.uint64)
    if kind == "sparse":
        x = rng.poisson(0.3, size=n).astype(np.uint64)
        x[rng.integers(0, n, size=8)] = 100000
        return x
    if kind == "big":
        return rng.integers(0, 2 ** 40, size=n, dtype=np.uint64)
    if kind == "related":
        base = np.random.default_rng(seed // 10).poisson(200.0, size=n)
        return (base + rng.poisson(20.0, size=n)).astype(np.uint64)
    raise ValueError(kind)


def test_golden_distance_file():
    assert len(GOLD) >= 8 and all("r" in c for c in GOLD)


@pytest.mark.gpu
@pytest.mark.parametrize("case", GOLD, ids=["%s_k%d" % (c["kind"], c["k"]) for c in GOLD])
def test_correlation_matches_reference(case):
    n = 4 ** case["k"]
    if case["kind"] == "constant_vs_poisson":
        r = distance.correlation(np.full(n, 7, dtype=np.uint64), _profile(1, n, "poisson"), n)
        assert math.isnan(r)                                               # 0 / 0 in the reference's division
        return
    a, b = _profile(case["seeds"][0], n, case["kind"]), _profile(case["seeds"][1], n, case["kind"])
    assert int(a.sum()) == case["sum_a"] and int(b.sum()) == case["sum_b"]           # same inputs as the golden run
    r = distance.correlation(a, b, n)
    assert abs(r - case["r"]) <= 1e-12 * max(1.0, abs(case["r"])), (r, case["r"])
    exact = distance.correlation(a, b, n, exact=True)
    want = float(np.corrcoef(a.astype(np.float64), b.astype(np.float64))[0, 1])
    assert abs(exact - want) <= 1e-10 * max(1.0, abs(want)), (exact, want)


@pytest.mark.gpu
def test_pearson_matrix_and_errors():
    n = 4 ** 7
    rows = np.stack([_profile(s, n, "related") for s in (20, 21, 22, 35, 36, 47)])           # 6 samples: a ragged 4 x 4 block grid
    m = distance.pearson_matrix(rows)
    assert m.shape == (6, 6) and np.allclose(m, m.T, rtol=0, atol=0) and np.all(np.diag(m) == 1.0)
    for i in range(6):
        for j in range(i + 1, 6):
            assert m[i, j] == distance.correlation(rows[i], rows[j], n)
    assert m[0, 1] > 0.8 and abs(m[0, 3]) < 0.1                                             # same / different base profile
    with pytest.raises(ValueError):
        distance.correlation(rows[0], rows[1], n + 1)
    with pytest.raises(TypeError):
        distance.correlation(rows[0], rows[1][:-1])


@pytest.mark.gpu
def test_spearman_matches_scipy():
    from scipy.stats import spearmanr
    n = 4 ** 8
    x, y = _profile(5, n, "sparse"), _profile(6, n, "sparse")                               # heavy ties
    got = distance.spearman(x, y)
    want = float(spearmanr(x, b=y)[0])
    assert abs(got - want) <= 1e-12, (got, want)
    x, y = _profile(90, n, "related"), _profile(91, n, "related")
    assert abs(distance.spearman(x, y) - float(spearmanr(x, b=y)[0])) <= 1e-12
