#!/usr/bin/env python
"""Freeze golden Pearson coefficients from the UNMODIFIED reference `kmerdb/distance.pyx` (`correlation`, :108-152).
The .pyx is compiled OUT OF TREE (a scratch directory under /tmp; Cython + gcc of this image) from the source where it lies
under /root/reference — nothing of it is copied into the repo.  Run in the build container only:

    python tests/golden/make_golden_distance.py

Inputs are regenerated from seeds by the test (numpy Generator, PCG64), so the JSON only holds seeds, shapes and results."""
import json
import os
import shutil
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def build_reference(tmp):
    shutil.copy("/root/reference/kmerdb/distance.pyx", os.path.join(tmp, "distance.pyx"))      # scratch copy for the compiler only
    with open(os.path.join(tmp, "setup.py"), "w") as fh:
        fh.write("from setuptools import setup, Extension\nfrom Cython.Build import cythonize\nimport numpy\n"
                 "setup(ext_modules=cythonize([Extension('distance', ['distance.pyx'], include_dirs=[numpy.get_include()])], language_level=3))\n")
    subprocess.run([sys.executable, "setup.py", "build_ext", "--inplace"], cwd=tmp, check=True, capture_output=True)
    sys.path.insert(0, tmp)
    import distance
    return distance


def profile(seed, n, kind):
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


def main():
    with tempfile.TemporaryDirectory(prefix="kdist_") as tmp:
        ref = build_reference(tmp)
        cases = []
        for k, kind, seeds in ((6, "poisson", (1, 2)), (8, "poisson", (3, 4)), (8, "sparse", (5, 6)), (7, "big", (7, 8)), (9, "related", (90, 91)),
                               (10, "related", (100, 101)), (5, "sparse", (11, 12))):
            n = 4 ** k
            a, b = profile(seeds[0], n, kind), profile(seeds[1], n, kind)
            r = float(ref.correlation(a, b, n))
            cases.append({"k": k, "kind": kind, "seeds": list(seeds), "r": r, "sum_a": int(a.sum()), "sum_b": int(b.sum())})
            print(k, kind, seeds, repr(r), "exact:", repr(float(np.corrcoef(a.astype(float), b.astype(float))[0, 1])))
        const = np.full(4 ** 5, 7, dtype=np.uint64)
        r = float(ref.correlation(const, profile(1, 4 ** 5, "poisson"), 4 ** 5))
        cases.append({"k": 5, "kind": "constant_vs_poisson", "seeds": [0, 1], "r": None if r != r else r})
        print("constant:", r)
    import Cython
    with open(os.path.join(HERE, "golden_distance.json"), "w") as fh:
        json.dump({"source": "kmerdb v0.9.6 distance.pyx correlation(), compiled out of tree with Cython %s" % Cython.__version__, "cases": cases}, fh, indent=1)
    print("wrote golden_distance.json with", len(cases), "cases")


if __name__ == "__main__":
    main()
