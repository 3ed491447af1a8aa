"""The C-ABI library loads without a GPU and exports every entry point include/kmerdb_b200.h declares; the
product has no CPU fallback (no compute call is made here)."""
import os
import re

import pytest

from kmerdb_b200 import _native

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "kmerdb_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(kdb_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    L = _native.lib()
    names = _declared()
    assert len(names) >= 30
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing
    assert sorted(_native.EXPORTS) == names, "the ctypes binding list and the header disagree"


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "kmerdb_b200")
    for f in os.listdir(pkg):
        if f.endswith(".py"):
            src = open(os.path.join(pkg, f)).read()
            assert "import oracle" not in src and "from oracle" not in src, f


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(_native.NativeError):
        _native.Profiler(8)                      # kdb_create fails loudly: there is no CPU counting path
