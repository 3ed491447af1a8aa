"""Whole-file parity of the CLI mirror with the UNMODIFIED reference subcommands (tests/golden/make_golden_cli.py ->
golden_cli.json): the gunzipped text of every .kdb / .kdbg (YAML header + rows) must be byte-identical, and failing
inputs must raise the reference's exception class without leaving an output file."""
import argparse
import base64
import builtins
import gzip
import json
import os

import pytest

from kmerdb_b200 import cli

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = json.load(open(os.path.join(HERE, "golden", "golden_cli.json")))["cases"]
FUNCS = {"profile": cli.profile, "make_graph": cli.make_graph}


def test_golden_cli_file():
    assert len(GOLD) >= 13 and {c["func"] for c in GOLD} == {"profile", "make_graph"}


@pytest.mark.gpu
@pytest.mark.parametrize("case", GOLD, ids=[c["name"] for c in GOLD])
def test_cli_output_is_byte_identical(case, tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    for fn, b64 in case["files"].items():
        with open(fn, "wb") as fh:
            fh.write(base64.b64decode(b64))
    ns = argparse.Namespace(device=0, **case["args"])
    if "raises" in case:
        with pytest.raises(getattr(builtins, case["raises"])):
            FUNCS[case["func"]](ns)
        assert not [f for f in os.listdir(".") if f.endswith(".kdb") or f.endswith(".kdbg")]      # no partial output
        return
    FUNCS[case["func"]](ns)
    for out, want in case["outputs"].items():
        got = gzip.open(out, "rt").read()
        if got != want:
            gl, wl = got.splitlines(), want.splitlines()
            diff = [(i, a, b) for i, (a, b) in enumerate(zip(gl, wl)) if a != b][:5]
            raise AssertionError("%s differs: %d vs %d lines; first differences %r" % (out, len(gl), len(wl), diff))
