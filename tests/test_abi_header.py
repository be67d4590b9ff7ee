"""The C header is the contract: every exported b200dp_* symbol is declared in include/b200dp.h and every
declaration is exported by the built library (skipped when the library has not been built)."""
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "neural-networks-parallel-training-with-mpi_b200", "lib", "libb200dp.so")


@pytest.mark.skipif(not os.path.exists(LIB), reason="libb200dp.so not built")
def test_header_and_library_agree():
    out = subprocess.run(["nm", "-D", "--defined-only", LIB], capture_output=True, text=True).stdout
    exported = {l.split()[-1] for l in out.splitlines() if " T b200dp_" in l}
    header = open(os.path.join(ROOT, "include", "b200dp.h")).read()
    declared = set(re.findall(r"\b(b200dp_[a-z0-9_]+)\s*\(", header))
    assert exported == declared, (sorted(exported - declared), sorted(declared - exported))


def test_every_entry_point_cites_the_reference_or_says_it_has_no_counterpart():
    header = open(os.path.join(ROOT, "include", "b200dp.h")).read()
    assert header.count("ref.py:") >= 10
    assert "no counterpart in the reference" in header
