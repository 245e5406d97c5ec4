import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def golden():
    with open(os.path.join(GOLDEN_DIR, "golden.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def golden_prog():
    """Progressive fixtures (tests/golden/prog, made by make_prog_golden.py): digests the REFERENCE BINARY produced on the
    baseline twin of each progressive file (same quantised coefficients, written as one scan)."""
    with open(os.path.join(GOLDEN_DIR, "golden_prog.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def oracle_mod():
    """The parity oracle (test infrastructure): C port always, reference binary when built."""
    from oracle import oracle
    oracle.build()
    return oracle


@pytest.fixture(scope="session")
def harness():
    from jpeg_decoder_b200 import harness as h
    h.build_harness()
    return h


@pytest.fixture(scope="session")
def lib_so():
    """libb200jpeg.so, (re)built in-tree with nvcc when stale (cross-compiles without a GPU)."""
    from jpeg_decoder_b200 import build
    return build.build_lib()


@pytest.fixture(scope="session")
def emul():
    """Kernel sources compiled against the SIMT emulation (tests/emul) -- CPU-tier logic checks."""
    from tests.emul import emul_harness
    emul_harness.build()
    return emul_harness


def fixture_bytes(name):
    with open(os.path.join(GOLDEN_DIR, name), "rb") as f:
        return f.read()
