import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _build_oracle():
    """Build the oracle's C restatement (and the reference build when /root/reference is present)."""
    import subprocess
    odir = os.path.join(ROOT, "oracle")
    if not os.path.exists(os.path.join(odir, "libgloracle.so")):
        subprocess.run(["make", "-C", odir, "port"], check=True, capture_output=True)
    if os.path.isdir("/root/reference/src/goldilocks/src") and not os.path.exists(os.path.join(odir, "_ref", "libglref.so")):
        subprocess.run(["make", "-C", odir, "ref"], check=True, capture_output=True)


@pytest.fixture(scope="session")
def port():
    _build_oracle()
    from oracle import refbind
    return refbind.port()


@pytest.fixture(scope="session")
def reference():
    """The compiled reference (oracle/_ref); tests that need it are skipped when it was never built."""
    _build_oracle()
    from oracle import refbind
    r = refbind.reference()
    if r is None:
        pytest.skip("oracle/_ref/libglref.so not built (needs /root/reference at build time)")
    return r


@pytest.fixture(scope="session")
def oracle(port):
    """Reference when available, else the pinned port."""
    from oracle import refbind
    return refbind.reference() or port


@pytest.fixture(scope="session")
def golden():
    import json
    with open(os.path.join(ROOT, "tests", "golden", "kat.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def fixtures():
    return np.load(os.path.join(ROOT, "tests", "golden", "fixtures.npz"))


@pytest.fixture(scope="session")
def ctx():
    """A context on cuda:0 of the real library.  Fails loudly if the extension is missing."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from xlayer_prover_b200 import Context
    c = Context(0)
    yield c
    c.close()


@pytest.fixture(scope="session")
def rng():
    return np.random.default_rng(20261104)
