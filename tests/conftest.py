import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    from oracle import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def ref():
    """The compiled reference header, or None where oracle/_ref was not built."""
    from oracle import Ref
    return Ref() if Ref.available() else None


@pytest.fixture(scope="session")
def readme_kats():
    """The worked examples of the reference's README (lines 100-410) in the KAT format."""
    with open(os.path.join(GOLDEN, "readme_examples.json")) as f:
        return json.load(f)["examples"]


@pytest.fixture(scope="session")
def kats():
    with open(os.path.join(GOLDEN, "kat_pint_test.json")) as f:
        return json.load(f)["kats"]


class RefVectors:
    def __init__(self, path):
        self.z = np.load(path)
        self.meta = json.loads(bytes(self.z["meta"]).decode())

    def layouts(self):
        return [(m["key"], m["elem_bytes"], tuple(m["bits"])) for m in self.meta]

    def __getitem__(self, k):
        return self.z[k]

    def has(self, k):
        return k in self.z.files


@pytest.fixture(scope="session")
def refvec():
    return RefVectors(os.path.join(GOLDEN, "ref_vectors.npz"))
