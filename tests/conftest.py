import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session", autouse=True)
def _oracle_built():
    from oracle import oracle
    oracle.build()
    yield


@pytest.fixture(scope="session")
def golden():
    """name -> (npz arrays, json meta) of the committed reference-formula fixtures."""
    import json
    import numpy as np
    gdir = os.path.join(ROOT, "tests", "golden")
    out = {}
    for name in ("sequence_pool", "cvm", "fusion_seqpool_cvm_concat", "lookup_table_v2"):
        with open(os.path.join(gdir, name + ".json")) as f:
            meta = json.load(f)
        out[name] = (np.load(os.path.join(gdir, name + ".npz")), meta)
    return out
