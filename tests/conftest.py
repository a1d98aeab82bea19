import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle():
    from oracle import pyoracle as po
    po.build(ref=True)
    return po.Oracle()


@pytest.fixture(scope="session")
def golden():
    with open(os.path.join(ROOT, "tests", "golden", "phase1_ref_shares.json")) as f:
        g = json.load(f)
    for c in g["cases"]:
        c["X"] = np.array([float(v) for v in c["X"]]).reshape(c["n"], c["d"])
        c["y"] = np.array([float(v) for v in c["y"]])
        c["shares"] = np.array([[int(v, 16) for v in row] for row in c["shares"]], dtype=np.uint64)
    g["ti_key"] = bytes.fromhex(g["ti_key"])
    g["ot_key"] = bytes.fromhex(g["ot_key"])
    return g


@pytest.fixture(scope="session")
def p1lib():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    import linreg_mpc_b200 as pkg
    return pkg.P1Lib()  # raises if libp1gpu.so is missing: no fallback
