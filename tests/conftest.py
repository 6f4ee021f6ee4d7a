import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def golden_search():
    z = np.load(os.path.join(GOLDEN, "search_cases.npz"))
    names = sorted({k.split("/")[0] for k in z.files})
    return {n: {k.split("/")[1]: z[k] for k in z.files if k.startswith(n + "/")} for n in names}


@pytest.fixture(scope="session")
def golden_edge():
    return dict(np.load(os.path.join(GOLDEN, "edge_cases.npz")))


@pytest.fixture(scope="session")
def golden_config1():
    return dict(np.load(os.path.join(GOLDEN, "config1_lcsm.npz")))


def case_inputs(spec):
    """Rebuild (db, queries, k) of a golden search case from its spec (see make_golden.py)."""
    import synth
    n, d, k, nq, seed, dist = [int(v) for v in spec]
    dist = "rand" if dist == 0 else "randn"
    db = synth.unit_rows(seed, n, d, dist)
    qs = synth.queries_for(seed + 1000, db, nq, dist)
    return db, qs, k


def config1_inputs(n=1000, d=128):
    import synth
    local = synth.unit_rows(1, n, d, "rand")
    other = synth.unit_rows(2, n, d, "rand")
    eps = synth.normal(3, (n, d)).astype(np.float32)
    for j in range(0, n, 10):
        v = local[(3 * j) % n] + 0.05 * eps[j]
        other[j] = v / np.linalg.norm(v)
    return local, other
