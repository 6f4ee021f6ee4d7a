"""CPU: the ScanContext oracle (oracle/sc_oracle.py) against golden vectors made by running the reference
(tests/golden/make_golden_sc.py), and the library's ScanContext symbols."""
import os
import re

import numpy as np
import pytest

from conftest import GOLDEN, ROOT
import synth
from oracle import sc_oracle


@pytest.fixture(scope="module")
def golden_sc():
    return dict(np.load(os.path.join(GOLDEN, "sc_cases.npz")))


@pytest.mark.parametrize("name", ["n7", "n120", "n1300"])
def test_oracle_search_matches_reference(golden_sc, name):
    n, nq, seed = [int(v) for v in golden_sc[name + "/spec"]]
    db = synth.scancontexts(seed, n)
    qs = synth.sc_queries_for(seed + 1000, db, nq)
    items = {i: 5000 + i for i in range(n)}
    # ring keys are bit-identical (same np.mean call)
    assert np.array_equal(np.mean(db.reshape(n, 20, 60), axis=2), golden_sc[name + "/ringkeys"])
    for j in range(nq):
        it, s = sc_oracle.search(db, items, qs[j])
        assert it[0] == golden_sc[name + "/items"][j]
        assert abs(s[0] - golden_sc[name + "/sims"][j]) <= 1e-12


def test_oracle_distance_matches_reference(golden_sc):
    db = synth.scancontexts(41, 6)
    qs = synth.sc_queries_for(1041, db, 4)
    db[5] = 0.0
    for (a, b), (dist, yaw) in zip(golden_sc["dist/pairs"], golden_sc["dist/values"]):
        d, y = sc_oracle.distance_sc(db[a].reshape(20, 60), qs[b].reshape(20, 60))
        assert abs(d - dist) <= 1e-12 and y == int(yaw)


def test_library_exports_every_scancontext_symbol():
    from cslam_b200 import _native
    text = open(os.path.join(ROOT, "include", "cslam_scancontext.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    declared = sorted(set(re.findall(r"\b(cslam_sc_[a-z_]+)\s*\(", text)))
    lib = _native.load()
    assert len(declared) == 8
    for name in declared:
        assert hasattr(lib, name), name
    assert set(declared) == set(_native.SC_SIGNATURES)
