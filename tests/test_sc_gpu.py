"""GPU parity of the ScanContext matcher through the C-ABI: against the golden vectors made
by the reference and, on other seeds, against the oracle.  Tolerance: similarities within 1e-12 absolute
(double arithmetic on both sides; only the summation order inside a 20-element dot product differs),
items exact."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
import synth
from oracle import sc_oracle

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def SC():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    from cslam_b200.lidar_pr import ScanContextMatching
    return ScanContextMatching


def _fill(SC, db, first_item=5000):
    m = SC()
    for i in range(db.shape[0]):
        m.add_item(db[i], first_item + i)
    return m


@pytest.mark.parametrize("name", ["n7", "n120", "n1300"])
def test_golden_cases(SC, name):
    g = dict(np.load(os.path.join(GOLDEN, "sc_cases.npz")))
    n, nq, seed = [int(v) for v in g[name + "/spec"]]
    db = synth.scancontexts(seed, n)
    qs = synth.sc_queries_for(seed + 1000, db, nq)
    m = _fill(SC, db)
    assert m.nb_items == n
    for j in range(nq):
        it, s = m.search(qs[j], 1)
        assert it[0] == g[name + "/items"][j]
        assert abs(s[0] - g[name + "/sims"][j]) <= 1e-12
        best, sim = m.search_best(qs[j])
        assert best == it[0] and sim == s[0]


def test_candidates_against_oracle(SC):
    n = 2500   # crosses the 1000 -> 2000 -> 4000 capacity doublings
    db = synth.scancontexts(77, n)
    qs = synth.sc_queries_for(1077, db, 8)
    m = _fill(SC, db, 0)
    scs = db.reshape(n, 20, 60)
    rks = np.mean(scs, axis=2)
    for q in qs:
        rows, dists, yaws = m._candidates(q)
        want = sc_oracle.shortlist(rks, sc_oracle.sc2rk(q.reshape(20, 60)), 10)
        assert np.array_equal(rows, want)
        for r, d, y in zip(rows, dists, yaws):
            dd, yy = sc_oracle.distance_sc(scs[r], q.reshape(20, 60))
            assert abs(d - dd) <= 1e-12 and y == yy


def test_rotated_revisit_is_found_with_its_yaw(SC):
    db = synth.scancontexts(5, 400)
    m = _fill(SC, db, 0)
    src = db[123].reshape(20, 60)
    q = np.roll(src, 17, axis=1)
    it, s = m.search(q.reshape(-1), 1)
    assert it[0] == 123 and s[0] > 1 - 1e-12
    assert m.last_yaw_diff_deg == pytest.approx(17 * 6.0)  # rolling the candidate by 17 sectors gives the query


def test_empty_and_degenerate(SC):
    m = SC()
    assert m.search(np.zeros(1200), 1) == ([None], [None])     # reference :55-56
    assert m.search_best(np.zeros(1200)) == (None, None)         # reference :98-99
    db = synth.scancontexts(9, 3)
    m = _fill(SC, db, 40)
    # all-zero query: no column is engaged, every distance is 1, nothing beats the initial 1.0 ->
    # the reference returns items[0] with similarity 0.0 (:77-80)
    assert m.search(np.zeros(1200), 1) == ([40], [0.0])
    it, s = m.search(db[2], 1)                                   # fewer rows than num_candidates
    assert it[0] == 42 and abs(s[0] - 1.0) <= 1e-12
