"""CPU tests: the oracle (numpy + C restatement) against the golden vectors produced by running
the unmodified reference (tests/golden/make_golden.py)."""
import ctypes
import os

import numpy as np
import pytest

from conftest import ROOT, case_inputs
from oracle import nns_oracle as oracle

RTOL, ATOL = 1e-5, 1e-6  # parity rule, BASELINE.md §5


def _c_oracle():
    path = os.path.join(ROOT, "oracle", "_build", "liboracle.so")
    if not os.path.exists(path):
        import subprocess
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle")])
    lib = ctypes.CDLL(path)
    lib.oracle_search.restype = ctypes.c_int
    lib.oracle_search.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_int64, ctypes.c_int64,
                                  ctypes.c_void_p, ctypes.c_int64, ctypes.c_int, ctypes.c_void_p,
                                  ctypes.c_void_p]
    return lib


def c_search(lib, db, qs, k):
    db = np.ascontiguousarray(db, np.float32)
    qs = np.ascontiguousarray(qs, np.float32).reshape(-1, db.shape[1])
    idx = np.empty((len(qs), k), np.int64)
    sims = np.empty((len(qs), k), np.float32)
    rc = lib.oracle_search(db.ctypes.data, db.shape[0], db.shape[1], db.shape[1], qs.ctypes.data,
                           len(qs), k, idx.ctypes.data, sims.ctypes.data)
    assert rc == 0
    return idx, sims


def _same_or_tie(db, q, got_idx, ref_idx, ref_sims):
    """Index equality except where the reference's own similarities tie within tolerance."""
    s64 = oracle.cosine_similarities(db, q, np.float64)
    for j, (g, r) in enumerate(zip(got_idx, ref_idx)):
        if g != r:
            assert abs(s64[g] - s64[r]) <= RTOL * abs(s64[r]) + ATOL, (j, g, r)


def test_numpy_oracle_matches_reference_golden(golden_search):
    for name, c in golden_search.items():
        db, qs, k = case_inputs(c["spec"])
        for j, q in enumerate(qs):
            idx, sims = oracle.search(db, q, k)
            assert len(idx) == c["idx"].shape[1], name
            np.testing.assert_allclose(sims, c["sims"][j], rtol=RTOL, atol=ATOL, err_msg=name)
            _same_or_tie(db, q, idx, c["idx"][j], c["sims"][j])
            assert oracle.check_parity(db, q, k, c["idx"][j], c["sims"][j]) is None, name
        # search_best == first of search
        np.testing.assert_array_equal(c["best_item"] - 1000, c["idx"][:, 0])


def test_c_oracle_matches_reference_golden(golden_search):
    lib = _c_oracle()
    for name, c in golden_search.items():
        db, qs, k = case_inputs(c["spec"])
        idx, sims = c_search(lib, db, qs, k)
        kk = c["idx"].shape[1]
        np.testing.assert_allclose(sims[:, :kk], c["sims"], rtol=RTOL, atol=ATOL, err_msg=name)
        assert (idx[:, kk:] == -1).all()
        for j, q in enumerate(qs):
            _same_or_tie(db, q, idx[j, :kk], c["idx"][j], c["sims"][j])


def test_oracle_edge_cases(golden_edge):
    g = golden_edge
    db, qs = g["db"], g["queries"]
    lib = _c_oracle()
    cidx, csims = c_search(lib, db, qs, len(db))
    for j, q in enumerate(qs):
        idx, sims = oracle.search(db, q, len(db))
        np.testing.assert_allclose(sims, g["sims"][j], rtol=RTOL, atol=ATOL)
        np.testing.assert_allclose(csims[j], g["sims"][j], rtol=RTOL, atol=ATOL)
        _same_or_tie(db, q, idx, g["idx"][j], g["sims"][j])
        _same_or_tie(db, q, cidx[j], g["idx"][j], g["sims"][j])
    # zero row -> similarity exactly 0; zero query -> all 0
    assert np.all(oracle.cosine_similarities(db, qs[0])[3] == 0.0)
    assert np.all(oracle.cosine_similarities(db, qs[2]) == 0.0)
    # float64 torch query -> float64 sims in the reference
    assert g["f64_sims"].dtype == np.float64


def test_oracle_capacity_sequence(golden_search):
    # reference capacity: 1000 rows then doubling (nns_matching.py:36-38)
    for name, c in golden_search.items():
        n = int(c["spec"][0])
        cap = 1000
        while cap < n:
            cap *= 2
        assert int(c["capacity"]) == cap, name


def test_check_parity_flags_real_mismatch():
    rng = np.random.default_rng(0)
    db = rng.random((64, 8)).astype(np.float32)
    q = rng.random(8).astype(np.float32)
    idx, sims = oracle.search(db, q, 5)
    assert oracle.check_parity(db, q, 5, idx, sims) is None
    bad = idx.copy()
    bad[0] = oracle.search(db, q, 64)[0][-1]
    assert oracle.check_parity(db, q, 5, bad, sims) is not None
    assert oracle.check_parity(db, q, 5, idx, sims + 1e-3) is not None


def test_nan_sorts_first():
    db = np.ones((4, 3), np.float32)
    db[2, 0] = np.nan
    idx, _ = oracle.search(db, np.ones(3, np.float32), 2)
    assert idx[0] == 2


def test_caller_oracle_matches_reference_swarm_golden(golden_swarm):
    """oracle/lcsm_oracle.py (the callers' threshold / self-hit / min-gap logic)
    against the unmodified reference LoopClosureSparseMatching on 4 robots x 300 keyframes."""
    import synth
    from conftest import SWARM_PARAMS, check_swarm
    from oracle.lcsm_oracle import OracleLCSM
    descs = synth.swarm_inputs()
    lcsm = OracleLCSM(SWARM_PARAMS)
    intra, nsims, first, local_edges = [], [], [], []
    for i in range(len(descs[0])):
        local_edges.append(len(lcsm.add_local_global_descriptor(descs[0][i], i)))
        kf, sims = lcsm.match_local_loop_closures(descs[0][i], i)
        intra.append(-1 if kf is None else kf)
        nsims.append(len(sims))
        first.append(sims[0] if len(sims) else 0.0)
        for r in range(1, 4):
            lcsm.add_other_robot_global_descriptor(r, i, descs[r][i].tolist())
    check_swarm(golden_swarm, lcsm.candidate_edges, intra, nsims, first, local_edges)
    assert sum(1 for x in intra if x >= 0) >= 70 and 400 <= len(lcsm.candidate_edges) <= 460
