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


@pytest.fixture(scope="session")
def golden_swarm():
    return dict(np.load(os.path.join(GOLDEN, "swarm_lcsm.npz")))


SWARM_PARAMS = {"robot_id": 0, "max_nb_robots": 4, "frontend.sensor_type": "stereo",
                "frontend.similarity_threshold": 0.85, "frontend.nb_best_matches": 10,
                "frontend.intra_loop_min_inbetween_keyframes": 20}


def intra_mismatch_is_reference_tie(local, i, ours, ref, k, min_gap, thr, rtol=1e-5, atol=1e-6):
    """Proof that a differing intra-robot match at keyframe ``i`` is a tie in the reference's own arithmetic.
    At that moment the local database is ``local[:i+1]`` (items = row numbers) and the query is ``local[i]``
    (loop_closure_sparse_matching.py:78-97).  With the fp64 similarities of that search, the two outcomes may
    differ only because (a) the two returned rows' similarities are within tolerance of each other, or
    (b) the row one side returned sits within tolerance of the threshold, of the top-k cut, or the self hit
    ties for first place (then `kfs[0] == kf_id` is decided by the tie).  Returns None or the reason it is NOT."""
    from oracle import nns_oracle
    s64 = nns_oracle.cosine_similarities(local[:i + 1], local[i], np.float64)
    order = np.argsort(-s64, kind="stable")

    def tol(v):
        return rtol * abs(v) + atol
    if ours >= 0 and ref >= 0:
        if abs(s64[ours] - s64[ref]) <= tol(s64[ref]):
            return None
        return "rows %d / %d have fp64 similarities %.9g / %.9g: not a tie" % (ours, ref, s64[ours], s64[ref])
    x = ours if ours >= 0 else ref
    if abs(s64[x] - thr) <= tol(thr):
        return None
    if len(order) > k and abs(s64[x] - s64[order[k - 1]]) <= tol(s64[x]) and \
            abs(s64[order[k - 1]] - s64[order[k]]) <= tol(s64[x]):
        return None
    if len(order) > 1 and abs(s64[order[0]] - s64[order[1]]) <= tol(s64[order[0]]):
        return None
    return "row %d (fp64 similarity %.9g) returned by one side only, not near the threshold %.3g nor a cut" % (
        x, s64[x], thr)


def check_swarm(g, edges, intra, nsims, first_sim, local_edges, rtol=1e-5, atol=1e-6):
    """Compare one swarm run (edge key -> weight, per-keyframe intra match / len(sims) / sims[0] / number of
    edges created by the local keyframe) with the reference's golden.  Every difference must be PROVEN to be a
    case where the reference's own decision sits within tolerance of the threshold or of a tie: there is no
    allowance for unexplained mismatches."""
    import synth
    thr = SWARM_PARAMS["frontend.similarity_threshold"]
    ref = {tuple(k): w for k, w in zip(g["edge_keys"].tolist(), g["edge_weights"])}
    proven = 0
    for k in set(ref) ^ set(edges):
        w = ref.get(k, edges.get(k))
        other = edges if k in ref else ref
        twins = [t for t in other if t not in (ref if k in ref else edges)
                 and ((t[0], t[1]) == (k[0], k[1]) or (t[2], t[3]) == (k[2], k[3]))
                 and abs(other[t] - w) <= 2e-6]     # the reference's own top-2 tie within tolerance
        assert abs(w - thr) <= 2e-6 or twins, ("edge only on one side and not a tie", k, w)
        proven += 1
    for k in set(ref) & set(edges):
        assert abs(ref[k] - edges[k]) <= rtol * abs(ref[k]) + atol, (k, ref[k], edges[k])
    local = synth.swarm_inputs()[0]
    for i, (a, b) in enumerate(zip(intra, g["intra"].tolist())):
        if a != b:
            why = intra_mismatch_is_reference_tie(
                local, i, a, b, SWARM_PARAMS["frontend.nb_best_matches"],
                SWARM_PARAMS["frontend.intra_loop_min_inbetween_keyframes"], thr, rtol, atol)
            assert why is None, ("intra match of keyframe %d differs from the reference's" % i, a, b, why)
    np.testing.assert_array_equal(np.asarray(nsims), g["nsims"])
    np.testing.assert_allclose(np.asarray(first_sim, np.float32), g["first_sim"], rtol=rtol, atol=atol)
    # the per-keyframe edge counts can only differ through the edges proven above to be reference ties
    assert int(np.abs(np.asarray(local_edges) - g["local_edges"]).sum()) <= proven
