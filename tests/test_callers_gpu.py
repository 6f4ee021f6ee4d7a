"""GPU parity tests of the callers' side of the path: the multi-database search
(`cslam_nns_search_multi`) and the device-side candidate filter (`cslam_nns_search_filtered`), through the
Python drop-in and therefore through the C-ABI, against the reference's golden (4-robot swarm) and the CPU
oracle of the callers (oracle/lcsm_oracle.py)."""
from collections import namedtuple

import numpy as np
import pytest
import torch

import synth
from conftest import SWARM_PARAMS, check_swarm
from oracle import lcsm_oracle, nns_oracle as oracle

pytestmark = pytest.mark.gpu
GD = namedtuple("GlobalDescriptor", ["keyframe_id", "robot_id", "descriptor"])


def _run_swarm(lcsm, descs, batched_msgs=False):
    intra, nsims, first, local_edges = [], [], [], []
    for i in range(len(descs[0])):
        local_edges.append(len(lcsm.add_local_global_descriptor(descs[0][i], i)))
        kf, sims = lcsm.match_local_loop_closures(descs[0][i], i)
        intra.append(-1 if kf is None else kf)
        nsims.append(len(sims))
        first.append(sims[0] if len(sims) else 0.0)
        msgs = [GD(i, r, descs[r][i].tolist()) for r in range(1, 4)]
        if batched_msgs:
            lcsm.add_other_robot_global_descriptors(msgs)
        else:
            for m in msgs:
                lcsm.add_other_robot_global_descriptor(m)
    edges = {k: float(e.weight) for k, e in lcsm.candidate_selector.candidate_edges.items()}
    return edges, intra, nsims, first, local_edges


@pytest.mark.parametrize("batched_msgs", [False, True])
def test_swarm_golden_through_fused_callers(golden_swarm, batched_msgs):
    """4 robots x 300 keyframes, threshold 0.85, min-gap 20: candidate edges, intra matches and returned
    similarity arrays equal the unmodified reference's; the multi-database search and the device filter
    are the code that ran (launch counter + last variant)."""
    from cslam_b200.loop_closure_sparse_matching import CandidateEdges, LoopClosureSparseMatching
    lcsm = LoopClosureSparseMatching(SWARM_PARAMS, candidate_selector=CandidateEdges(0, 4))
    res = _run_swarm(lcsm, synth.swarm_inputs(), batched_msgs)
    check_swarm(golden_swarm, *res)
    assert lcsm.local_nnsm.ids_on_device          # the fused filter path was eligible throughout
    assert lcsm.local_nnsm._h.lib.cslam_nns_launch_count() > 0


def test_search_best_multi_equals_per_database_search_best():
    """One query against 7 databases of different sizes (one empty, one never allocated): same items and
    bit-identical similarities as seven search_best calls (reference loop :47-55)."""
    from cslam_b200 import NearestNeighborsMatching as M
    d = 256
    sizes = [0, 1, 37, 1000, 4099, 0, 20000]
    ms = []
    for j, n in enumerate(sizes):
        m = M()
        if n:
            m.add_items(synth.unit_rows(300 + j, n, d, "randn"), [10 * j + i for i in range(n)])
        ms.append(m)
    ms[5] = M(dim=d)   # allocated but empty: the reference raises IndexError in search_best
    qs = synth.unit_rows(399, 6, d, "randn")
    for q in qs:
        live = [m for m in ms if m.n > 0]
        multi = M.search_best_multi(live + [ms[0]], torch.from_numpy(q))
        assert multi[-1] == (None, None)
        for m, (item, sim) in zip(live, multi):
            it1, s1 = m.search_best(q)
            assert item == it1 and isinstance(sim, np.float32)
            assert sim == s1                                  # same kernel, same bits
            db = m.data[:m.n].cpu().numpy()
            assert oracle.check_parity(db, q, 1, np.asarray([item - 10 * ms.index(m)]), np.asarray([sim])) is None
    with pytest.raises(IndexError):
        M.search_best_multi([ms[5]], qs[0])
    a, b = M(), M()
    a.add_item(np.ones(8, np.float32), 0)
    b.add_item(np.ones(9, np.float32), 0)
    with pytest.raises(Exception):
        M.search_best_multi([a, b], np.ones(8, np.float32))   # databases of different dim


@pytest.mark.parametrize("n,d,k", [(1, 16, 10), (9, 16, 10), (500, 128, 10), (5000, 64, 10), (3000, 512, 32)])
def test_search_candidates_intra_vs_oracle(n, d, k):
    """search + device filter == search followed by the reference's host filter (:83-97), for every
    combination of self hit, min-gap window and threshold, with keyframe ids that are not the row index."""
    from cslam_b200 import NearestNeighborsMatching as M
    db = synth.unit_rows(500 + n, n, d, "rand")
    ids = [3 * i + 7 for i in range(n)]
    m = M()
    if n // 2:
        m.add_items(db[: n // 2], ids[: n // 2])
    for i in range(n // 2, n):              # ids arrive through both add paths
        m.add_item(db[i], ids[i])
    qs = synth.queries_for(600 + n, db, 12, "rand", every=2, noise=0.02)
    for j, q in enumerate(qs):
        items_ref, sims_ref = m.search(q, k)
        for min_gap in (0, 5, 40):
            for thr in (-1.0, float(np.median(sims_ref)), float(sims_ref[0]), 2.0):
                for qid in (items_ref[0], items_ref[min(1, len(items_ref) - 1)] + 1, 10 ** 12):
                    items, sims, pos = m.search_candidates(q, k, query_id=qid, min_gap=min_gap, threshold=thr)
                    assert items == items_ref and np.array_equal(sims, sims_ref)
                    want, want_sims = lcsm_oracle.filter_intra(items_ref, sims_ref, qid, min_gap, thr)
                    got = items[pos] if pos >= 0 else None
                    assert got == want, (j, min_gap, thr, qid, pos)
    # batch form: B queries with their own ids
    qid = np.asarray([ids[(7 * j + 1) % n] for j in range(len(qs))], np.int64)
    items, sims, pos = m.search_candidates(qs, k, query_id=qid, min_gap=5, threshold=0.5)
    assert sims.shape == (len(qs), min(k, n)) and pos.dtype == np.int32
    for j, q in enumerate(qs):
        assert oracle.check_parity(db, q, k, (np.asarray(items[j]) - 7) // 3, sims[j]) is None
        want, _ = lcsm_oracle.filter_intra(items[j], sims[j], int(qid[j]), 5, 0.5)
        assert (items[j][pos[j]] if pos[j] >= 0 else None) == want


def test_search_candidates_best_mode_and_errors():
    from cslam_b200 import NearestNeighborsMatching as M
    from cslam_b200._native import FILTER_BEST
    db = synth.unit_rows(701, 2000, 128, "randn")
    m = M()
    m.add_items(db, list(range(2000)))
    qs = synth.queries_for(702, db, 40, "randn", every=2, noise=0.03)
    items, sims, pos = m.search_candidates(qs, 1, mode=FILTER_BEST, threshold=0.7, batched=True)
    ref_items, ref_sims = m.search_batch(qs, 1)
    assert items == ref_items and np.array_equal(sims, ref_sims)
    for j in range(len(qs)):
        assert (pos[j] == 0) == lcsm_oracle.filter_best(ref_items[j][0], ref_sims[j][0], 0.7)
    assert 0 < int((pos == 0).sum()) < len(qs)
    # a NaN row sorts first (torch.argsort) and does not pass `>=`, but is not skipped by `<` (:93)
    m2 = M()
    rows = synth.unit_rows(703, 50, 16, "rand")
    rows[11, 3] = np.nan
    m2.add_items(rows, list(range(100, 150)))
    it, s, p = m2.search_candidates(rows[0], 1, mode=FILTER_BEST, threshold=0.0)
    assert it == [111] and np.isnan(s[0]) and p == -1
    it, s, p = m2.search_candidates(rows[0], 5, query_id=0, min_gap=0, threshold=0.0)
    assert it[0] == 111 and p == 0
    # non-integer items: the id table does not exist; asking for an id filter is an error, not a guess
    m3 = M()
    m3.add_item(rows[0], "kf-a")
    m3.add_item(rows[1], 1)
    assert not m3.ids_on_device
    with pytest.raises(TypeError):
        m3.search_candidates(rows[0], 2, query_id=1)
    it, s, p = m3.search_candidates(rows[0], 2, mode=FILTER_BEST, threshold=0.5)
    assert it[0] == "kf-a" and p == 0


def test_row_ids_survive_capacity_growth():
    """ids are uploaded lazily (tail only) into a second VMM arena that grows with the database."""
    from cslam_b200 import NearestNeighborsMatching as M
    d = 32
    db = synth.unit_rows(801, 2600, d, "rand")
    m = M()
    for hi in (10, 900, 1100, 2600):                      # crosses 1000 and 2000 (two doublings)
        lo = m.n
        m.add_items(db[lo:hi], [5 * i for i in range(lo, hi)])
        q = db[hi - 1]
        items, sims, pos = m.search_candidates(q, 10, query_id=5 * (hi - 1), min_gap=3, threshold=-1.0)
        assert items[0] == 5 * (hi - 1)
        want, _ = lcsm_oracle.filter_intra(items, sims, 5 * (hi - 1), 3, -1.0)
        assert (items[pos] if pos >= 0 else None) == want
        assert pos >= 1                                    # the self hit at position 0 was dropped
