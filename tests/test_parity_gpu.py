"""GPU parity tests (run with ``-m gpu`` on a B200): the CUDA path, called through the Python
drop-in and therefore through the C-ABI, against (a) the golden vectors produced by the unmodified
reference and (b) the CPU oracle on the same seeded inputs.

Parity rule: indices equal to the reference's, except where the reference's own
neighbouring similarities are closer than the tolerance; similarities within 1e-5*|s| + 1e-6.
"""
import os
from collections import namedtuple

import numpy as np
import pytest
import torch

import synth
from conftest import case_inputs, config1_inputs, intra_mismatch_is_reference_tie
from oracle import nns_oracle as oracle

pytestmark = pytest.mark.gpu
RTOL, ATOL = 1e-5, 1e-6


@pytest.fixture(scope="module")
def M():
    from cslam_b200 import NearestNeighborsMatching
    return NearestNeighborsMatching


def build(M, db, items=None, bulk=False):
    m = M()
    if bulk:
        m.add_items(db, list(range(len(db))) if items is None else items)
    else:
        for i, row in enumerate(db):
            m.add_item(row, i if items is None else items[i])
    return m


def assert_parity(db, q, k, idx, sims):
    msg = oracle.check_parity(db, q, k, idx, sims, RTOL, ATOL)
    assert msg is None, msg


# ---- golden vectors (reference outputs) -----------------------------------------------------------

def test_golden_search_cases(M, golden_search):
    for name, c in golden_search.items():
        db, qs, k = case_inputs(c["spec"])
        m = build(M, db, items=[1000 + i for i in range(len(db))])
        assert m.n == len(db) and len(m.data) == int(c["capacity"]), name
        for j, q in enumerate(qs):
            items, sims = m.search(q, k)
            idx = np.asarray(items) - 1000
            assert sims.dtype == np.float32 and isinstance(items, list)
            np.testing.assert_allclose(sims, c["sims"][j], rtol=RTOL, atol=ATOL, err_msg=name)
            assert_parity(db, q, k, idx, sims)
            # against the reference's own indices, modulo reference ties
            s64 = oracle.cosine_similarities(db, q, np.float64)
            for a, b in zip(idx, c["idx"][j]):
                assert a == b or abs(s64[a] - s64[b]) <= RTOL * abs(s64[b]) + ATOL, (name, j)
            item, sim = m.search_best(q)
            assert isinstance(sim, np.float32)
            assert item == c["best_item"][j] or abs(sim - c["best_sim"][j]) <= ATOL
            assert abs(sim - c["best_sim"][j]) <= RTOL * abs(c["best_sim"][j]) + ATOL


def test_golden_edge_cases(M, golden_edge):
    g = golden_edge
    db, qs = g["db"], g["queries"]
    m = build(M, db)
    for j, q in enumerate(qs):
        items, sims = m.search(q, len(db))
        np.testing.assert_allclose(sims, g["sims"][j], rtol=RTOL, atol=ATOL)
        assert sorted(items) == list(range(len(db)))
        assert np.all(sims[:-1] >= sims[1:])
    # zero row -> 0 similarity; zero query -> all zeros, rows in index order (deterministic ties)
    items, sims = m.search(qs[2], len(db))
    assert np.all(sims == 0.0) and items == list(range(len(db)))
    # float64 torch query -> float64 similarities (torch type promotion in the reference)
    items, sims = m.search(torch.from_numpy(qs[0].astype(np.float64)), 4)
    assert sims.dtype == np.float64
    assert items == g["f64_items"].tolist()
    np.testing.assert_allclose(sims, g["f64_sims"], rtol=RTOL, atol=ATOL)


def test_golden_config1_loop_closure_sparse_matching(golden_config1):
    """BASELINE config 1: 2 robots x 1k keyframes x 128-d; the candidate-edge set must equal the one
    the reference produced (keys exactly; weights within tolerance)."""
    from cslam_b200.loop_closure_sparse_matching import CandidateEdges, LoopClosureSparseMatching
    GD = namedtuple("GlobalDescriptor", ["keyframe_id", "robot_id", "descriptor"])
    params = {"robot_id": 0, "max_nb_robots": 2, "frontend.sensor_type": "stereo",
              "frontend.similarity_threshold": 0.0, "frontend.nb_best_matches": 10,
              "frontend.intra_loop_min_inbetween_keyframes": 10}
    local, other = config1_inputs()
    lcsm = LoopClosureSparseMatching(params, candidate_selector=CandidateEdges(0, 2))
    intra = []
    for i in range(len(local)):
        lcsm.add_local_global_descriptor(local[i], i)
        kf, _ = lcsm.match_local_loop_closures(local[i], i)
        intra.append(-1 if kf is None else kf)
        lcsm.add_other_robot_global_descriptor(GD(i, 1, other[i].tolist()))
    edges = lcsm.candidate_selector.candidate_edges
    keys = np.asarray(sorted(edges.keys()), np.int64)
    g = golden_config1
    if not np.array_equal(keys, g["edge_keys"]):
        # any difference must be a tie in the reference's own similarities
        ours, ref = set(map(tuple, keys.tolist())), set(map(tuple, g["edge_keys"].tolist()))
        ref_w = {tuple(k): w for k, w in zip(g["edge_keys"].tolist(), g["edge_weights"])}
        for k in ours - ref:
            twins = [r for r in ref - ours if (r[0], r[1]) == (k[0], k[1]) or (r[2], r[3]) == (k[2], k[3])]
            assert twins and min(abs(ref_w[t] - edges[k].weight) for t in twins) <= 2e-6, k
    else:
        w = np.asarray([edges[tuple(k)].weight for k in keys.tolist()])
        np.testing.assert_allclose(w, g["edge_weights"], rtol=RTOL, atol=ATOL)
    for i, (a, b) in enumerate(zip(intra, g["intra"].tolist())):
        if a != b:   # allowed only where the reference's own similarities PROVE a tie (no blanket allowance)
            why = intra_mismatch_is_reference_tie(local, i, a, b, params["frontend.nb_best_matches"],
                                                  params["frontend.intra_loop_min_inbetween_keyframes"],
                                                  params["frontend.similarity_threshold"], RTOL, ATOL)
            assert why is None, (i, a, b, why)


def test_batched_messages_equal_looped():
    """a GlobalDescriptors message carries several descriptors; handing them to
    add_other_robot_global_descriptors (ONE batched search of the local database) must give the same
    matches and the same candidate edges as the reference's per-descriptor loop
    (global_descriptor_loop_closure_detection.py:388-404 -> loop_closure_sparse_matching.py:58-76)."""
    from cslam_b200.loop_closure_sparse_matching import CandidateEdges, LoopClosureSparseMatching
    GD = namedtuple("GlobalDescriptor", ["keyframe_id", "robot_id", "descriptor"])
    params = {"robot_id": 0, "max_nb_robots": 3, "frontend.sensor_type": "stereo",
              "frontend.similarity_threshold": 0.7, "frontend.nb_best_matches": 10,
              "frontend.intra_loop_min_inbetween_keyframes": 10}
    local = synth.unit_rows(51, 3000, 128, "randn")
    remote = synth.queries_for(52, local, 90, "randn", every=2, noise=0.03)
    msgs = [GD(1000 + i, 1 + (i % 2), remote[i].tolist()) for i in range(len(remote))]

    def run(batched):
        lcsm = LoopClosureSparseMatching(params, candidate_selector=CandidateEdges(0, 3))
        for i in range(len(local)):
            lcsm.add_local_global_descriptor(local[i], i)
        out = []
        for c in range(0, len(msgs), 10):   # the reference publishes descriptors in chunks (:192-221)
            chunk = msgs[c:c + 10]
            out += lcsm.add_other_robot_global_descriptors(chunk) if batched else \
                [lcsm.add_other_robot_global_descriptor(m) for m in chunk]
        return lcsm, out

    la, a = run(False)
    lb, b = run(True)
    assert sum(m is not None for m in a) == 45          # exactly the planted revisits pass the threshold
    for ma, mb in zip(a, b):
        assert (ma is None) == (mb is None)
        if ma is not None:
            assert (ma.robot0_keyframe_id, ma.robot1_id, ma.robot1_keyframe_id) == \
                (mb.robot0_keyframe_id, mb.robot1_id, mb.robot1_keyframe_id)
            assert abs(ma.weight - mb.weight) <= 1e-5 * abs(ma.weight) + 1e-6
    assert sorted(la.candidate_selector.candidate_edges) == sorted(lb.candidate_selector.candidate_edges)
    for rid in (1, 2):   # the other robots' databases received the same rows in the same order
        assert la.other_robots_nnsm[rid].n == lb.other_robots_nnsm[rid].n == 45
        assert la.other_robots_nnsm[rid].items == lb.other_robots_nnsm[rid].items


# ---- oracle on seeded inputs ------------------------------------------------------------------------

@pytest.mark.parametrize("n,d,k,dist", [
    (1, 8, 1, "rand"), (5, 10, 3, "rand"), (33, 100, 33, "rand"), (1000, 128, 10, "rand"),
    (1001, 128, 1, "randn"), (4097, 512, 10, "rand"), (3000, 256, 1, "randn"),
    (2500, 12, 7, "rand"), (300, 2048, 10, "rand"), (130, 4096, 5, "randn"),
    (64, 9000, 4, "rand"), (20000, 64, 128, "rand"), (777, 36, 100, "randn"),
])
def test_single_query_vs_oracle(M, n, d, k, dist):
    db = synth.unit_rows(100 + n, n, d, dist)
    qs = synth.queries_for(200 + n, db, 5, dist)
    m = build(M, db, bulk=True)
    for q in qs:
        items, sims = m.search(q, k)
        assert_parity(db, q, k, np.asarray(items), sims)
        assert np.all(sims[:-1] >= sims[1:])
    assert m._h.lib.cslam_nns_last_variant(m._h.h) == 1


@pytest.mark.parametrize("n,d", [(30000, 96), (150000, 128), (6000, 2100), (40, 64)])
def test_single_query_register_lists_33_to_128(M, n, d):
    """32 < k <= 128: four keys per lane (offer_big / warp_merge_desc_big) on the register path, the chunked
    path (rows longer than a stage), a grid of one CTA per SM, fewer rows than k, duplicated rows and a NaN row."""
    db = synth.unit_rows(700 + n, n, d)
    if n > 100:
        db[17] = db[5]            # exact duplicates: the lower row must come first
        db[n - 3] = db[5]
        db[99] = 0.0              # becomes the NaN row below
    qs = synth.queries_for(800 + n, db, 2)
    qs[1] = db[5] * 3.0
    m = build(M, db, bulk=True)
    for k in (33, 40, 64, 65, 96, 127, 128):
        for q in qs:
            items, sims = m.search(q, k)
            assert len(items) == min(k, n)
            assert_parity(db, q, k, np.asarray(items), sims)
    assert m._h.lib.cslam_nns_last_variant(m._h.h) == 1
    if n > 100:   # a NaN similarity sorts first (torch.argsort descending), the rest is unchanged
        want_items, want_sims = m.search(qs[0], 39)
        db[99, 0] = np.nan
        m2 = build(M, db, bulk=True)
        items, sims = m2.search(qs[0], 40)
        assert items[0] == 99 and np.isnan(sims[0])
        assert list(items[1:]) == list(want_items) and np.array_equal(sims[1:], want_sims)


def test_unnormalised_rows_and_queries(M):
    db = (synth.uniform(5, (600, 40)) * 7.0 - 2.0).astype(np.float32)
    q = (synth.uniform(6, (40,)) * 3.0).astype(np.float32)
    m = build(M, db, bulk=True)
    items, sims = m.search(q, 10)
    assert_parity(db, q, 10, np.asarray(items), sims)


def test_two_pass_large_k(M):
    n, d = 5000, 96
    db = synth.unit_rows(31, n, d)
    q = synth.unit_rows(32, 1, d)[0]
    m = build(M, db, bulk=True)
    for k in (129, 1000, n, n + 50):
        items, sims = m.search(q, k)
        assert len(items) == min(k, n)
        assert_parity(db, q, k, np.asarray(items), sims)
    assert m._h.lib.cslam_nns_last_variant(m._h.h) == 3


def test_batched_matches_single(M):
    n, d, B, k = 3000, 128, 37, 10
    db = synth.unit_rows(41, n, d)
    qs = synth.queries_for(42, db, B)
    m = build(M, db, bulk=True)
    items_b, sims_b = m.search_batch(qs, k)
    assert sims_b.shape == (B, k)
    for j in range(B):
        assert_parity(db, qs[j], k, np.asarray(items_b[j]), sims_b[j])
    # 2-D query through search() takes the batched route as well
    items2, sims2 = m.search(qs, k)
    assert items2 == items_b and np.array_equal(sims2, sims_b)


@pytest.mark.parametrize("n,d,B,k,dist", [
    (300, 32, 2, 3, "rand"), (1000, 128, 16, 10, "rand"), (5000, 100, 37, 10, "randn"),
    (20000, 256, 256, 10, "rand"), (2500, 4096, 64, 5, "rand"), (40000, 512, 200, 1, "randn"),
    (129, 36, 3, 32, "rand"), (9000, 10, 10, 10, "rand"), (3000, 130, 300, 8, "rand"),
])
def test_batched_tcgen05_vs_oracle(M, n, d, B, k, dist):
    """K2 (tcgen05 3xTF32 + fused top-k) against the oracle; B > 256 runs in passes of 256."""
    db = synth.unit_rows(300 + n, n, d, dist)
    qs = synth.queries_for(400 + n, db, B, dist)
    m = build(M, db, bulk=True)
    lib, h = m._h.lib, m._h.h
    lib.cslam_nns_set_variant(h, 2)
    idx, sims = m._search_block(torch.from_numpy(qs), k)
    assert lib.cslam_nns_last_variant(h) == 2
    for j in range(B):
        assert_parity(db, qs[j], k, idx[j], sims[j])
    # same answer as the single-query kernel wherever the oracle sees no tie
    lib.cslam_nns_set_variant(h, 1)
    idx1, sims1 = m._search_block(torch.from_numpy(qs), k)
    np.testing.assert_allclose(sims, sims1, rtol=2e-5, atol=2e-6)


def test_batched_after_incremental_adds(M):
    """The batched kernel must see rows appended after an earlier batched search (tensor map is
    rebuilt from the current row count on every call)."""
    d = 64
    db = synth.unit_rows(501, 700, d)
    m = M()
    m.add_items(db[:300], list(range(300)))
    items, _ = m.search_batch(db[290:310], 1)
    assert [it[0] for it in items[:10]] == list(range(290, 300))
    m.add_items(db[300:], list(range(300, 700)))
    items, sims = m.search_batch(db[290:310], 1)
    assert [it[0] for it in items] == list(range(290, 310))
    assert np.all(np.abs(sims[:, 0] - 1.0) < 1e-5)


def test_device_resident_inputs(M):
    n, d = 2048, 64
    db = synth.unit_rows(51, n, d)
    m = M()
    m.add_items(torch.from_numpy(db).cuda(), list(range(n)))
    q = synth.unit_rows(52, 1, d)[0]
    items, sims = m.search(torch.from_numpy(q).cuda(), 5)
    assert_parity(db, q, 5, np.asarray(items), sims)


# ---- reference semantics (tests/test_sparse_matching.py of the reference) -------------

def test_reference_test_similarity(M):
    """Same assertions as the reference's test_similarity (tests/test_sparse_matching.py:51-81)."""
    rng = np.random.default_rng(7)
    nnsm = M()
    for i in range(100):
        v = rng.random(100)
        nnsm.add_item(v / np.linalg.norm(v), i)
    for _ in range(100):
        query = rng.random(100)
        query = query / np.linalg.norm(query)
        ds = np.linalg.norm(query[np.newaxis, :] - nnsm.data.cpu().numpy()[:nnsm.n], axis=1)
        ns_dist = np.argsort(ds)[:100]
        ns_sim, sims = nnsm.search(query, 100)
        assert np.all(sims[:-1] >= sims[1:])
        for j in range(100):
            id_d, id_s = ns_dist[j], ns_sim[j]
            if id_d != id_s:
                if abs(sims[id_d] - sims[id_s]) < 1e-6 or abs(ds[id_d] - ds[id_s]) < 1e-6:
                    continue
            assert ns_dist[j] == ns_sim[j]
        best, _ = nnsm.search_best(query)
        assert ns_dist[0] == best


def test_storage_is_exact_fp32_and_data_view(M):
    rng = np.random.default_rng(8)
    m = M()
    v = rng.random(10)
    v = v / np.linalg.norm(v)
    m.add_item(v, "kf-a")                      # items are arbitrary python objects (:40)
    assert isinstance(m.data, torch.Tensor) and m.data.is_cuda
    assert tuple(m.data.shape) == (1000, 10) and m.dim == 10 and m.n == 1
    for i in range(10):                        # reference tests :35-36
        assert abs(v[i] - float(m.data[0, i])) < 1e-7
    assert np.array_equal(m.data[0].cpu().numpy(), v.astype(np.float32))
    assert float(m.data[1:].abs().sum()) == 0.0   # rows past n are zeros, as torch.zeros gives
    assert m.search_best(v)[0] == "kf-a"


def test_capacity_doubles_without_moving_rows(M):
    d = 20
    db = synth.unit_rows(61, 4100, d)
    m = M()
    caps = set()
    for i, row in enumerate(db):
        m.add_item(row, i)
        caps.add(len(m.data))
    assert caps == {1000, 2000, 4000, 8000}     # reference sequence (nns_matching.py:36-38)
    assert np.array_equal(m.data.cpu().numpy()[:m.n], db)
    items, sims = m.search(db[4099], 1)
    assert items == [4099] and abs(sims[0] - 1.0) < 1e-6


def test_empty_and_quirks(M):
    m = M()
    assert m.search(np.ones(4), 3) == ([], []) and m.search_best(np.ones(4)) == (None, None)
    m2 = M(dim=6)                               # constructed with dim, still empty
    assert tuple(m2.data.shape) == (1000, 6)
    items, sims = m2.search(np.ones(6), 3)
    assert items == [] and sims.shape == (0,) and sims.dtype == np.float32
    with pytest.raises(IndexError):             # reference quirk
        m2.search_best(np.ones(6))
    m2.add_item(np.arange(6, dtype=np.float64), 5)
    assert m2.search(np.ones(6), 0)[0] == []
    items, sims = m2.search(np.ones(6), 7)      # k > n -> n results
    assert items == [5] and sims.shape == (1,)
    with pytest.raises(RuntimeError):           # wrong length
        m2.add_item(np.ones(5), 1)
    with pytest.raises(RuntimeError):
        m2.search(np.ones(7), 1)


def test_nan_row_sorts_first(M):
    db = synth.unit_rows(71, 50, 8)
    db[17, 3] = np.nan
    m = build(M, db, bulk=True)
    items, sims = m.search(synth.unit_rows(72, 1, 8)[0], 3)
    assert items[0] == 17 and np.isnan(sims[0])


def test_interleaved_add_and_search(M):
    """Config-4 style: every search must see every earlier add (stream order)."""
    d = 256
    db = synth.unit_rows(81, 1500, d)
    m = M()
    m.add_items(db[:1000], list(range(1000)))
    for i in range(1000, 1500):
        m.add_item(db[i], i)
        item, sim = m.search_best(db[i])
        assert item == i and abs(sim - 1.0) < 1e-5


def test_no_vmm_fallback_path(M, monkeypatch):
    monkeypatch.setenv("CSLAM_NNS_NO_VMM", "1")
    db = synth.unit_rows(91, 2500, 32)
    m = build(M, db)
    assert np.array_equal(m.data.cpu().numpy()[:m.n], db)
    items, sims = m.search(db[2400], 2)
    assert items[0] == 2400


# ---- wire-format ingest and persistence ------------------------------------------------

def test_message_ingest_and_save_load_round_trip(M, tmp_path):
    GD = namedtuple("GlobalDescriptor", ["keyframe_id", "robot_id", "descriptor"])
    db = synth.unit_rows(91, 1500, 96)
    qs = synth.queries_for(92, db, 5)
    one_by_one = M()
    for i in range(len(db)):
        one_by_one.add_item(db[i], 10 + i)
    by_msg = M()
    for lo in range(0, len(db), 10):   # <= 10 descriptors per message, like the reference's publisher
        by_msg.add_descriptor_msgs([GD(10 + i, 1, db[i].tolist()) for i in range(lo, min(lo + 10, len(db)))])
    assert by_msg.n == one_by_one.n and by_msg.items == one_by_one.items
    assert torch.equal(by_msg.data[:by_msg.n].cpu(), one_by_one.data[:one_by_one.n].cpu())
    path = str(tmp_path / "db.npz")
    by_msg.save(path)
    back = M.load(path)
    assert back.n == by_msg.n and back.items == by_msg.items and back.dim == by_msg.dim
    assert torch.equal(back.data[:back.n].cpu(), by_msg.data[:by_msg.n].cpu())
    for q in qs:
        a, sa = one_by_one.search(q, 10)
        b, sb = back.search(q, 10)
        assert a == b and np.array_equal(sa, sb)


def test_pinned_message_staging_is_not_reused_early(M):
    """cslam_nns_add must not return while the copy engine still reads a PINNED source (include/cslam_nns.h):
    add_descriptor_msgs packs every message into a pinned staging block that torch's pinned cache hands straight
    to the next message.  Two different 10-descriptor messages alternate 1000 times; every stored row must be
    the row of ITS message."""
    GD = namedtuple("GlobalDescriptor", ["keyframe_id", "robot_id", "descriptor"])
    d = 512
    a = synth.unit_rows(93, 10, d)
    b = synth.unit_rows(94, 10, d, "randn")
    msg_a = [GD(i, 1, a[i].tolist()) for i in range(10)]
    msg_b = [GD(100 + i, 1, b[i].tolist()) for i in range(10)]
    m = M()
    for r in range(1000):
        m.add_descriptor_msgs(msg_a if r % 2 == 0 else msg_b)
    got = m.data[:m.n].cpu().numpy().reshape(1000, 10, d)
    assert np.array_equal(got[0::2], np.broadcast_to(a, (500, 10, d)))
    assert np.array_equal(got[1::2], np.broadcast_to(b, (500, 10, d)))
    # the same through the raw C-ABI with a caller-owned pinned buffer that is overwritten right after the call
    m2 = M(dim=d)
    h = m2._open()
    pin = torch.empty((10, d), dtype=torch.float32).pin_memory()
    for r in range(300):
        pin.copy_(torch.from_numpy(a if r % 2 == 0 else b))
        rc = h.lib.cslam_nns_add(h.h, pin.data_ptr(), 10, d, 0, m2._stream(), None)
        assert rc == 0
        pin.zero_()                           # legal: the source may be reused as soon as the call returned
        m2.n += 10
    got = m2.data[:m2.n].cpu().numpy().reshape(300, 10, d)
    assert np.array_equal(got[0::2], np.broadcast_to(a, (150, 10, d)))
    assert np.array_equal(got[1::2], np.broadcast_to(b, (150, 10, d)))


def test_save_load_refuses_pickle_by_default_and_data_setter(M, tmp_path):
    db = synth.unit_rows(95, 50, 24)
    m = M()
    for i, row in enumerate(db):
        m.add_item(row, "kf-%d" % i)              # arbitrary python items
    path = str(tmp_path / "objs.npz")
    with pytest.raises(TypeError):
        m.save(path)                              # non-integer items need an explicit opt-in
    m.save(path, allow_pickle=True)
    with pytest.raises(ValueError):
        M.load(path)
    back = M.load(path, allow_pickle=True)
    assert back.items == m.items and torch.equal(back.data[:50].cpu(), m.data[:50].cpu())
    # `data` is assignable like the reference's plain attribute: the rows are copied into the arena
    m2 = M()
    m2.add_items(db, list(range(50)))
    m2.data = torch.from_numpy(db[::-1].copy())   # n and items stay as they are, the stored rows change
    assert tuple(m2.data.shape) == (1000, 24) and m2.n == 50
    items, sims = m2.search(db[7], 1)
    assert items == [42] and abs(sims[0] - 1.0) < 1e-6


# ---- both batched tilings on every shape --------------------------------------------------------------

@pytest.mark.parametrize("route", ["qm", "rows", "qm_noscreen", "qm_screen_all", "qm_seeded"])
def test_both_batched_kernels_on_all_shapes(route):
    """The router (k2_qm.cu: q_route) normally sends small batches (<= 96 queries, fewer at small
    dimensions) to the row-major kernel and the rest to the query-major CTA-pair kernel; here each kernel is forced (the switch is read once per process,
    hence the subprocess) over shapes that cover padding on every axis: n < 256 and not a multiple of 256,
    dim not a multiple of 32, B = 2 .. 256, k = 1 .. 32."""
    import subprocess
    import sys
    from conftest import ROOT
    shapes = ["300,32,4,3", "1000,128,16,10", "5000,100,37,10", "20000,256,256,10", "3000,4096,64,5",
              "70000,512,200,1", "129,36,2,32", "257,40,97,7", "4096,64,255,32", "777,200,130,10"]
    if route == "rows":   # the row-major kernel holds <= 192 queries (tensor-memory columns); larger forced
        # batches fall through to the query-major kernel, which the other parameter covers
        shapes = [s for s in shapes if int(s.split(",")[2]) <= 192] + ["9000,96,192,4", "2500,72,130,32"]
    env = dict(os.environ, CSLAM_NNS_K2_FORCE=route.split("_")[0])
    if route == "qm_noscreen":      # every batch through the 3xTF32 pass, also the big databases
        env["CSLAM_NNS_SCREEN"] = "0"
    elif route == "qm_screen_all":  # the screening route (one product + exact re-scoring) on EVERY shape,
        env["CSLAM_NNS_SCREEN_MIN_ROWS"] = "0"   # also n < 256, k = 32, dim not a multiple of 4 or 32
    elif route == "qm_seeded":      # ... and with the bound seeded from a pass over every 4th row on every shape
        env.update(CSLAM_NNS_SCREEN_MIN_ROWS="0", CSLAM_NNS_SEED_MIN_ROWS="0", CSLAM_NNS_SEED_STRIDE="4",
                   CSLAM_NNS_SEED_MAX_DIM="100000")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "k2_check.py")] + shapes, env=env,
                         capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and "K2 CHECK OK" in out.stdout, out.stdout[-2000:] + out.stderr[-2000:]


def test_batched_nan_zero_and_duplicate_rows(M):
    """Rows that are NaN (sort first, like torch.argsort descending), all-zero (similarity 0) and exact
    duplicates (the lower row wins) through the query-major kernel (B = 128 > 96)."""
    n, d, B, k = 3000, 64, 128, 6
    db = synth.unit_rows(71, n, d)
    db[5] = np.nan
    db[7] = 0.0
    db[100] = db[50]
    db[2000] = db[50]
    qs = synth.queries_for(72, db, B)
    qs[3] = db[50]
    m = M()
    m.add_items(db, list(range(n)))
    items, sims = m.search_batch(qs, k)
    for j in range(B):
        assert items[j][0] == 5 and np.isnan(sims[j][0])            # the NaN row first
        ref_items, ref_sims = m.search(qs[j], k)                     # single-query kernel
        assert items[j] == ref_items
        np.testing.assert_allclose(sims[j][1:], ref_sims[1:], rtol=RTOL, atol=ATOL)
    assert items[3][1:4] == [50, 100, 2000]                          # duplicates: lower row first


@pytest.mark.gpu
def test_search_batches_pipeline_equals_search_batch():
    """The pipelined generator (uploads, searches and result reads of consecutive blocks overlapping) returns what
    one synchronous search_batch per block returns — bitwise — for ragged block sizes, a reused matcher and an
    early stop of the consumer; the sharded class (one rank) goes through the same pipeline."""
    import numpy as np
    import torch
    import synth
    from cslam_b200 import NearestNeighborsMatching
    from cslam_b200.sharded import ShardedNearestNeighborsMatching
    n, d, k = 30_000, 256, 10
    db = synth.unit_rows(4242, n, d)
    m = NearestNeighborsMatching()
    m.add_items(db, [1000 + i for i in range(n)])
    sizes = [256, 256, 64, 256, 1, 130, 256, 256]
    blocks = [synth.queries_for(500 + j, db, b) for j, b in enumerate(sizes)]
    want = [m.search_batch(b, k) for b in blocks]
    for depth in (1, 2, 3):
        got = list(m.search_batches(iter(blocks), k, depth=depth))
        assert len(got) == len(want)
        for (gi, gs), (wi, ws) in zip(got, want):
            assert gi == wi and np.array_equal(gs, ws)
    # pinned torch blocks, and a consumer that stops after two blocks; the matcher stays usable
    pinned = [torch.from_numpy(b).pin_memory() for b in blocks]
    gen = m.search_batches(iter(pinned), k)
    first, second = next(gen), next(gen)
    gen.close()
    assert first[0] == want[0][0] and second[0] == want[1][0]
    again = list(m.search_batches(iter(pinned[:3]), k))
    assert [a[0] for a in again] == [w[0] for w in want[:3]]
    # the sharded class with one rank
    sm = ShardedNearestNeighborsMatching(dim=d)
    sm.add_local_shard(torch.from_numpy(db).cuda())
    sm.finalize_bulk()
    got = list(sm.search_batches(iter(blocks), k))
    for (gi, gs), (wi, ws) in zip(got, want):
        assert np.array_equal(np.asarray(gi) + 1000, np.asarray(wi)) and np.array_equal(gs, ws)
    # a prepared device call returns the bits of the ordinary one, also after the query buffer's contents change
    qd = torch.from_numpy(blocks[0]).cuda()
    oi = torch.empty((256, k), dtype=torch.int64, device="cuda")
    os_ = torch.empty((256, k), dtype=torch.float32, device="cuda")
    plan = sm.plan_search_device(qd, k, oi, os_)
    for j in (0, 1, 3):
        qd.copy_(torch.from_numpy(blocks[j]))
        plan()
        ri, rs = sm.search_batch_device(qd, k)
        torch.cuda.synchronize()
        assert torch.equal(ri, oi) and torch.equal(rs, os_)
