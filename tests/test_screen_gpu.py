"""The screening route of the batched search (csrc/k2_qm.cu SCREEN + csrc/k6_rescore.cu): one tf32 product
selects candidates, an exact fp32 re-scoring with a completeness proof returns the reference's top-k, and the
queries the proof does not cover are recomputed by the masked 3xTF32 pass.  Results must obey the same parity
rule as every other path (oracle = CPU restatement of cslam/nns_matching.py:57-63)."""
import ctypes

import numpy as np
import pytest
import torch

import synth
from oracle import nns_oracle as oracle

pytestmark = pytest.mark.gpu


def _stats(m):
    a, b = ctypes.c_int64(), ctypes.c_int64()
    m._h.lib.cslam_nns_screen_stats(m._h.h, ctypes.byref(a), ctypes.byref(b))
    return a.value, b.value


def _matcher(db, min_rows=None):
    from cslam_b200 import NearestNeighborsMatching
    m = NearestNeighborsMatching()
    m.add_items(db, list(range(len(db))))
    if min_rows is not None:
        m._h.lib.cslam_nns_set_screen(m._h.h, 1, min_rows)
    return m


@pytest.mark.parametrize("n,d,B,k,dist", [
    (20000, 256, 256, 10, "rand"), (40000, 512, 200, 1, "randn"), (9000, 100, 130, 10, "rand"),
    (3000, 130, 128, 8, "rand"), (600, 36, 97, 32, "randn"), (30000, 64, 256, 16, "rand"),
    (12000, 1024, 256, 10, "rand"), (5000, 10, 100, 5, "rand"),
])
def test_screened_search_vs_oracle(n, d, B, k, dist):
    db = synth.unit_rows(900 + n, n, d, dist)
    qs = synth.queries_for(901 + n, db, B, dist)
    m = _matcher(db, min_rows=0)
    idx, sims = m._search_block(torch.from_numpy(qs), k)
    assert m._h.lib.cslam_nns_last_variant(m._h.h) == 2
    screened, fallback = _stats(m)
    assert screened == B, "the screening route did not run"
    for j in range(B):
        msg = oracle.check_parity(db, qs[j], k, idx[j], sims[j])
        assert msg is None, (j, msg)
    assert np.all(sims[:, :-1] >= sims[:, 1:])
    if n >= 9000:
        assert fallback <= B // 8, (screened, fallback)     # generic data: the proof covers (nearly) every query
    # the re-scored similarities are plain fp32 dot products, far inside the tolerance (the 3xTF32 pass that
    # answers the queries the proof does not cover is an order of magnitude less precise, still inside it)
    s64 = np.stack([oracle.cosine_similarities(db[idx[j]], qs[j], np.float64) for j in range(0, B, 7)])
    assert np.abs(sims[::7] - s64).max() < (1e-6 if fallback == 0 else 8e-6)
    # same rows as the 3xTF32 pass wherever the oracle sees no tie
    m._h.lib.cslam_nns_set_screen(m._h.h, 0, -1)
    idx3, sims3 = m._search_block(torch.from_numpy(qs), k)
    assert _stats(m)[0] == screened                              # ... which did not screen
    np.testing.assert_allclose(sims, sims3, rtol=1e-5, atol=2e-6)
    differ = idx != idx3
    for j, i in zip(*np.nonzero(differ)):
        a = oracle.cosine_similarities(db[[idx[j, i], idx3[j, i]]], qs[j], np.float64)
        assert abs(a[0] - a[1]) <= 1e-5 * abs(a[0]) + 1e-6


def test_neighbours_in_consecutive_rows_are_proven_without_fallback():
    """SLAM data: the keyframes of a revisited place are consecutive rows, i.e. they all land in ONE of the
    screening pass's lists.  The lists hold 2k + 4 entries so that this does not push the list's published
    bound up to the k-th best (which would send the query to the 3xTF32 pass every time)."""
    n, d, B, k = 30000, 128, 128, 10
    db = synth.unit_rows(83, n, d)
    qs = synth.unit_rows(84, B, d)
    eps = synth.normal(85, (B, 9, d)).astype(np.float32)
    for j in range(B):                       # query j revisits a place seen at rows 200 j + 100 .. + 108
        for t in range(9):
            v = qs[j] + (0.010 + 0.003 * t) * eps[j, t]
            db[200 * j + 100 + t] = v / np.linalg.norm(v)
    m = _matcher(db)
    idx, sims = m._search_block(torch.from_numpy(qs), k)
    screened, fallback = _stats(m)
    assert screened == B and fallback == 0, (screened, fallback)
    for j in range(B):
        assert sorted(idx[j][:9].tolist()) == list(range(200 * j + 100, 200 * j + 109))
        msg = oracle.check_parity(db, qs[j], k, idx[j], sims[j])
        assert msg is None, (j, msg)


def test_screened_search_soak():
    """Many screened searches back to back with changing batch sizes, k and databases: every call must agree
    with the single-query kernel (same rows; similarities to fp32 rounding)."""
    from cslam_b200 import NearestNeighborsMatching
    for rep, (n, d) in enumerate([(9000, 100), (20000, 256), (33000, 48)]):
        db = synth.unit_rows(600 + rep, n, d)
        m = NearestNeighborsMatching()
        m.add_items(db, list(range(n)))
        m._h.lib.cslam_nns_set_screen(m._h.h, 1, 0)
        for it, (B, k) in enumerate([(130, 10), (256, 1), (97, 5), (200, 16), (256, 10), (128, 32), (255, 3)] * 3):
            qs = synth.queries_for(700 + 10 * rep + it, db, B)
            idx, sims = m._search_block(torch.from_numpy(qs), k)
            m._h.lib.cslam_nns_set_variant(m._h.h, 1)
            sub = list(range(0, B, 16))
            i1, s1 = m._search_block(torch.from_numpy(qs[sub]), k)
            m._h.lib.cslam_nns_set_variant(m._h.h, 0)
            np.testing.assert_allclose(sims[sub], s1, rtol=1e-5, atol=2e-6)
            assert (idx[sub] == i1).mean() > 0.98
        assert _stats(m)[0] > 0


def test_near_duplicate_database_takes_the_fallback_and_stays_exact():
    """Hundreds of rows within 1e-4 of each other: the screened similarities cannot separate them, the band
    overflows or the bound is too close, and the masked 3xTF32 pass must answer those queries — exactly."""
    n, d, B, k = 20000, 128, 160, 10
    db = synth.unit_rows(77, n, d)
    noise = synth.normal(78, (600, d)).astype(np.float32)
    db[5000:5600] = db[17] + 2e-5 * noise
    db[5000:5600] /= np.linalg.norm(db[5000:5600], axis=1, keepdims=True)
    qs = synth.queries_for(79, db, B)
    qs[0:40] = db[17] + 1e-5 * synth.normal(80, (40, d)).astype(np.float32)   # 40 queries into the cluster
    m = _matcher(db)
    idx, sims = m._search_block(torch.from_numpy(qs), k)
    screened, fallback = _stats(m)
    assert screened == B and 40 <= fallback < B, (screened, fallback)
    for j in range(B):
        msg = oracle.check_parity(db, qs[j], k, idx[j], sims[j])
        assert msg is None, (j, msg)
    # a second call starts from a clean mask
    idx2, sims2 = m._search_block(torch.from_numpy(qs[60:]), k)
    assert np.array_equal(idx2, idx[60:]) and np.array_equal(sims2, sims[60:])
    assert _stats(m)[1] == fallback


def test_screening_with_nan_zero_and_duplicate_rows():
    n, d, B, k = 9000, 64, 128, 6
    db = synth.unit_rows(71, n, d)
    db[5] = np.nan
    db[7] = 0.0
    db[100] = db[50]
    db[8000] = db[50]
    qs = synth.queries_for(72, db, B)
    qs[3] = db[50]
    qs[9] = 0.0
    m = _matcher(db)
    items, sims = m.search_batch(qs, k)
    assert _stats(m)[0] == B
    for j in range(B):
        assert items[j][0] == 5 and np.isnan(sims[j][0])          # NaN sorts first (torch.argsort descending)
    assert items[3][1:4] == [50, 100, 8000]                        # exact duplicates: lower row first
    assert items[9][1:] == [0, 1, 2, 3, 4] and np.all(sims[9][1:] == 0.0)   # zero query: all sims 0, row order
    for j in (0, 3, 64):
        msg = oracle.check_parity(np.delete(db, 5, axis=0), qs[j], k - 1,
                                  np.asarray([r - (r > 5) for r in items[j][1:]]), sims[j][1:])
        assert msg is None, (j, msg)


def test_screening_sees_interleaved_adds_and_sharded_keys():
    """Rows appended between two screened searches are found; the key output used by the row-sharded search
    (cslam_nns_search_keys, global row = local * world + rank) goes through the re-scoring kernel too."""
    d, B, k = 96, 128, 10
    db = synth.unit_rows(501, 24000, d)
    from cslam_b200 import NearestNeighborsMatching
    m = NearestNeighborsMatching()
    m.add_items(db[:16000], list(range(16000)))
    qs = np.ascontiguousarray(db[15990:15990 + B])
    items, _ = m.search_batch(qs, 1)
    assert [it[0] for it in items[:10]] == list(range(15990, 16000))
    m.add_items(db[16000:], list(range(16000, 24000)))
    items, sims = m.search_batch(qs, 1)
    assert [it[0] for it in items] == list(range(15990, 15990 + B)) and np.all(np.abs(sims[:, 0] - 1) < 1e-6)
    assert _stats(m)[0] == 2 * B
    keys = torch.zeros((B, k), dtype=torch.int64, device="cuda")
    q_dev = torch.from_numpy(qs).cuda()
    rc = m._h.lib.cslam_nns_search_keys(m._h.h, ctypes.c_void_p(q_dev.data_ptr()), B, d, 1, k, 4, 3,
                                        ctypes.c_void_p(keys.data_ptr()), m._stream())
    assert rc == 0
    idx, sims = m._search_block(q_dev, k)
    kk = keys.cpu().numpy().astype(np.uint64)
    rows = (~kk & np.uint64(0xFFFFFFFF)).astype(np.int64)
    assert np.array_equal(rows, idx * 4 + 3)
