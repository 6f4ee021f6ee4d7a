"""Parity at the FULL sizes BASELINE.json names (last sentence): the library's results on a
subsample of queries against the reference's own three torch calls (cslam/nns_matching.py:61-63, restated in
tests/refops.py) on the SAME resident tensors — 100k x 512 (k = 1, 10), 2M x 256 with interleaved adds,
1M x 4096 with batches of 256 (a 16.4 GB database: byte offsets beyond 2^32, 3906 super tiles) and 10M x 256
with batches of 256.  The databases are generated on the device (seeded) and bulk-appended through the C-ABI.
"""
import numpy as np
import pytest
import torch

import refops

pytestmark = pytest.mark.gpu


def _bulk(rows, dim, seed, chunk_bytes=1 << 30):
    """Matcher with `rows` seeded unit rows appended through cslam_nns_add (device source); items = row ids
    are NOT materialised (results are compared as row numbers)."""
    from cslam_b200 import NearestNeighborsMatching
    dev = torch.device("cuda", 0)
    m = NearestNeighborsMatching(dim=dim)
    h = m._open()
    gen = torch.Generator(device=dev)
    gen.manual_seed(seed)
    chunk = max(1, min(rows, chunk_bytes // (dim * 4)))
    done, probe = 0, None
    while done < rows:
        c = min(chunk, rows - done)
        x = torch.rand((c, dim), generator=gen, device=dev, dtype=torch.float32)
        x /= x.norm(dim=1, keepdim=True)
        if probe is None:
            probe = x[:1024].clone()
        m._append(h, x)
        m.n += c
        done += c
        del x
    return m, probe


def _queries(dim, nq, probe, seed):
    dev = probe.device
    gq = torch.Generator(device=dev)
    gq.manual_seed(seed)
    q = torch.rand((nq, dim), generator=gq, device=dev, dtype=torch.float32)
    q[::10] = probe[:nq][::10] + 0.02 * torch.randn((q[::10].shape[0], dim), generator=gq, device=dev)
    q /= q.norm(dim=1, keepdim=True)
    return q


def _check(m, q, k, idx, sims, nsub=32):
    step = max(1, q.shape[0] // nsub)
    sub = torch.arange(0, q.shape[0], step, device=q.device)[:nsub]
    rep = refops.check_against_reference(m.data[:m.n], q[sub].contiguous(), k, np.asarray(idx)[sub.cpu().numpy()],
                                         np.asarray(sims)[sub.cpu().numpy()])
    assert rep.ok, rep.as_dict()
    return rep


def test_c2_100k_x_512_single_query_k1_and_k10():
    m, probe = _bulk(100_000, 512, 2)
    q = _queries(512, 40, probe, 12)
    for k in (1, 10):
        res = [m._search_block(q[j:j + 1], k) for j in range(q.shape[0])]
        assert m._h.lib.cslam_nns_last_variant(m._h.h) == 1
        idx = np.concatenate([r[0] for r in res])
        sims = np.concatenate([r[1] for r in res])
        _check(m, q, k, idx, sims)
    # planted near-duplicates are found (query 0 is a noisy copy of row 0)
    i0, s0 = m._search_block(q[0:1], 1)
    assert int(i0[0, 0]) == 0 and s0[0, 0] > 0.9


def test_c4_2m_x_256_with_interleaved_adds():
    m, probe = _bulk(2_000_000, 256, 4)
    n0 = m.n
    q = _queries(256, 32, probe, 14)
    new = _queries(256, 24, probe, 15).cpu().numpy()
    for j in range(len(new)):             # config 4: an add_item between the searches; every search sees it
        m.add_item(new[j], n0 + j)
        item, sim = m.search_best(new[j])
        assert item == n0 + j and abs(float(sim) - 1.0) < 1e-5
        m.items.pop(n0 + j)               # keep `items` sparse: results below are compared as row numbers
    assert m.n == n0 + len(new) and len(m.data) == 2_048_000      # capacity: reference sequence 1000 * 2^11
    res = [m._search_block(q[j:j + 1], 1) for j in range(q.shape[0])]
    _check(m, q, 1, np.concatenate([r[0] for r in res]), np.concatenate([r[1] for r in res]))


@pytest.mark.parametrize("rows,dim,seed", [(1_000_000, 4096, 3), (10_000_000, 256, 5)])
def test_c3_c5_batched_256_top10_full_size(rows, dim, seed):
    m, probe = _bulk(rows, dim, seed)
    B, k = 256, 10
    q = _queries(dim, B, probe, seed + 10)
    idx, sims = m._search_block(q, k)
    assert m._h.lib.cslam_nns_last_variant(m._h.h) == 2
    assert idx.shape == (B, k) and np.all(sims[:, :-1] >= sims[:, 1:])
    assert np.all(idx[::10, 0] == np.arange(0, B, 10))          # every 10th query is a noisy copy of row j
    _check(m, q, k, idx, sims)
    # size-independent properties on ALL 256 queries: an fp64 re-scoring of the returned rows reproduces the
    # returned similarities, and the single-query kernel returns the same rows for a few of them
    view = m.data[:m.n]
    for j in range(0, B, 16):
        s64 = refops.fp64_similarity(view, torch.as_tensor(idx[j], device=view.device), q[j]).cpu().numpy()
        assert np.all(np.abs(sims[j] - s64) <= 1e-5 * np.abs(s64) + 1e-6)
    lib, h = m._h.lib, m._h.h
    lib.cslam_nns_set_variant(h, 1)
    for j in (0, 7, 130):
        i1, s1 = m._search_block(q[j:j + 1], k)
        np.testing.assert_allclose(s1[0], sims[j], rtol=2e-5, atol=2e-6)
        differ = i1[0] != idx[j]
        if differ.any():   # only where the two rows tie in fp64
            a = refops.fp64_similarity(view, torch.as_tensor(i1[0][differ], device=view.device), q[j])
            b = refops.fp64_similarity(view, torch.as_tensor(idx[j][differ], device=view.device), q[j])
            assert float((a - b).abs().max()) <= 2e-6
    lib.cslam_nns_set_variant(h, 0)
