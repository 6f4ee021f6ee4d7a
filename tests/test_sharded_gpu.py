"""Row-sharded search on the GPU.  A single box GPU plays every rank in turn (one kernel pass per
shard, sequentially — no two launches wait on each other), the per-rank keys are stacked the way the
all-gather would deliver them and merged by kernel K4 through ``cslam_nns_merge_keys``.  The result
must be bit-identical to the unsharded search and agree with the oracle."""
import ctypes
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

pytestmark = pytest.mark.gpu


def _shard_keys(M, native, shard_rows, q_dev, k, world, rank):
    m = M.NearestNeighborsMatching(dim=int(q_dev.shape[1]))
    if len(shard_rows):
        m.add_items(torch.from_numpy(shard_rows), list(range(len(shard_rows))))
    B = int(q_dev.shape[0])
    keys = torch.zeros((B, k), dtype=torch.int64, device=q_dev.device)
    h = m._h
    rc = h.lib.cslam_nns_search_keys(h.h, ctypes.c_void_p(q_dev.data_ptr()), B, q_dev.shape[1], 1, k,
                                     world, rank, ctypes.c_void_p(keys.data_ptr()), m._stream())
    native.check(rc, h.h)
    return keys, m


@pytest.mark.parametrize("world,n,d,B,k", [(2, 4001, 256, 32, 10), (8, 20000, 128, 256, 10),
                                           (4, 3000, 100, 1, 1), (3, 7, 64, 5, 10)])
def test_sharded_equals_unsharded_and_oracle(world, n, d, B, k):
    import synth
    from oracle import nns_oracle
    from cslam_b200 import _native, nns_matching as M
    db = synth.unit_rows(31, n, d, "rand")
    qs = synth.queries_for(32, db, B, "rand")
    q_dev = torch.from_numpy(qs).cuda()
    parts, keep = [], []
    for r in range(world):
        keys, m = _shard_keys(M, _native, db[r::world], q_dev, k, world, r)
        parts.append(keys)
        keep.append(m)
    gathered = torch.stack(parts).contiguous()
    out_idx = torch.empty((B, k), dtype=torch.int64, device="cuda")
    out_sim = torch.empty((B, k), dtype=torch.float32, device="cuda")
    lib = _native.load()
    rc = lib.cslam_nns_merge_keys(0, ctypes.c_void_p(gathered.data_ptr()), world, B, k,
                                  ctypes.c_void_p(out_idx.data_ptr()), ctypes.c_void_p(out_sim.data_ptr()),
                                  1, ctypes.c_void_p(torch.cuda.current_stream().cuda_stream), None)
    _native.check(rc, None)
    idx, sims = out_idx.cpu().numpy(), out_sim.cpu().numpy()

    whole = M.NearestNeighborsMatching(dim=d)
    whole.add_items(torch.from_numpy(db), list(range(n)))
    widx, wsims = whole._search_block(q_dev, k)
    kk = widx.shape[1]
    for b in range(B):
        err = nns_oracle.check_parity(db, qs[b], k, idx[b, :kk], sims[b, :kk])
        assert err is None, err
    # same kernels on a different partition: identical rows unless two sims are within tolerance
    same = idx[:, :kk] == widx
    if not same.all():
        assert np.allclose(sims[:, :kk][~same], wsims[~same], rtol=1e-5, atol=1e-6)
    assert (idx[:, kk:] == -1).all()


@pytest.mark.parametrize("world,n,d,B,k", [(2, 4001, 256, 32, 10), (8, 20000, 128, 256, 10), (4, 3000, 100, 1, 1),
                                           (3, 7, 64, 5, 10), (8, 9000, 64, 300, 32)])
def test_peer_exchange_kernel_equals_allgather_merge(world, n, d, B, k):
    """K5 (csrc/k5_exchange.cu), the fused exchange + merge over peer memory, against the all-gather + K4
    route on the same per-rank keys.  One GPU plays every rank: `world` exchange objects in this process are
    connected by pointer, every rank pushes (phase 1), then every rank waits + merges (phase 2) — the flags
    are already there, so no launch waits for another.  Two consecutive calls exercise both buffer parities
    and a changed batch."""
    import synth
    from cslam_b200 import _native, nns_matching as M
    lib = _native.load()
    xs = (ctypes.c_void_p * world)()
    for r in range(world):
        x = ctypes.c_void_p()
        assert lib.cslam_nns_xchg_create(0, world, r, ctypes.byref(x)) == 0, lib.cslam_nns_xchg_last_error(None)
        xs[r] = x
    try:
        for r in range(world):
            assert lib.cslam_nns_xchg_connect_local(xs[r], xs) == 0, lib.cslam_nns_xchg_last_error(xs[r])
        db = synth.unit_rows(41, n, d, "rand")
        stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        for call, Bc in enumerate((B, max(1, B // 2), B)):
            qs = synth.queries_for(42 + call, db, Bc, "rand")
            q_dev = torch.from_numpy(qs).cuda()
            parts, keep = [], []
            for r in range(world):
                keys, m = _shard_keys(M, _native, db[r::world], q_dev, k, world, r)
                parts.append(keys)
                keep.append(m)
            gathered = torch.stack(parts).contiguous()
            ref_idx = torch.empty((Bc, k), dtype=torch.int64, device="cuda")
            ref_sim = torch.empty((Bc, k), dtype=torch.float32, device="cuda")
            rc = lib.cslam_nns_merge_keys(0, ctypes.c_void_p(gathered.data_ptr()), world, Bc, k,
                                          ctypes.c_void_p(ref_idx.data_ptr()), ctypes.c_void_p(ref_sim.data_ptr()),
                                          1, stream, None)
            _native.check(rc, None)
            for r in range(world):
                assert lib.cslam_nns_xchg_merge(xs[r], ctypes.c_void_p(parts[r].data_ptr()), Bc, k, 1, None,
                                                None, stream) == 0, lib.cslam_nns_xchg_last_error(xs[r])
            for r in range(world):
                idx = torch.full((Bc, k), -7, dtype=torch.int64, device="cuda")
                sim = torch.full((Bc, k), 7.0, dtype=torch.float32, device="cuda")
                assert lib.cslam_nns_xchg_merge(xs[r], ctypes.c_void_p(parts[r].data_ptr()), Bc, k, 2,
                                                ctypes.c_void_p(idx.data_ptr()), ctypes.c_void_p(sim.data_ptr()),
                                                stream) == 0, lib.cslam_nns_xchg_last_error(xs[r])
                torch.cuda.synchronize()
                assert lib.cslam_nns_xchg_status(xs[r]) == 0
                assert torch.equal(idx, ref_idx), (call, r)
                assert torch.equal(sim, ref_sim), (call, r)
        # limits are refused, not truncated
        assert lib.cslam_nns_xchg_merge(xs[0], ctypes.c_void_p(parts[0].data_ptr()), 1, 33, 3, None, None,
                                        stream) == _native.EINVAL
    finally:
        torch.cuda.synchronize()
        for r in range(world):
            lib.cslam_nns_xchg_destroy(xs[r])


def test_peer_exchange_timeout_is_an_error_not_a_hang(monkeypatch):
    """A rank whose peer never publishes must get an error on its next call, and the GPU must stay usable."""
    from cslam_b200 import _native
    monkeypatch.setenv("CSLAM_NNS_XCHG_TIMEOUT_S", "0.2")
    lib = _native.load()
    xs = (ctypes.c_void_p * 2)()
    for r in range(2):
        x = ctypes.c_void_p()
        assert lib.cslam_nns_xchg_create(0, 2, r, ctypes.byref(x)) == 0
        xs[r] = x
    try:
        for r in range(2):
            assert lib.cslam_nns_xchg_connect_local(xs[r], xs) == 0
        keys = torch.zeros((4, 3), dtype=torch.int64, device="cuda")
        idx = torch.empty((4, 3), dtype=torch.int64, device="cuda")
        sim = torch.empty((4, 3), dtype=torch.float32, device="cuda")
        stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        args = (ctypes.c_void_p(keys.data_ptr()), 4, 3, 3, ctypes.c_void_p(idx.data_ptr()),
                ctypes.c_void_p(sim.data_ptr()), stream)
        assert lib.cslam_nns_xchg_merge(xs[0], *args) == 0      # rank 1 never calls
        torch.cuda.synchronize()
        assert lib.cslam_nns_xchg_status(xs[0]) == _native.ECUDA
        assert lib.cslam_nns_xchg_merge(xs[0], *args) == _native.ECUDA
        assert b"did not publish" in lib.cslam_nns_xchg_last_error(xs[0])
        assert float(torch.ones(8, device="cuda").sum()) == 8.0   # the device is alive
    finally:
        for r in range(2):
            lib.cslam_nns_xchg_destroy(xs[r])


@pytest.mark.parametrize("world,n,d,B,k", [(2, 4001, 256, 32, 10), (4, 3000, 100, 1, 1), (3, 2, 64, 5, 10)])
def test_search_sharded_single_call_equals_unsharded(world, n, d, B, k):
    """cslam_nns_search_sharded (local search + exchange + merge in ONE C-ABI call) with one GPU
    playing every rank: phase 1 (search + push) on every rank, then phase 2 (wait + merge) on every rank.
    Every rank must end with the unsharded search's rows and bit-identical similarities; a rank without rows
    (n < world) takes part with an empty shard."""
    import synth
    from cslam_b200 import _native, nns_matching as M
    lib = _native.load()
    db = synth.unit_rows(61, n, d, "rand")
    qs = synth.queries_for(62, db, B, "rand")
    q_dev = torch.from_numpy(qs).cuda()
    whole = M.NearestNeighborsMatching(dim=d)
    whole.add_items(torch.from_numpy(db), list(range(n)))
    widx, wsims = whole._search_block(q_dev, k)
    kk = widx.shape[1]
    shards = []
    for r in range(world):
        m = M.NearestNeighborsMatching(dim=d)
        if len(db[r::world]):
            m.add_items(torch.from_numpy(db[r::world]), list(range(len(db[r::world]))))
        shards.append(m)
    xs = (ctypes.c_void_p * world)()
    for r in range(world):
        x = ctypes.c_void_p()
        assert lib.cslam_nns_xchg_create(0, world, r, ctypes.byref(x)) == 0, lib.cslam_nns_xchg_last_error(None)
        xs[r] = x
    try:
        for r in range(world):
            assert lib.cslam_nns_xchg_connect_local(xs[r], xs) == 0
        stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        outs = []
        for phase in (1, 2):
            for r in range(world):
                idx = torch.full((B, k), -7, dtype=torch.int64, device="cuda")
                sim = torch.full((B, k), 7.0, dtype=torch.float32, device="cuda")
                h = shards[r]._h
                rc = lib.cslam_nns_search_sharded(h.h, xs[r], q_dev.data_ptr(), B, d, 1, k, r, world, phase,
                                                  idx.data_ptr(), sim.data_ptr(), stream)
                _native.check(rc, h.h)
                if phase == 2:
                    outs.append((idx, sim))
        torch.cuda.synchronize()
        for r, (idx, sim) in enumerate(outs):
            assert lib.cslam_nns_xchg_status(xs[r]) == 0
            assert np.array_equal(idx.cpu().numpy()[:, :kk], widx), r
            assert np.array_equal(sim.cpu().numpy()[:, :kk], wsims), r
            assert (idx.cpu().numpy()[:, kk:] == -1).all()
        h = shards[0]._h
        assert lib.cslam_nns_search_sharded(h.h, xs[0], q_dev.data_ptr(), B, d, 1, k, world, world, 3,
                                            None, None, stream) == _native.EINVAL     # rank out of range
    finally:
        torch.cuda.synchronize()
        for r in range(world):
            lib.cslam_nns_xchg_destroy(xs[r])
