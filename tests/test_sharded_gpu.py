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
