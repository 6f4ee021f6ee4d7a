"""cslam_nns_search_static — the single-query search that allocates nothing (include/cslam_nns.h; no counterpart
in the reference) — against the existing entry point
cslam_nns_search(B = 1) on the same rows, through the C-ABI, on two calls in a row.

Collected last on purpose (file name).
"""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

RTOL, ATOL = 1e-5, 1e-6      # the tolerance of the other parity tests


def _static_search(lib, db, q, k, out_idx, out_sim, stream=0):
    err = ctypes.create_string_buffer(512)
    rc = lib.cslam_nns_search_static(db.device.index, db.data_ptr(), db.shape[0], db.stride(0), db.shape[1],
                                      q.data_ptr(), k, out_idx.data_ptr(), out_sim.data_ptr(), stream, err, 512)
    return rc, err.value.decode()


@pytest.mark.parametrize("rows,dim,k", [(20_000, 512, 10), (3_001, 256, 1), (5_000, 4096, 32), (7, 128, 10)])
def test_static_scratch_search_equals_search_twice_in_a_row(rows, dim, k):
    from cslam_b200 import NearestNeighborsMatching, _native
    lib = _native.load()
    dev = torch.device("cuda", 0)
    gen = torch.Generator(device=dev)
    gen.manual_seed(rows + dim)
    db = torch.rand((rows, dim), generator=gen, device=dev, dtype=torch.float32)
    db /= db.norm(dim=1, keepdim=True)
    qs = torch.rand((2, dim), generator=gen, device=dev, dtype=torch.float32)
    qs[1] = db[rows // 2] + 0.01 * torch.randn((dim,), generator=gen, device=dev)   # a planted neighbour
    m = NearestNeighborsMatching(dim=dim)
    h = m._open()
    m._append(h, db)
    m.n += rows

    # garbage in the outputs, and the scratch still holds whatever the previous call (or test) left there
    out_idx = torch.full((2, k), 12345, dtype=torch.int64, device=dev)
    out_sim = torch.full((2, k), 7.0, dtype=torch.float32, device=dev)
    for j in range(2):                        # two calls in a row on one stream share the static scratch
        rc, msg = _static_search(lib, db, qs[j], k, out_idx[j], out_sim[j])
        assert rc == 0, msg
    torch.cuda.synchronize()
    first = (out_idx.clone(), out_sim.clone())
    # again, now that the kernels' module is loaded: the device's free memory must not move (no allocation)
    out_idx.fill_(54321)
    out_sim.fill_(-3.0)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info(dev)[0]
    for j in range(2):
        rc, msg = _static_search(lib, db, qs[j], k, out_idx[j], out_sim[j])
        assert rc == 0, msg
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info(dev)[0] >= free_before
    assert torch.equal(first[0], out_idx) and torch.equal(first[1], out_sim)
    for j in range(2):
        ref_idx, ref_sim = m._search_block(qs[j:j + 1], k)
        got_idx, got_sim = out_idx[j].cpu().numpy(), out_sim[j].cpu().numpy()
        kk = min(k, rows)
        assert np.all(got_idx[kk:] == -1) and np.all(np.isneginf(got_sim[kk:]))
        assert np.all(np.abs(got_sim[:kk] - ref_sim[0][:kk]) <= RTOL * np.abs(ref_sim[0][:kk]) + ATOL)
        differ = got_idx[:kk] != ref_idx[0][:kk]
        # the same kernel on the same bytes: identical rows (a difference could only be a tie in similarity)
        assert not differ.any() or np.all(np.abs(got_sim[:kk][differ] - ref_sim[0][:kk][differ]) <= ATOL)
    assert int(out_idx[1, 0]) == rows // 2


def test_static_scratch_search_rejects_sizes_beyond_its_maximum():
    from cslam_b200 import _native
    lib = _native.load()
    dev = torch.device("cuda", 0)
    db = torch.zeros((64, 64), device=dev)
    q = torch.zeros((64,), device=dev)
    oi = torch.zeros((64,), dtype=torch.int64, device=dev)
    os_ = torch.zeros((64,), device=dev)
    rc, msg = _static_search(lib, db, q, 33, oi, os_)
    assert rc == -1 and "CSLAM_NNS_STATIC_MAX_K" in msg and "33" in msg
    rc, msg = _static_search(lib, db, q, 0, oi, os_)
    assert rc == -1 and "k = 0" in msg
    err = ctypes.create_string_buffer(256)
    rc = lib.cslam_nns_search_static(0, db.data_ptr(), 1 << 32, 64, 64, q.data_ptr(), 1, oi.data_ptr(),
                                      os_.data_ptr(), 0, err, 256)
    assert rc == -1 and b"2^32" in err.value
    rc = lib.cslam_nns_search_static(0, db.data_ptr() + 4, 8, 64, 64, q.data_ptr(), 1, oi.data_ptr(),
                                      os_.data_ptr(), 0, err, 256)
    assert rc == -1 and b"aligned" in err.value
