#!/usr/bin/env python
"""Small-shape walk through every kernel of libcslam_nns.so, meant to run under compute-sanitizer
(tools/sanitize.sh): K1 in all its row-width variants and all three select tails (k = 1, <= 32, <= 128), the
two-pass K3, both batched tilings (row-major k2_kernel, query-major k2q_kernel pair), the screening pass with its
re-scoring kernel and the masked fallback, K4, the candidate filter, the multi-database call, K5's two phases
(one GPU plays every rank: push on every rank, then wait + merge — no launch waits for another) and the
ScanContext kernels.  Every result is checked against the CPU oracle, so a sanitizer-clean run is also a correct
one.  Shapes are tiny because the tools slow kernels down 10-100x."""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import synth  # noqa: E402
from cslam_b200 import NearestNeighborsMatching as M, _native  # noqa: E402
from oracle import nns_oracle as oracle  # noqa: E402


def check(db, q, k, items, sims):
    msg = oracle.check_parity(db, q, k, np.asarray(items), sims)
    assert msg is None, msg


def main():
    only = set(sys.argv[1:])
    lib = _native.load()
    torch.cuda.set_device(0)

    def want(name):
        return not only or name in only

    if want("k1"):
        for n, d, ks in [(700, 8, (1, 5)), (900, 40, (1, 10)), (1500, 128, (1, 10, 40)), (600, 512, (1, 33)),
                         (300, 2048, (1, 10)), (2000, 36, (129,))]:
            db = synth.unit_rows(100 + d, n, d)
            m = M()
            m.add_items(db[: n // 2], list(range(n // 2)))
            for i in range(n // 2, n // 2 + 8):
                m.add_item(db[i], i)                          # single-row appends, capacity doubling
            m.add_items(db[n // 2 + 8:], list(range(n // 2 + 8, n)))
            for q in synth.queries_for(200 + d, db, 2):
                for k in ks:
                    items, sims = m.search(q, k)
                    check(db, q, k, items, sims)
        print("k1/k3 ok", flush=True)
    if want("k2"):
        for n, d, B, k in [(3000, 100, 8, 10), (2500, 64, 37, 5), (4000, 96, 128, 10), (1500, 200, 130, 32),
                           (5000, 128, 256, 10), (40000, 64, 256, 10)]:
            db = synth.unit_rows(300 + n, n, d)
            qs = synth.queries_for(400 + n, db, B)
            m = M()
            m.add_items(db, list(range(n)))
            items, sims = m.search_batch(qs, k)
            for j in range(0, B, max(1, B // 8)):
                check(db, qs[j], k, items[j], sims[j])
        # near-duplicate rows: the screening pass cannot prove its candidate set complete -> masked fallback
        db = synth.unit_rows(77, 6000, 64)
        db[1000:1400] = db[17] + 1e-4 * synth.normal(78, (400, 64)).astype(np.float32)
        qs = synth.queries_for(79, db, 128)
        qs[5] = db[17]
        m = M()
        m.add_items(db, list(range(len(db))))
        items, sims = m.search_batch(qs, 10)
        for j in (0, 5, 64, 127):
            check(db, qs[j], 10, items[j], sims[j])
        print("k2/k4/k6 ok", flush=True)
    if want("callers"):
        db = synth.unit_rows(1, 2000, 128)
        qs = synth.queries_for(2, db, 4)
        m = M()
        m.add_items(db, list(range(len(db))))
        others = []
        for r in range(3):
            o = M()
            o.add_items(db[r::3], list(range(r, len(db), 3)))
            others.append(o)
        for q in qs:
            for o, (item, sim) in zip(others, M.search_best_multi(others, q)):
                assert (item, sim) == o.search_best(q)
            items, sims = m.search(q, 10)
            got_items, got_sims, pos = m.search_candidates(q, 10, query_id=items[0], min_gap=5, threshold=0.5)
            assert got_items == items
        its, sims, pos = m.search_candidates(torch.from_numpy(qs), 1, mode=_native.FILTER_BEST, batched=True,
                                             threshold=0.8)
        assert len(its) == 4
        print("filter/multi ok", flush=True)
    if want("k5"):
        world, n, d, B, k = 4, 3000, 64, 32, 10
        db = synth.unit_rows(41, n, d)
        q_dev = torch.from_numpy(synth.queries_for(42, db, B)).cuda()
        xs = (ctypes.c_void_p * world)()
        for r in range(world):
            x = ctypes.c_void_p()
            assert lib.cslam_nns_xchg_create(0, world, r, ctypes.byref(x)) == 0
            xs[r] = x
        for r in range(world):
            assert lib.cslam_nns_xchg_connect_local(xs[r], xs) == 0
        stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        ms, outs = [], []
        for r in range(world):
            m = M(dim=d)
            m.add_items(torch.from_numpy(db[r::world]), list(range(len(db[r::world]))))
            ms.append(m)
        for call in range(2):
            outs = []
            for phase in (1, 2):
                for r in range(world):
                    idx = torch.empty((B, k), dtype=torch.int64, device="cuda")
                    sim = torch.empty((B, k), dtype=torch.float32, device="cuda")
                    h = ms[r]._h
                    rc = lib.cslam_nns_search_sharded(h.h, xs[r], q_dev.data_ptr(), B, d, 1, k, r, world, phase,
                                                      idx.data_ptr(), sim.data_ptr(), stream)
                    _native.check(rc, h.h)
                    if phase == 2:
                        outs.append((idx, sim))
            torch.cuda.synchronize()
            whole = M(dim=d)
            whole.add_items(torch.from_numpy(db), list(range(n)))
            wi, ws = whole._search_block(q_dev, k)
            for idx, sim in outs:
                assert np.array_equal(idx.cpu().numpy(), wi) and np.array_equal(sim.cpu().numpy(), ws)
        for r in range(world):
            lib.cslam_nns_xchg_destroy(xs[r])
        print("k5 ok", flush=True)
    if want("sc"):
        from cslam_b200.lidar_pr import ScanContextMatching
        base = synth.scancontexts(61, 300)
        sc = ScanContextMatching()
        for i in range(300):
            sc.add_item(base[i].reshape(20, 60), i)
        for q in synth.sc_queries_for(1061, base, 3):
            sc.search(q, 1)
        print("sc ok", flush=True)
    torch.cuda.synchronize()
    print("SANITIZE DRIVER OK", flush=True)


if __name__ == "__main__":
    main()
