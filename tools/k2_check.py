#!/usr/bin/env python
"""Batched tcgen05 kernel (variant 2) against the single-query kernel (variant 1) and the oracle."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import synth  # noqa: E402
from cslam_b200 import NearestNeighborsMatching  # noqa: E402
from oracle import nns_oracle as oracle  # noqa: E402

shapes = [(300, 32, 4, 3), (1000, 128, 16, 10), (5000, 100, 37, 10), (20000, 256, 256, 10),
          (3000, 4096, 64, 5), (70000, 512, 200, 1), (129, 36, 2, 32)]
if len(sys.argv) > 1:
    shapes = [tuple(int(v) for v in a.split(",")) for a in sys.argv[1:]]
bad = 0
for n, d, B, k in shapes:
    db = synth.unit_rows(n + d, n, d)
    qs = synth.queries_for(n + d + 1, db, B)
    m = NearestNeighborsMatching()
    m.add_items(db, list(range(n)))
    lib, h = m._h.lib, m._h.h
    lib.cslam_nns_set_variant(h, 2)
    t0 = time.time()
    idx2, s2 = m._search_block(torch.from_numpy(qs), k)
    torch.cuda.synchronize()
    t2 = time.time() - t0
    lib.cslam_nns_set_variant(h, 1)
    idx1, s1 = m._search_block(torch.from_numpy(qs), k)
    nbad = 0
    worst = 0.0
    for j in range(B):
        msg = oracle.check_parity(db, qs[j], k, idx2[j], s2[j])
        if msg is not None:
            nbad += 1
            if nbad <= 3:
                print("   q%d: %s" % (j, msg))
        worst = max(worst, float(np.max(np.abs(s2[j] - s1[j]) / (np.abs(s1[j]) + 1e-6))))
    same = int((idx1 == idx2).all(axis=1).sum())
    print("n=%d d=%d B=%d k=%d: parity failures %d/%d, identical to K1 on %d/%d, max rel diff vs K1 %.2e (%.1f ms)"
          % (n, d, B, k, nbad, B, same, B, worst, t2 * 1e3), flush=True)
    bad += nbad
print("K2 CHECK", "FAILED" if bad else "OK")
sys.exit(1 if bad else 0)
