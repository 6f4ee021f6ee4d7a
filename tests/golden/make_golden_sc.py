#!/usr/bin/env python
"""Golden vectors for the ScanContext matcher, produced by RUNNING THE UNMODIFIED REFERENCE
(/root/reference/cslam/lidar_pr/scancontext_matching.py, numpy + scipy KD-tree).  Build container only:

    python tests/golden/make_golden_sc.py

Inputs come from tests/synth.py (counter-based generator), so only outputs are stored.
"""
import os
import sys

import numpy as np

sys.path.insert(0, "/root/reference")
from cslam.lidar_pr.scancontext_matching import ScanContextMatching as RefSC  # noqa: E402
import cslam.lidar_pr.scancontext_utils as ref_utils  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import synth  # noqa: E402

# name, n items, n queries, seed
CASES = [("n7", 7, 4, 31), ("n120", 120, 6, 32), ("n1300", 1300, 6, 33)]


def main():
    out = {}
    for name, n, nq, seed in CASES:
        db = synth.scancontexts(seed, n)
        qs = synth.sc_queries_for(seed + 1000, db, nq)
        m = RefSC()
        for i in range(n):
            m.add_item(db[i], 5000 + i)
        items = np.empty(nq, np.int64)
        sims = np.empty(nq, np.float64)
        for j in range(nq):
            it, s = m.search(qs[j], 1)
            items[j], sims[j] = it[0], s[0]
        out[name + "/spec"] = np.array([n, nq, seed], np.int64)
        out[name + "/items"] = items
        out[name + "/sims"] = sims
        out[name + "/ringkeys"] = np.array(m.ringkeys[:n])
    # distance_sc on its own: pairs (db row a, query b) incl. a pair with empty columns and an all-zero scan
    db = synth.scancontexts(41, 6)
    qs = synth.sc_queries_for(1041, db, 4)
    db[5] = 0.0
    pairs = [(0, 0), (1, 1), (3, 0), (2, 3), (5, 1)]
    out["dist/pairs"] = np.array(pairs, np.int64)
    out["dist/values"] = np.array([ref_utils.distance_sc(db[a].reshape(20, 60), qs[b].reshape(20, 60))
                                   for a, b in pairs], np.float64)
    np.savez_compressed(os.path.join(HERE, "sc_cases.npz"), **out)
    print("wrote sc_cases.npz:", {k: v.shape for k, v in out.items()})


# NOTE: LoopClosureSparseMatching with frontend.sensor_type == "lidar" cannot be pinned: it hands a float32
# TORCH tensor to ScanContextMatching.add_item (loop_closure_sparse_matching.py:45-46), whose
# sc2rk = np.mean(tensor, axis=1) (scancontext_utils.py:76) raises TypeError (numpy forwards dtype=/out=
# keywords torch's mean() does not accept) — probed here with numpy 2.3 / torch 2.11.  The standalone matcher
# with numpy inputs (what GlobalDescriptorLoopClosureDetection's ScanContext.compute_embedding produces) works
# and is what the cases above pin.


if __name__ == "__main__":
    main()
