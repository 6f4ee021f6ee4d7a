"""CPU restatement of the reference's ScanContext matcher — TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this module; the product
(cslam_b200/) never does.  Parity is PINNED: tests/test_sc_oracle.py checks every function here against
golden vectors produced by running the unmodified reference (tests/golden/make_golden_sc.py imports
/root/reference/cslam/lidar_pr/scancontext_matching.py).

Follows, line by line:
  sc2rk          scancontext_utils.py:75-76
  distance_sc    scancontext_utils.py:78-113 (vectorised over columns; the sum over columns is taken in the
                 reference's order j = 0..S-1 with a Python loop so that rounding matches)
  shortlist      scancontext_matching.py:58-63 (scipy KDTree.query(k) == the k Euclidean-nearest rows, nearest
                 first; brute force here)
  search         scancontext_matching.py:47-87
"""
import numpy as np


def sc2rk(sc):
    return np.mean(sc, axis=1)


def distance_sc(sc1, sc2):
    """(dist, yaw_diff) of scancontext_utils.distance_sc."""
    r, s = sc1.shape
    n1 = np.sqrt(np.sum(sc1 * sc1, axis=0))
    n2 = np.sqrt(np.sum(sc2 * sc2, axis=0))
    any1 = np.any(sc1, axis=0)
    any2 = np.any(sc2, axis=0)
    with np.errstate(divide="ignore", invalid="ignore"):
        cos = (sc1.T @ sc2) / np.outer(n1, n2)       # [a, j]: candidate column a vs query column j
    sims = np.zeros(s)
    for i in range(s):
        total = 0.0
        engaged = 0
        for j in range(s):
            a = (j - (i + 1)) % s                      # np.roll by one column per iteration (:87)
            if not any1[a] or not any2[j]:
                continue
            total = total + cos[a, j]
            engaged += 1
        sims[i] = 0.0 if engaged == 0 else total / engaged
    yaw_diff = int(np.argmax(sims)) + 1
    return 1 - np.max(sims), yaw_diff


def shortlist(ringkeys, query_rk, num_candidates):
    """Rows of the `num_candidates` nearest ring keys, nearest first (ties: lower row); -1 pads."""
    d2 = np.sum((ringkeys - query_rk[None, :]) ** 2, axis=1)
    order = np.lexsort((np.arange(len(d2)), d2))[:num_candidates]
    out = np.full(num_candidates, -1, np.int64)
    out[:len(order)] = order
    return out


def search(scancontexts, items, query, shape=(20, 60), num_candidates=10):
    """([item], [similarity]) like ScanContextMatching.search; scancontexts is (n, R*S) float64."""
    n = scancontexts.shape[0]
    if n < 1:
        return [None], [None]
    scs = scancontexts.reshape(n, shape[0], shape[1])
    rks = np.mean(scs, axis=2)
    q = np.asarray(query, np.float64).reshape(shape)
    cand = shortlist(rks, sc2rk(q), num_candidates)
    nn_dist, nn_idx = 1.0, None
    for c in cand:
        if c < 0:
            continue
        dist, _yaw = distance_sc(scs[c], q)
        if dist < nn_dist:
            nn_dist, nn_idx = dist, int(c)
    if nn_idx is None:
        return [items[0]], [0.0]
    return [items[nn_idx]], [1 - nn_dist]
