"""Deterministic synthetic descriptors shared by the golden generator, the tests and bench.py.

A counter-based generator (splitmix64 over the element index) in plain numpy integer arithmetic:
bit-identical on every machine and numpy version, so fixtures only need to store OUTPUTS.
"""
import numpy as np

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def _splitmix64(x):
    x = (x + np.uint64(0x9E3779B97F4A7C15)) & _M64
    z = x
    z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M64
    z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M64
    return z ^ (z >> np.uint64(31))


def uniform(seed, shape, offset=0):
    """U[0,1) float64 array; element i depends only on (seed, offset + i)."""
    n = int(np.prod(shape))
    with np.errstate(over="ignore"):
        idx = np.arange(offset, offset + n, dtype=np.uint64)
        bits = _splitmix64(idx ^ _splitmix64(np.uint64(seed) + np.uint64(0x1234567)))
    return ((bits >> np.uint64(11)).astype(np.float64) * (1.0 / (1 << 53))).reshape(shape)


def normal(seed, shape):
    """Standard normal via Box-Muller on two uniform streams."""
    u1 = uniform(seed, shape)
    u2 = uniform(seed ^ 0x5DEECE66D, shape)
    return np.sqrt(-2.0 * np.log(1.0 - u1)) * np.cos(2.0 * np.pi * u2)


def unit_rows(seed, n, d, dist="rand"):
    """(n, d) float32 rows, L2-normalised like tests/test_sparse_matching.py:57-58 of the reference."""
    x = uniform(seed, (n, d)) if dist == "rand" else normal(seed, (n, d))
    x = x / np.maximum(np.linalg.norm(x, axis=1, keepdims=True), 1e-30)
    return x.astype(np.float32)


def queries_for(seed, db, nq, dist="rand", every=3, noise=0.05):
    """Queries from the same distribution; every `every`-th one is a noisy copy of a DB row so that
    true near-duplicates exist."""
    n, d = db.shape
    qs = unit_rows(seed, nq, d, dist)
    eps = normal(seed + 77, (nq, d)).astype(np.float32)
    for j in range(0, nq, every):
        v = db[(7 * j + 1) % n] + noise * eps[j]
        qs[j] = v / np.linalg.norm(v)
    return qs


def swarm_inputs(robots=4, n=300, d=64):
    """4 robots x 300 keyframes x 64-d with planted inter-robot overlaps and intra-robot revisits."""
    descs = [unit_rows(70 + r, n, d, "rand") for r in range(robots)]
    eps = normal(79, (robots, n, d)).astype(np.float32)
    for i in range(100, n, 7):          # robot 0 revisits its own place of 90 keyframes ago
        v = descs[0][i - 90] + 0.04 * eps[0][i]
        descs[0][i] = v / np.linalg.norm(v)
    for i in range(4, n, 9):            # ... and a place of 3 keyframes ago (inside the min-gap window)
        v = descs[0][i - 3] + 0.04 * eps[0][i]
        descs[0][i] = v / np.linalg.norm(v)
    for r in range(1, robots):          # the other robots see places of robot 0
        for i in range(r, n, 11):
            v = descs[0][(5 * i + r) % n] + 0.05 * eps[r][i]
            descs[r][i] = v / np.linalg.norm(v)
    return descs


def scancontexts(seed, n, shape=(20, 60), empty_frac=0.15):
    """(n, rings*sectors) float64 ScanContext-like descriptors: non-negative "max heights" with a share of
    empty bins and a few completely empty sectors (columns), as ptcloud2sc produces for real scans
    (reference cslam/lidar_pr/scancontext_utils.py:45-73)."""
    r, s = shape
    h = 4.0 * uniform(seed, (n, r, s))
    hole = uniform(seed + 101, (n, r, s)) < empty_frac
    h[hole] = 0.0
    dead = uniform(seed + 202, (n, 1, s)) < 0.08          # sectors with no return at all
    h = np.where(dead, 0.0, h)
    return h.reshape(n, r * s)


def sc_queries_for(seed, db, nq, shape=(20, 60), every=2, noise=0.05):
    """Queries: fresh descriptors, every `every`-th one a DB scan seen again under a yaw rotation (a column
    roll) with a little height noise — the case the column-shift distance exists for."""
    r, s = shape
    n = db.shape[0]
    qs = scancontexts(seed, nq, shape)
    eps = normal(seed + 77, (nq, r, s))
    for j in range(0, nq, every):
        src = db[(5 * j + 3) % n].reshape(r, s)
        rolled = np.roll(src, (7 * j + 1) % s, axis=1)
        v = np.where(rolled > 0, np.maximum(rolled + noise * eps[j], 0.0), 0.0)
        qs[j] = v.reshape(-1)
    return qs
