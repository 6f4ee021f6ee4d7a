"""CPU ORACLE — test infrastructure only.  Nothing in ``cslam_b200/`` may import this module.

numpy restatement of the reference's nearest-neighbour search,
``/root/reference/cslam/nns_matching.py:57-63``:

    similarity = torch.nn.functional.cosine_similarity(view, qTensor)        (:61)
    ns = torch.argsort(similarity, descending=True)[:k]                      (:62)
    return [items[n] for n in ns], gather(similarity, 0, ns)                 (:63)

The arithmetic lives in a third-party dependency that is not vendored under /root/reference:
PyTorch (the reference env pins pytorch-1.12.0, ``spec-file.txt:115``; this image has 2.11.0+cu128,
whose behaviour is the de-facto oracle).  Its published formula (torch 2.x) is

    cos(x, q) = sum_i (x_i / max(|x|, eps)) * (q_i / max(|q|, eps)),   eps = 1e-8

i.e. a per-vector clamp of the norm.  Parity is PINNED: ``tests/golden/make_golden.py`` ran the
unmodified reference class (device forced to CPU) from /root/reference in the build container and
committed its outputs under ``tests/golden/``; ``tests/test_oracle.py`` checks this restatement
against every one of those vectors.
"""
import numpy as np

EPS = 1e-8


def cosine_similarities(db, q, dtype=np.float32):
    """db: (n, d), q: (d,) -> (n,) cosine similarities, computed in ``dtype``  (:61)."""
    db = np.asarray(db, dtype=dtype)
    q = np.asarray(q, dtype=dtype).reshape(-1)
    nx = np.maximum(np.sqrt((db * db).sum(axis=1, dtype=dtype)), dtype(EPS))
    nq = max(np.sqrt((q * q).sum(dtype=dtype)), dtype(EPS))
    return ((db @ q) / (nx * nq)).astype(dtype)


def search(db, q, k, dtype=np.float32):
    """Indices of the k most similar rows (descending, ties -> lower row) and their sims (:62-63)."""
    s = cosine_similarities(db, q, dtype)
    key = np.where(np.isnan(s), np.inf, s)  # NaN sorts first under torch.argsort(descending=True)
    order = np.argsort(-key, kind="stable")[:max(int(k), 0)]
    return order.astype(np.int64), s[order]


def search_batch(db, queries, k, dtype=np.float32):
    """Loop of the single-query search — the reference API is single-query only (:60)."""
    queries = np.asarray(queries).reshape(-1, np.asarray(db).shape[1])
    out = [search(db, q, k, dtype) for q in queries]
    return np.stack([o[0] for o in out]), np.stack([o[1] for o in out])


def search_sharded(db, queries, k, world, dtype=np.float32):
    """Row-striped sharding has no reference counterpart; its oracle is the search on the whole DB
    (SURVEY §8e).  Provided so tests read symmetrically."""
    del world
    return search_batch(db, queries, k, dtype)


def neighbour_gaps(db, q, k):
    """fp64 similarities sorted descending: used to decide where an index mismatch is a tie under the
    parity rule (|delta sim| <= 1e-5*|s| + 1e-6)."""
    s = cosine_similarities(db, q, np.float64)
    order = np.argsort(-s, kind="stable")
    return order, s[order]


def check_parity(db, q, k, got_idx, got_sim, rtol=1e-5, atol=1e-6):
    """Parity rule of BASELINE.md §5.  Returns None when satisfied, else a message.

    * sims: |got - ref| <= rtol*|ref| + atol, position by position;
    * indices: equal to the fp64 reference order, except that a returned row may differ when its own
      fp64 similarity is within tolerance of the reference row's at that position (a tie)."""
    got_idx = np.asarray(got_idx).reshape(-1)
    got_sim = np.asarray(got_sim, dtype=np.float64).reshape(-1)
    s64 = cosine_similarities(db, q, np.float64)
    order = np.argsort(-s64, kind="stable")
    kk = min(int(k), len(s64))
    if len(got_idx) != kk or len(got_sim) != kk:
        return f"expected {kk} results, got {len(got_idx)}/{len(got_sim)}"
    ref_sim = s64[order[:kk]]
    tol = rtol * np.abs(ref_sim) + atol
    bad = np.abs(got_sim - ref_sim) > tol
    if bad.any():
        j = int(np.argmax(bad))
        return f"sim[{j}] = {got_sim[j]!r} vs reference {ref_sim[j]!r}"
    if len(set(got_idx.tolist())) != kk:
        return "duplicate rows returned"
    if (got_idx < 0).any() or (got_idx >= len(s64)).any():
        return "row index out of range"
    own = s64[got_idx]
    bad = (got_idx != order[:kk]) & (np.abs(own - ref_sim) > tol)
    if bad.any():
        j = int(np.argmax(bad))
        return (f"idx[{j}] = {got_idx[j]} (sim {own[j]!r}) vs reference {order[j]} "
                f"(sim {ref_sim[j]!r}): not a tie")
    return None
