"""The reference's search arithmetic on device tensors, and the full-size parity gate built on it.

TEST INFRASTRUCTURE (used by ``tests/`` and by ``bench.py``'s parity / ``gpu_reference`` legs only; never by
the product).  ``/root/reference`` does not exist on the benchmark box, so the three calls of
``/root/reference/cslam/nns_matching.py:61-63`` are restated here verbatim with stock torch ops:

    similarity = torch.nn.functional.cosine_similarity(view, qTensor)        (:61)
    ns = torch.argsort(similarity, descending=True)[:k]                      (:62)
    sims = torch.gather(similarity, 0, ns)                                   (:63)

on the SAME tensors the library searches (the zero-copy ``.data`` view of the HBM database).  cosine_similarity
is row-wise independent, so the view is walked in row chunks to keep torch's two (rows x dim) temporaries small;
the values are the ones an unchunked call gives.

Parity rule: a returned row must equal the reference's row at that position unless the
REFERENCE's own similarities of the two rows differ by less than ``rtol*|s| + atol`` (a tie by the reference's
own arithmetic); returned similarities must agree with the reference's within the same bound and with an fp64
re-scoring of the returned rows.
"""
import numpy as np
import torch

RTOL, ATOL = 1e-5, 1e-6


def reference_similarity(view, q, chunk_bytes=1 << 30):
    """nns_matching.py:61 on device: view (n, dim) fp32 (any row pitch), q (dim,) -> (n,) fp32."""
    n, dim = int(view.shape[0]), int(view.shape[1])
    rows = max(1, chunk_bytes // (dim * 4))
    qt = q.reshape(1, dim)
    if n <= rows:
        return torch.nn.functional.cosine_similarity(view, qt)
    out = torch.empty((n,), dtype=torch.float32, device=view.device)
    for lo in range(0, n, rows):
        out[lo:lo + rows] = torch.nn.functional.cosine_similarity(view[lo:lo + rows], qt)
    return out


def reference_search(view, q, k):
    """nns_matching.py:61-63 on device: (rows int64[k], sims fp32[k], similarity[n])."""
    sim = reference_similarity(view, q)
    ns = torch.argsort(sim, descending=True)[:k]
    return ns, torch.gather(sim, 0, ns), sim


def fp64_similarity(view, rows, q):
    """fp64 cosine similarity (per-vector eps clamp as torch) of the given rows of `view` with q."""
    x = view[rows].double()
    qd = q.double().reshape(1, -1)
    nx = x.norm(dim=1).clamp_min(1e-8)
    nq = qd.norm().clamp_min(1e-8)
    return (x * qd).sum(dim=1) / (nx * nq)


class ParityReport(object):
    """Accumulates the outcome of the parity rule over queries; ``as_dict()`` is what bench.py prints."""

    def __init__(self, what=""):
        self.what = what
        self.queries = 0
        self.non_tie = 0          # rows that differ from the reference's and are NOT a reference tie
        self.ties = 0             # rows that differ but the reference's own similarities tie within tolerance
        self.bad_sims = 0         # similarities outside rtol*|s|+atol of the reference's
        self.bad_fp64 = 0         # similarities outside the bound of an fp64 re-scoring of the returned row
        self.max_rel = 0.0
        self.max_abs64 = 0.0
        self.first = None

    def add(self, ref_rows, ref_sims, ref_sim_of_ours, our_rows, our_sims, our_fp64):
        """All arguments are 1-D arrays of k entries for one query (numpy, float64/int64)."""
        ref_rows, our_rows = np.asarray(ref_rows, np.int64), np.asarray(our_rows, np.int64)
        ref_sims = np.asarray(ref_sims, np.float64)
        our_sims = np.asarray(our_sims, np.float64)
        own = np.asarray(ref_sim_of_ours, np.float64)
        tol = RTOL * np.abs(ref_sims) + ATOL
        self.queries += 1
        with np.errstate(invalid="ignore"):
            both_nan = np.isnan(ref_sims) & np.isnan(our_sims)
            d = np.where(both_nan, 0.0, np.abs(our_sims - ref_sims))
            bad_s = ~(d <= tol)
            differ = our_rows != ref_rows
            tie = differ & (np.where(np.isnan(own) & np.isnan(ref_sims), 0.0, np.abs(own - ref_sims)) <= tol)
            non_tie = differ & ~tie
            d64 = np.where(np.isnan(our_fp64) & np.isnan(our_sims), 0.0, np.abs(our_sims - np.asarray(our_fp64)))
            bad64 = ~(d64 <= RTOL * np.abs(our_fp64) + ATOL)
        dup = len(set(our_rows.tolist())) != len(our_rows)
        self.bad_sims += int(bad_s.sum())
        self.ties += int(tie.sum())
        self.non_tie += int(non_tie.sum()) + (1 if dup else 0)
        self.bad_fp64 += int(bad64.sum())
        rel = d / np.maximum(np.abs(ref_sims), 1e-30)
        if rel.size:
            self.max_rel = max(self.max_rel, float(np.nanmax(rel)))
            self.max_abs64 = max(self.max_abs64, float(np.nanmax(d64)))
        if self.first is None and (bad_s.any() or non_tie.any() or bad64.any() or dup):
            j = int(np.argmax(bad_s | non_tie | bad64))
            self.first = "query %d pos %d: ours (row %d, sim %.9g, fp64 %.9g) reference (row %d, sim %.9g)%s" % (
                self.queries - 1, j, our_rows[j], our_sims[j], float(np.asarray(our_fp64)[j]), ref_rows[j],
                ref_sims[j], " duplicate rows" if dup else "")

    @property
    def ok(self):
        return self.queries > 0 and self.non_tie == 0 and self.bad_sims == 0 and self.bad_fp64 == 0

    def as_dict(self):
        d = {"queries": self.queries, "non_tie_mismatches": self.non_tie, "reference_ties": self.ties,
             "sims_out_of_tolerance": self.bad_sims, "sims_vs_fp64_out_of_tolerance": self.bad_fp64,
             "max_rel_sim_err": self.max_rel, "max_abs_err_vs_fp64": self.max_abs64,
             "rule": "rows equal the reference's unless the reference's own similarities tie within "
                     "1e-5*|s|+1e-6; sims within the same bound of the reference's and of an fp64 re-scoring",
             "against": self.what, "ok": self.ok}
        if self.first:
            d["first_failure"] = self.first
        return d


def check_against_reference(view, queries, k, our_rows, our_sims, rank=0, world=1, group=None,
                            what="torch cosine_similarity + argsort + gather on the same .data view"):
    """Full-size parity gate.  `view`: this rank's (n_local, dim) device view — global row g lives on rank
    g % world as local row g // world (world == 1: the whole database).  `queries` (nq, dim) device tensor,
    identical on every rank; `our_rows` (nq, k) GLOBAL rows and `our_sims` (nq, k) as the library returned them
    (device or host; identical on every rank).  Every rank scores its shard with the reference's own ops; the
    per-shard top-k are merged on the host like the reference would order the concatenated database.  Returns a
    ParityReport (meaningful on every rank)."""
    import torch.distributed as dist
    dev = view.device
    nq = int(queries.shape[0])
    our_rows_t = torch.as_tensor(np.asarray(our_rows.cpu() if torch.is_tensor(our_rows) else our_rows),
                                 dtype=torch.int64, device=dev).reshape(nq, -1)
    our_sims_t = torch.as_tensor(np.asarray(our_sims.cpu() if torch.is_tensor(our_sims) else our_sims),
                                 dtype=torch.float32, device=dev).reshape(nq, -1)
    kk = int(our_rows_t.shape[1])
    n_local = int(view.shape[0])
    kl = min(kk, n_local)
    loc_rows = torch.full((nq, kk), -1, dtype=torch.int64, device=dev)
    loc_sims = torch.full((nq, kk), -float("inf"), dtype=torch.float32, device=dev)
    own_ref = torch.zeros((nq, kk), dtype=torch.float64, device=dev)   # reference's sim of OUR rows (this rank's)
    own_64 = torch.zeros((nq, kk), dtype=torch.float64, device=dev)    # fp64 re-scoring of OUR rows (this rank's)
    for j in range(nq):
        q = queries[j]
        if n_local:
            ns, s, sim = reference_search(view, q, kl)
            loc_rows[j, :kl] = ns * world + rank
            loc_sims[j, :kl] = s
            mine = (our_rows_t[j] % world == rank) & (our_rows_t[j] >= 0)
            if bool(mine.any()):
                lr = torch.div(our_rows_t[j][mine], world, rounding_mode="floor")
                own_ref[j, mine] = sim[lr].double()
                own_64[j, mine] = fp64_similarity(view, lr, q)
            del sim
    if world > 1:
        all_rows = [torch.empty_like(loc_rows) for _ in range(world)]
        all_sims = [torch.empty_like(loc_sims) for _ in range(world)]
        dist.all_gather(all_rows, loc_rows, group=group)
        dist.all_gather(all_sims, loc_sims, group=group)
        # NaN entries cannot be summed across ranks: mark them, sum, restore
        for t in (own_ref, own_64):
            nan = torch.isnan(t)
            t[nan] = 0.0
            flag = nan.double()
            dist.all_reduce(t, group=group)
            dist.all_reduce(flag, group=group)
            t[flag > 0] = float("nan")
        cat_rows = torch.cat(all_rows, dim=1).cpu().numpy()
        cat_sims = torch.cat(all_sims, dim=1).cpu().numpy()
    else:
        cat_rows, cat_sims = loc_rows.cpu().numpy(), loc_sims.cpu().numpy()
    own_ref, own_64 = own_ref.cpu().numpy(), own_64.cpu().numpy()
    ours_r, ours_s = our_rows_t.cpu().numpy(), our_sims_t.cpu().numpy()
    rep = ParityReport(what)
    for j in range(nq):
        valid = cat_rows[j] >= 0
        r, s = cat_rows[j][valid], cat_sims[j][valid]
        # the reference's order on the concatenated database: similarity descending (NaN first); ties are
        # unspecified by torch.argsort and are covered by the tie rule, lower row first here
        key = np.where(np.isnan(s), np.inf, s.astype(np.float64))
        order = np.lexsort((r, -key))[:kk]
        have = ours_r[j] >= 0
        if int(have.sum()) != len(order):
            rep.queries += 1
            rep.non_tie += 1
            rep.first = rep.first or "query %d: %d results, reference has %d" % (j, int(have.sum()), len(order))
            continue
        rep.add(r[order], s[order], own_ref[j][have], ours_r[j][have], ours_s[j][have], own_64[j][have])
    return rep
