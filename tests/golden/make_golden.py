#!/usr/bin/env python
"""Generate the golden vectors under tests/golden/ by RUNNING THE UNMODIFIED REFERENCE.

Run in the build container only (needs /root/reference, which does not exist on the GPU box):

    python tests/golden/make_golden.py

The reference class hard-codes torch.device('cuda') (cslam/nns_matching.py:16); there is no GPU in
the build container, so each instance is created with dim=None and its `device` attribute is set to
CPU before the first add — the arithmetic (torch cosine_similarity / argsort / gather) is the same
code path on the other device.  LoopClosureSparseMatching needs a `node` with get_logger().info
(cslam/loop_closure_sparse_matching.py:12,72); a stub is passed.
"""
import os
import sys
from collections import namedtuple

import numpy as np
import torch

sys.path.insert(0, "/root/reference")
from cslam.nns_matching import NearestNeighborsMatching as RefNNS  # noqa: E402
import cslam.loop_closure_sparse_matching as ref_lcsm  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import synth  # noqa: E402  (tests/synth.py: deterministic inputs, so only outputs are stored)

CPU = torch.device("cpu")


def ref_matcher():
    m = RefNNS()
    m.device = CPU
    return m


# name, distribution, N, D, k, number of queries, seed
SEARCH_CASES = [
    ("d10_n37_k5", "rand", 37, 10, 5, 8, 11),
    ("d100_n100_k100", "rand", 100, 100, 100, 6, 12),      # tests/test_sparse_matching.py:51-81 shape
    ("d128_n1500_k10", "rand", 1500, 128, 10, 8, 13),      # crosses the 1000-row doubling (:37-38)
    ("d128_n1500_k1_randn", "randn", 1500, 128, 1, 8, 14),
    ("d512_n700_k10", "rand", 700, 512, 10, 6, 15),        # CosPlace-shaped
    ("d256_n2100_k10_randn", "randn", 2100, 256, 10, 6, 16),
    ("d4096_n96_k10", "rand", 96, 4096, 10, 4, 17),        # NetVLAD-shaped rows (16 KB)
    ("d3_n9_k20", "rand", 9, 3, 20, 4, 18),                # k > n
    ("d130_n300_k300", "rand", 300, 130, 300, 3, 19),      # k > 128: full ranking
    ("d96_n3000_k40", "rand", 3000, 96, 40, 4, 31),        # 32 < k <= 128: the register lists of four keys per lane
    ("d64_n2500_k128_randn", "randn", 2500, 64, 128, 4, 32),
]


def gen_search_cases():
    out = {}
    for name, dist, n, d, k, nq, seed in SEARCH_CASES:
        db = synth.unit_rows(seed, n, d, dist)
        qs = synth.queries_for(seed + 1000, db, nq, dist)
        m = ref_matcher()
        for i in range(n):
            m.add_item(db[i], 1000 + i)  # item id != row index on purpose
        idx = np.full((nq, min(k, n)), -1, np.int64)
        sims = np.zeros((nq, min(k, n)), np.float32)
        for j in range(nq):
            items, s = m.search(qs[j], k)
            idx[j] = np.asarray(items) - 1000
            sims[j] = s
        best = [m.search_best(qs[j]) for j in range(nq)]
        assert np.array_equal(m.data[:n].numpy(), db)  # storage is the exact fp32 input (:41)
        out[name + "/spec"] = np.asarray([n, d, k, nq, seed, 0 if dist == "rand" else 1], np.int64)
        out[name + "/idx"] = idx
        out[name + "/sims"] = sims
        out[name + "/best_item"] = np.asarray([b[0] for b in best], np.int64)
        out[name + "/best_sim"] = np.asarray([b[1] for b in best], np.float32)
        out[name + "/capacity"] = np.int64(len(m.data))
    np.savez_compressed(os.path.join(HERE, "search_cases.npz"), **out)
    print("search_cases.npz:", len(SEARCH_CASES), "cases")


def gen_edge_cases():
    """Zero row, duplicate rows, tiny-norm row, un-normalised rows, float64 torch query."""
    rng = np.random.default_rng(21)
    d = 16
    db = rng.random((12, d)).astype(np.float32)
    db[3] = 0.0                      # zero row -> similarity 0
    db[5] = db[2]                    # exact duplicate -> tie
    db[7] = db[2] * 3.0              # same direction, other norm -> tie in cosine
    db[9] = 1e-12                    # norm below eps -> clamped
    m = ref_matcher()
    for i in range(len(db)):
        m.add_item(db[i], i)
    qs = rng.random((5, d)).astype(np.float32)
    qs[1] = db[2]
    qs[2] = 0.0                      # zero query
    sims_all = np.stack([m.search(q, len(db))[1] for q in qs])
    idx_all = np.stack([np.asarray(m.search(q, len(db))[0]) for q in qs])
    q64 = torch.from_numpy(qs[0].astype(np.float64))
    it64, s64 = m.search(q64, 4)
    np.savez_compressed(os.path.join(HERE, "edge_cases.npz"), db=db, queries=qs, idx=idx_all,
                        sims=sims_all, f64_items=np.asarray(it64), f64_sims=s64)
    print("edge_cases.npz: sims dtype for float64 torch query:", s64.dtype)


class _Logger:
    def info(self, *_a, **_k):
        pass


class _Node:
    def get_logger(self):
        return _Logger()


GlobalDescriptor = namedtuple("GlobalDescriptor", ["keyframe_id", "robot_id", "descriptor"])


def config1_inputs(n=1000, d=128):
    local = synth.unit_rows(1, n, d, "rand")
    other = synth.unit_rows(2, n, d, "rand")
    eps = synth.normal(3, (n, d)).astype(np.float32)
    for j in range(0, n, 10):  # true overlaps between the two robots
        v = local[(3 * j) % n] + 0.05 * eps[j]
        other[j] = v / np.linalg.norm(v)
    return local, other


def gen_config1():
    """BASELINE config 1: 2 robots x 1k keyframes x 128-d through LoopClosureSparseMatching."""
    ref_lcsm.NearestNeighborsMatching = lambda *a, **k: ref_matcher()  # CPU-forced reference class
    params = {"robot_id": 0, "max_nb_robots": 2, "frontend.sensor_type": "stereo",
              "frontend.similarity_threshold": 0.0, "frontend.nb_best_matches": 10,
              "frontend.intra_loop_min_inbetween_keyframes": 10,
              "frontend.enable_sparsification": True,
              "evaluation.enable_sparsification_comparison": False}
    n, d = 1000, 128
    local, other = config1_inputs(n, d)
    lcsm = ref_lcsm.LoopClosureSparseMatching(params, _Node())
    intra = []
    for i in range(n):
        lcsm.add_local_global_descriptor(local[i], i)
        kf, sims = lcsm.match_local_loop_closures(local[i], i)
        intra.append(-1 if kf is None else kf)
        lcsm.add_other_robot_global_descriptor(GlobalDescriptor(i, 1, other[i].tolist()))
    edges = lcsm.candidate_selector.candidate_edges
    keys = np.asarray(sorted(edges.keys()), np.int64)
    weights = np.asarray([edges[tuple(k)].weight for k in keys.tolist()], np.float64)
    np.savez_compressed(os.path.join(HERE, "config1_lcsm.npz"), edge_keys=keys, edge_weights=weights, intra=np.asarray(intra, np.int64))
    print("config1_lcsm.npz:", len(keys), "candidate edges;", sum(1 for x in intra if x >= 0),
          "intra matches")


def gen_swarm():
    """4 robots, a real threshold and min-gap window, through the reference's
    LoopClosureSparseMatching: candidate edges, intra matches and the similarity arrays it returns."""
    ref_lcsm.NearestNeighborsMatching = lambda *a, **k: ref_matcher()
    params = {"robot_id": 0, "max_nb_robots": 4, "frontend.sensor_type": "stereo",
              "frontend.similarity_threshold": 0.85, "frontend.nb_best_matches": 10,
              "frontend.intra_loop_min_inbetween_keyframes": 20,
              "frontend.enable_sparsification": True,
              "evaluation.enable_sparsification_comparison": False}
    descs = synth.swarm_inputs()
    n = len(descs[0])
    lcsm = ref_lcsm.LoopClosureSparseMatching(params, _Node())
    intra, nsims, first_sim, local_edges = [], [], [], []
    for i in range(n):
        ms = lcsm.add_local_global_descriptor(descs[0][i], i)
        local_edges.append(len(ms))
        kf, sims = lcsm.match_local_loop_closures(descs[0][i], i)
        intra.append(-1 if kf is None else kf)
        nsims.append(len(sims))
        first_sim.append(sims[0] if len(sims) else np.float32(0))
        for r in range(1, 4):
            lcsm.add_other_robot_global_descriptor(GlobalDescriptor(i, r, descs[r][i].tolist()))
    edges = lcsm.candidate_selector.candidate_edges
    keys = np.asarray(sorted(edges.keys()), np.int64)
    weights = np.asarray([edges[tuple(k)].weight for k in keys.tolist()], np.float64)
    np.savez_compressed(os.path.join(HERE, "swarm_lcsm.npz"), edge_keys=keys, edge_weights=weights,
                        intra=np.asarray(intra, np.int64), nsims=np.asarray(nsims, np.int64),
                        first_sim=np.asarray(first_sim, np.float32),
                        local_edges=np.asarray(local_edges, np.int64))
    print("swarm_lcsm.npz:", len(keys), "candidate edges;", sum(1 for x in intra if x >= 0),
          "intra matches;", sum(local_edges), "edges from local keyframes")


if __name__ == "__main__":
    torch.manual_seed(0)
    if len(sys.argv) > 1 and sys.argv[1] == "swarm":
        gen_swarm()
        sys.exit(0)
    gen_search_cases()
    gen_edge_cases()
    gen_config1()
    gen_swarm()
