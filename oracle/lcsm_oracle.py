"""CPU ORACLE — test infrastructure only.  Nothing in ``cslam_b200/`` may import this module.

numpy restatement of the CALLERS of the nearest-neighbour search,
``/root/reference/cslam/loop_closure_sparse_matching.py:37-97``, on top
of ``oracle.nns_oracle.search``.  Parity is PINNED: ``tests/golden/make_golden.py`` ran the unmodified
reference ``LoopClosureSparseMatching`` (matcher device forced to CPU) on ``synth.swarm_inputs()`` and on
the config-1 inputs and committed its candidate edges, intra-robot matches and similarity-array lengths
(``tests/golden/swarm_lcsm.npz``, ``config1_lcsm.npz``); ``tests/test_oracle.py`` checks this file
against them.
"""
import numpy as np

from . import nns_oracle


def filter_best(kf, similarity, threshold):
    """:49-52 and :69-71 — the best hit is a match iff ``similarity >= threshold``."""
    return kf is not None and bool(similarity >= threshold)


def filter_intra(kfs, similarities, kf_id, min_gap, threshold):
    """:78-97 — returns ``(match or None, similarities as the reference returns them)``."""
    kfs, similarities = list(kfs), np.asarray(similarities)
    if len(kfs) > 0 and kfs[0] == kf_id:                 # :83-84
        kfs, similarities = kfs[1:], similarities[1:]
    if len(kfs) == 0:                                    # :85-86
        return None, similarities
    for kf, similarity in zip(kfs, similarities):        # :88
        if abs(kf - kf_id) < min_gap:                    # :89-91
            continue
        if similarity < threshold:                       # :93-94
            continue
        return kf, similarities                          # :96
    return None, similarities                            # :97


class OracleMatcher(object):
    """nns_matching.py:5-78 on numpy (rows kept as a python list of fp32 vectors)."""

    def __init__(self):
        self.rows, self.items, self.n = [], {}, 0

    def add_item(self, vector, item):                    # :23-42
        self.rows.append(np.asarray(vector, dtype=np.float32).reshape(-1))
        self.items[self.n] = item
        self.n += 1

    def search(self, query, k):                          # :44-63
        if self.n == 0:
            return [], []
        idx, sims = nns_oracle.search(np.stack(self.rows), np.asarray(query, np.float32), k)
        return [self.items[int(i)] for i in idx], sims

    def search_best(self, query):                        # :65-78
        if self.n == 0:
            return None, None
        items, sims = self.search(query, 1)
        return items[0], sims[0]


class OracleLCSM(object):
    """loop_closure_sparse_matching.py:12-97 with candidate edges recorded like
    algebraic_connectivity_maximization.py:558-571 (key -> best weight)."""

    def __init__(self, params):
        self.params = params
        self.local_nnsm = OracleMatcher()
        self.other_robots_nnsm = {i: OracleMatcher() for i in range(params['max_nb_robots'])
                                  if i != params['robot_id']}
        self.candidate_edges = {}

    def _add_match(self, key, weight):
        if key not in self.candidate_edges or weight > self.candidate_edges[key]:
            self.candidate_edges[key] = weight

    def add_local_global_descriptor(self, embedding, keyframe_id):          # :37-56
        q = np.asarray(embedding).astype(np.float32)
        self.local_nnsm.add_item(q, keyframe_id)
        matches = []
        for i in range(self.params['max_nb_robots']):
            if i != self.params['robot_id']:
                kf, similarity = self.other_robots_nnsm[i].search_best(q)
                if filter_best(kf, similarity, self.params['frontend.similarity_threshold']):
                    key = (self.params['robot_id'], keyframe_id, i, kf)
                    self._add_match(key, float(similarity))
                    matches.append(key)
        return matches

    def add_other_robot_global_descriptor(self, robot_id, keyframe_id, descriptor):   # :58-76
        q = np.asarray(descriptor).astype(np.float32)
        self.other_robots_nnsm[robot_id].add_item(q, keyframe_id)
        kf, similarity = self.local_nnsm.search_best(q)
        if filter_best(kf, similarity, self.params['frontend.similarity_threshold']):
            key = (self.params['robot_id'], kf, robot_id, keyframe_id)
            self._add_match(key, float(similarity))
            return key
        return None

    def match_local_loop_closures(self, descriptor, kf_id):                 # :78-97
        q = np.asarray(descriptor).astype(np.float32)
        kfs, sims = self.local_nnsm.search(q, self.params['frontend.nb_best_matches'])
        return filter_intra(kfs, sims, kf_id,
                            self.params['frontend.intra_loop_min_inbetween_keyframes'],
                            self.params['frontend.similarity_threshold'])
