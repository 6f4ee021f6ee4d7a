"""Host-side mirror of the reference caller of the hot path,
``cslam.loop_closure_sparse_matching.LoopClosureSparseMatching``
(``/root/reference/cslam/loop_closure_sparse_matching.py:6-115``), wired to the B200 matcher.

Only the three methods that drive ``NearestNeighborsMatching`` are mirrored;
candidate *selection* (``AlgebraicConnectivityMaximization`` + MAC) is out of scope and
is delegated to the reference's own class when the ``cslam`` package is importable.  When it is not
(e.g. on the benchmark box), matches are recorded by ``CandidateEdges``, which keeps exactly the
``candidate_edges`` bookkeeping of ``algebraic_connectivity_maximization.py:558-571`` so results can be
compared edge by edge.

Additive: ``add_other_robot_global_descriptors`` handles one whole
``GlobalDescriptors`` message (<= 10 descriptors, ``global_descriptor_loop_closure_detection.py:388-404``)
with ONE batched search instead of one search per descriptor, and ``add_local_global_descriptor``
searches the R-1 other robots' databases through ``search_best_multi`` (one staging, one synchronisation).
Additive: the threshold / self-hit / min-gap filter of ``:51,71,83-97`` runs on the device
(``NearestNeighborsMatching.search_candidates``) whenever the keyframe ids are ints; results are identical.
"""
from typing import NamedTuple

import numpy as np
import torch

from ._native import FILTER_BEST
from .nns_matching import NearestNeighborsMatching

try:  # the real selector, when this package is dropped into a cslam installation
    from cslam.algebraic_connectivity_maximization import (  # type: ignore
        AlgebraicConnectivityMaximization, EdgeInterRobot)
    _HAVE_REFERENCE_SELECTOR = True
except Exception:  # pragma: no cover - depends on the environment
    _HAVE_REFERENCE_SELECTOR = False

    class EdgeInterRobot(NamedTuple):
        """Inter-robot loop closure edge (algebraic_connectivity_maximization.py:8-15)."""
        robot0_id: int
        robot0_keyframe_id: int
        robot1_id: int
        robot1_keyframe_id: int
        weight: float


class CandidateEdges(object):
    """Records candidate edges exactly like ``add_match`` (…maximization.py:558-571) does."""

    def __init__(self, robot_id=0, max_nb_robots=1):
        self.robot_id = robot_id
        self.max_nb_robots = max_nb_robots
        self.candidate_edges = {}

    def add_match(self, match):
        key = (match.robot0_id, match.robot0_keyframe_id, match.robot1_id, match.robot1_keyframe_id)
        if key in self.candidate_edges:
            if match.weight > self.candidate_edges[key].weight:
                self.candidate_edges[key] = match
        else:
            self.candidate_edges[key] = match

    def select_candidates(self, *_args, **_kwargs):
        raise NotImplementedError(
            "candidate selection (AlgebraicConnectivityMaximization/MAC) is outside the hot path; "
            "install the reference `cslam` package to get it")


class LoopClosureSparseMatching(object):
    """Sparse matching for loop closure detection (reference :6-115, stereo/rgbd sensor types)."""

    def __init__(self, params, node=None, candidate_selector=None):
        self.params = params
        self.node = node
        lidar = self.params.get("frontend.sensor_type", "stereo") == "lidar"
        if lidar:  # reference :22-23,29-30: the ScanContext matcher implements the same three methods
            from .lidar_pr import ScanContextMatching
        make = ScanContextMatching if lidar else NearestNeighborsMatching
        self._dense = not lidar   # the dense matcher has the batched / multi-database / filtered entries
        self.local_nnsm = make()
        self.other_robots_nnsm = {}
        for i in range(self.params['max_nb_robots']):
            if i != self.params['robot_id']:
                self.other_robots_nnsm[i] = make()
        if candidate_selector is not None:
            self.candidate_selector = candidate_selector
        elif _HAVE_REFERENCE_SELECTOR:
            self.candidate_selector = AlgebraicConnectivityMaximization(
                self.params['robot_id'], self.params['max_nb_robots'])
        else:
            self.candidate_selector = CandidateEdges(self.params['robot_id'],
                                                     self.params['max_nb_robots'])

    def _log(self, text):
        if self.node is not None:
            self.node.get_logger().info(text)

    def add_local_global_descriptor(self, embedding, keyframe_id):
        """Reference :37-56 — 1 add on the local DB + (R-1) search_best on the other robots' DBs."""
        matches = []
        tensor = torch.from_numpy(embedding.astype(np.float32))
        self.local_nnsm.add_item(tensor, keyframe_id)
        robots = [i for i in range(self.params['max_nb_robots']) if i != self.params['robot_id']]
        if self._dense and len(robots) > 1:
            # f1: the R-1 databases are searched with the query staged once and ONE host synchronisation
            best = NearestNeighborsMatching.search_best_multi(
                [self.other_robots_nnsm[i] for i in robots], tensor)
        else:
            best = [self.other_robots_nnsm[i].search_best(tensor) for i in robots]
        for i, (kf, similarity) in zip(robots, best):
            if kf is not None:
                if similarity >= self.params['frontend.similarity_threshold']:
                    match = EdgeInterRobot(self.params['robot_id'], keyframe_id, i, kf, similarity)
                    self.candidate_selector.add_match(match)
                    matches.append(match)
        return matches

    def add_other_robot_global_descriptor(self, msg):
        """Reference :58-76 — 1 add on that robot's DB + 1 search_best on the local DB."""
        tensor = torch.from_numpy(np.asarray(msg.descriptor).astype(np.float32))
        self.other_robots_nnsm[msg.robot_id].add_item(tensor, msg.keyframe_id)
        match = None
        kf, similarity = self.local_nnsm.search_best(tensor)
        if kf is not None:
            if similarity >= self.params['frontend.similarity_threshold']:
                self._log(f"Found potential matching KF: ({kf},{msg.keyframe_id}): {similarity}")
                match = EdgeInterRobot(self.params['robot_id'], kf, msg.robot_id, msg.keyframe_id,
                                       similarity)
                self.candidate_selector.add_match(match)
        return match

    def add_other_robot_global_descriptors(self, msgs):
        """Additive (f1): same results as calling the method above for each message in order — the
        adds go to the other robots' DBs and the searches to the local DB, so they commute — but the
        local DB is scanned ONCE for the whole message."""
        msgs = list(msgs)
        if not msgs:
            return []
        if not hasattr(self.local_nnsm, "search_batch"):  # lidar: the ScanContext matcher is single-query
            return [self.add_other_robot_global_descriptor(m) for m in msgs]
        block = np.stack([np.asarray(m.descriptor).astype(np.float32) for m in msgs])
        by_robot = {}
        for j, m in enumerate(msgs):
            by_robot.setdefault(m.robot_id, []).append(j)
        for rid, rows in by_robot.items():
            self.other_robots_nnsm[rid].add_items(torch.from_numpy(block[rows]),
                                                  [msgs[j].keyframe_id for j in rows])
        if self.local_nnsm.n == 0:
            return [None] * len(msgs)
        # f2: the `similarity >= threshold` test of :71 runs on the device (FILTER_BEST); pos == 0 <=> passed
        items, sims, pos = self.local_nnsm.search_candidates(
            torch.from_numpy(block), 1, mode=FILTER_BEST, batched=True,
            threshold=self.params['frontend.similarity_threshold'])
        out = []
        for m, its, s, p in zip(msgs, items, sims, pos):
            match = None
            if p == 0:
                match = EdgeInterRobot(self.params['robot_id'], its[0], m.robot_id, m.keyframe_id, s[0])
                self.candidate_selector.add_match(match)
            out.append(match)
        return out

    def match_local_loop_closures(self, descriptor, kf_id):
        """Reference :78-97 — top `frontend.nb_best_matches` on the local DB, then host filtering."""
        tensor = torch.from_numpy(np.asarray(descriptor).astype(np.float32))
        nnsm = self.local_nnsm
        if self._dense and nnsm.n > 0 and nnsm.ids_on_device and isinstance(kf_id, (int, np.integer)):
            # f2: self-hit drop, min-gap window and threshold are evaluated on the device next to the top-k
            kfs, similarities, pos = nnsm.search_candidates(
                tensor, self.params['frontend.nb_best_matches'], query_id=kf_id,
                min_gap=self.params['frontend.intra_loop_min_inbetween_keyframes'],
                threshold=self.params['frontend.similarity_threshold'])
            match = kfs[pos] if pos >= 0 else None
            if len(kfs) > 0 and kfs[0] == kf_id:
                similarities = similarities[1:]
            return match, similarities
        kfs, similarities = self.local_nnsm.search(tensor, k=self.params['frontend.nb_best_matches'])
        if len(kfs) > 0 and kfs[0] == kf_id:
            kfs, similarities = kfs[1:], similarities[1:]
        if len(kfs) == 0:
            return None, similarities
        for kf, similarity in zip(kfs, similarities):
            if abs(kf - kf_id) < self.params['frontend.intra_loop_min_inbetween_keyframes']:
                continue
            if similarity < self.params['frontend.similarity_threshold']:
                continue
            return kf, similarities
        return None, similarities

    def select_candidates(self, number_of_candidates, is_neighbor_in_range,
                          greedy_initialization=True):
        """Reference :99-115 — delegated to the candidate selector (out of scope here)."""
        return self.candidate_selector.select_candidates(
            int(number_of_candidates), dict(is_neighbor_in_range), greedy_initialization)
