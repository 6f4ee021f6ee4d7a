"""Drop-in ``ScanContextMatching`` (reference: ``/root/reference/cslam/lidar_pr/scancontext_matching.py``).

Same constructor, same three methods, same return shapes and quirks as the reference class; the two
array-heavy steps of ``search`` run on the GPU through ``include/cslam_scancontext.h``:

* step 1 (reference :58-63) builds a scipy KD-tree over ALL ring keys on every search and queries it for
  ``num_candidates`` neighbours; here the ring keys live in HBM and one kernel brute-forces the Euclidean
  shortlist (nearest first, ties to the lower row);
* step 2 (reference :68-75 with ``scancontext_utils.distance_sc`` :78-113) is a Python triple loop of
  ``num_candidates x sectors x sectors`` column cosines; here one CTA per candidate computes the column-cosine
  table once and sums it per shift in the reference's order.

What stays on the host, exactly as in the reference: the ring key ``np.mean(sc, axis=1)``
(``scancontext_utils.sc2rk`` :75-76, so the stored keys are bit-identical), the selection loop over the
shortlist (first strictly smaller distance wins, :68-75), the ``items`` dict and the return values.
There is no CPU fallback: without the CUDA library / a B200 the first ``add_item`` raises.
"""
import ctypes

import numpy as np
import torch

from .. import _native


class _ScHandle(object):
    def __init__(self, device_index, shape, num_candidates):
        self.lib = _native.load()
        self.h = ctypes.c_void_p()
        rc = self.lib.cslam_sc_create(device_index, int(shape[0]), int(shape[1]), int(num_candidates), 0,
                                      ctypes.byref(self.h))
        if rc != 0:
            raise _native.NativeError(rc, (self.lib.cslam_sc_last_error(None) or b"").decode())

    def check(self, rc):
        if rc != 0:
            raise _native.NativeError(rc, (self.lib.cslam_sc_last_error(self.h) or b"").decode())

    def __del__(self):
        try:
            if self.h:
                self.lib.cslam_sc_destroy(self.h)
                self.h = None
        except Exception:
            pass


def sc2rk(sc):
    """Ring key, ``scancontext_utils.sc2rk`` (:75-76)."""
    return np.mean(sc, axis=1)


class ScanContextMatching(object):
    """Nearest Neighbor matching of description vectors (reference :6-104)."""

    def __init__(self, shape=[20, 60], num_candidates=10, threshold=0.15):
        self.shape = shape
        self.num_candidates = num_candidates
        self.threshold = threshold
        self.items = dict()
        self.nb_items = 0
        self._h = None  # device storage is created at the first add (the reference allocates numpy zeros)

    def _open(self):
        if self._h is None:
            dev = torch.cuda.current_device() if torch.cuda.is_available() else 0
            self._h = _ScHandle(dev, self.shape, self.num_candidates)
        return self._h

    @staticmethod
    def _stream():
        return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream if torch.cuda.is_available() else 0)

    def add_item(self, descriptor, item):
        """Reference :24-45."""
        # the same two host operations as the reference (:31,:39), then both are widened to the float64 the
        # reference's storage arrays have (:18-19,:41-42).  A torch tensor (what LoopClosureSparseMatching
        # hands over, loop_closure_sparse_matching.py:45) makes the reference's np.mean raise TypeError; it is
        # accepted here as the numpy array it wraps.
        descriptor = np.asarray(descriptor)
        sc = descriptor.reshape(self.shape)
        rk = np.ascontiguousarray(np.asarray(sc2rk(sc), dtype=np.float64))
        sc = np.ascontiguousarray(np.asarray(sc, dtype=np.float64))
        h = self._open()
        row = ctypes.c_int64()
        h.check(h.lib.cslam_sc_add(h.h, ctypes.c_void_p(sc.ctypes.data), ctypes.c_void_p(rk.ctypes.data),
                                   self._stream(), ctypes.byref(row)))
        self.items[self.nb_items] = item
        self.nb_items = self.nb_items + 1

    def _candidates(self, query):
        """(rows, dists, yaw shifts) of the ring-key shortlist, in shortlist order."""
        query_sc = np.asarray(query).reshape(self.shape)                                 # :61
        rk = np.ascontiguousarray(np.asarray(sc2rk(query_sc), dtype=np.float64))        # :62
        query_sc = np.ascontiguousarray(np.asarray(query_sc, dtype=np.float64))
        h = self._open()
        c = self.num_candidates
        rows = np.empty(c, np.int64)
        dists = np.empty(c, np.float64)
        yaws = np.empty(c, np.int32)
        h.check(h.lib.cslam_sc_search(h.h, ctypes.c_void_p(query_sc.ctypes.data), ctypes.c_void_p(rk.ctypes.data),
                                      ctypes.c_void_p(rows.ctypes.data), ctypes.c_void_p(dists.ctypes.data),
                                      ctypes.c_void_p(yaws.ctypes.data), self._stream()))
        return rows, dists, yaws

    def search(self, query, k):
        """Reference :47-87 (``k`` is accepted and ignored there as well: one match is returned)."""
        if self.nb_items < 1:
            return [None], [None]
        rows, dists, yaws = self._candidates(query)
        nn_dist = 1.0  # initialize with the largest value of distance (:66)
        nn_idx = None
        nn_yawdiff = None
        for ith in range(self.num_candidates):
            if rows[ith] < 0:
                continue  # fewer rows than candidates: the reference compares an all-zero row, dist 1 (:70-72)
            if dists[ith] < nn_dist:
                nn_dist = dists[ith]
                nn_yawdiff = yaws[ith]
                nn_idx = int(rows[ith])
        if nn_idx is None:
            nn_idx = 0
            similarity = 0.0
        else:
            similarity = 1 - nn_dist
        self.last_yaw_diff_deg = 0 if nn_yawdiff is None else nn_yawdiff * (360 / self.shape[1])
        return [self.items[nn_idx]], [similarity]

    def search_best(self, query):
        """Reference :89-104."""
        if self.nb_items < 1:
            return None, None
        idxs, sims = self.search(query, 1)
        return idxs[0], sims[0]
