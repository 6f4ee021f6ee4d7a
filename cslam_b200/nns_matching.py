"""Drop-in replacement for ``cslam.nns_matching.NearestNeighborsMatching``.

Same constructor, methods, return types and public attributes as the reference class
(``/root/reference/cslam/nns_matching.py:5-78``); underneath, the descriptor database lives in a
growable HBM arena owned by ``libcslam_nns.so`` and every search is one hand-written sm_100a kernel
(see ``include/cslam_nns.h`` and ``DESIGN.md``).  PyTorch is used only for the current CUDA stream,
for exposing ``.data`` as a tensor view and for converting inputs exactly like the reference does.

Additive API (does not change any reference signature): ``add_items``, ``search_batch`` and
``search`` accepting a ``(B, dim)`` query; ``search_best_multi`` (one query against several databases,
one synchronisation) and ``search_candidates`` (search + the caller's threshold / self-hit /
min-gap filter on the device).
"""
import ctypes
import operator

import numpy as np
import torch

from . import _native


def _raw_stream(device_index):
    """Handle of torch's current CUDA stream on ``device_index`` as an int (the fast private getter that
    torch's own extensions use, with the public API as the fallback)."""
    try:
        return torch._C._cuda_getCurrentRawStream(device_index)
    except AttributeError:  # pragma: no cover - older/newer torch without the private getter
        return torch.cuda.current_stream(device_index).cuda_stream


def _addr(a):
    """Address of a numpy array's first element (cheaper than ``a.ctypes.data``)."""
    return a.__array_interface__['data'][0]


class _CudaArrayView:
    """Minimal ``__cuda_array_interface__`` carrier so torch can alias the arena without a copy."""

    def __init__(self, ptr, shape, strides, owner):
        self.__cuda_array_interface__ = {
            "shape": tuple(shape),
            "typestr": "<f4",
            "data": (int(ptr), False),
            "strides": tuple(strides),
            "version": 3,
        }
        self._owner = owner  # keeps the native handle alive as long as the view is


class _Handle:
    """Owns the native handle; destroyed when the last Python reference goes away."""

    def __init__(self, device_index, dim):
        self.lib = _native.load()
        self.h = ctypes.c_void_p()
        rc = self.lib.cslam_nns_create(int(device_index), int(dim or 0), 0, ctypes.byref(self.h))
        _native.check(rc, None)

    def __del__(self):
        try:
            if getattr(self, "h", None) is not None and self.h.value:
                self.lib.cslam_nns_destroy(self.h)
                self.h = ctypes.c_void_p()
        except Exception:  # interpreter shutdown
            pass


class NearestNeighborsMatching(object):
    """Nearest Neighbor matching of description vectors (reference: nns_matching.py:5)."""

    def __init__(self, dim=None):
        """Reference: nns_matching.py:9-21."""
        self.n = 0
        self.device = torch.device('cuda')
        self.dim = dim
        self.items = dict()
        self._h = None
        self._view = None
        self._view_key = None
        # integer identities of the rows, for the device-side candidate filter (search_candidates):
        # items that are ints are queued here and handed to the library lazily, in one call
        self._ids_pending = []
        self._ids_first = 0        # row of _ids_pending[0]
        self._ids_ok = True        # False once a non-integer item was added
        self._multi_cache = None   # search_best_multi: (ids of the databases, handle array, result block)
        if dim is not None:
            self._open()

    # -- native handle ---------------------------------------------------------------------------
    def _device_index(self):
        dev = torch.device(self.device)
        if dev.type != 'cuda':
            raise RuntimeError(
                "cslam_b200.NearestNeighborsMatching runs on a CUDA device only (device=%s); "
                "there is no CPU path" % (dev,))
        return dev.index if dev.index is not None else torch.cuda.current_device()

    def _open(self):
        if self._h is None:
            self._dev_index = self._device_index()
            self._h = _Handle(self._dev_index, self.dim)
            # persistent result block of the single-query fast path (k <= 128): the library writes here,
            # the caller gets copies
            self._res_idx = np.empty(128, dtype=np.int64)
            self._res_sim = np.empty(128, dtype=np.float32)
            self._res_ptr = (_addr(self._res_idx), _addr(self._res_sim))
            self._qid_buf = np.zeros(1, dtype=np.int64)
            self._pos_buf = np.zeros(1, dtype=np.int32)
            self._qid_ptr, self._pos_ptr = _addr(self._qid_buf), _addr(self._pos_buf)
        return self._h

    def _stream(self):
        return ctypes.c_void_p(_raw_stream(self._dev_index if self._h is not None else self._device_index()))

    # -- public attribute `data` (reference: nns_matching.py:19-21, read by tests :36,:49,:66) -----
    @property
    def data(self):
        if self._h is None:
            return []  # reference: empty python list until the first allocation
        lib, h = self._h.lib, self._h.h
        cap, ld, ptr = lib.cslam_nns_capacity(h), lib.cslam_nns_ld(h), lib.cslam_nns_data_ptr(h)
        key = (cap, ptr)
        if self._view_key != key:
            carrier = _CudaArrayView(ptr, (cap, int(self.dim)), (ld * 4, 4), self._h)
            self._view = torch.as_tensor(carrier, device=torch.device('cuda', lib.cslam_nns_device(h)))
            self._view_key = key
        return self._view

    @data.setter
    def data(self, value):
        """The reference's ``data`` is a plain attribute; assigning a ``(rows, dim)`` tensor/array here copies it
        into the HBM arena (capacity grows to hold it, ``n`` is left alone exactly as a rebinding would)."""
        if self._h is None and isinstance(value, (list, tuple)) and len(value) == 0:
            return
        value = torch.as_tensor(value)
        if value.dim() != 2:
            raise ValueError("data must be a (rows, dim) matrix")
        if self.dim is None:
            self.dim = int(value.shape[1])
        if int(value.shape[1]) != self.dim:
            raise RuntimeError("data has %d columns, the matcher's dim is %d" % (value.shape[1], self.dim))
        h = self._open()
        _native.check(h.lib.cslam_nns_reserve(h.h, int(value.shape[0])), h.h)
        view = self.data
        view[:value.shape[0]].copy_(value.to(torch.float32))
        if view.shape[0] > value.shape[0]:
            view[value.shape[0]:].zero_()
        torch.cuda.current_stream(view.device).synchronize()

    # -- add_item (reference: nns_matching.py:23-42) ------------------------------------------------
    def add_item(self, vector, item):
        """Add item to the matching list.

        Args:
            vector (np.array): descriptor
            item: identification info (e.g., int)
        """
        if not torch.is_tensor(vector):
            vector = torch.from_numpy(vector.astype(np.float32))
        if self.dim is None:
            self.dim = len(vector)
        h = self._open()
        if (not vector.is_cuda and vector.dtype == torch.float32 and vector.is_contiguous()
                and vector.numel() == self.dim):
            # the production call (one CPU fp32 descriptor, reference :45,:64): straight to the C-ABI
            rc = h.lib.cslam_nns_add(h.h, vector.data_ptr(), 1, self.dim, 0, _raw_stream(self._dev_index), None)
            if rc:
                _native.check(rc, h.h)
        else:
            row = vector.reshape(1, self.dim)  # RuntimeError on a wrong length, like the reference (:41)
            row = row.to(torch.float32)
            self._append(h, row)
        self.items[self.n] = item
        if self._ids_ok:
            if type(item) is int and -0x8000000000000000 <= item <= 0x7fffffffffffffff:
                self._ids_pending.append(item)
            else:
                self._note_ids((item,))
        self.n += 1

    def add_items(self, matrix, items):
        """Additive API: append ``len(items)`` rows at once (one DMA instead of one per row)."""
        if not torch.is_tensor(matrix):
            matrix = torch.from_numpy(np.asarray(matrix).astype(np.float32))
        if matrix.dim() != 2:
            raise ValueError("matrix must be (count, dim)")
        if self.dim is None:
            self.dim = int(matrix.shape[1])
        if matrix.shape[1] != self.dim or matrix.shape[0] != len(items):
            raise RuntimeError("shape mismatch: matrix %s, dim %s, %d items" %
                               (tuple(matrix.shape), self.dim, len(items)))
        h = self._open()
        self._append(h, matrix.to(torch.float32))
        for j, item in enumerate(items):
            self.items[self.n + j] = item
        self._note_ids(items)
        self.n += len(items)

    # -- wire / on-disk format (additive) ------------------------------------------------
    def add_descriptor_msgs(self, msgs):
        """Ingest one ``GlobalDescriptors`` message: each element carries ``descriptor`` (the float32[]
        wire field, reference global_descriptor_loop_closure_detection.py:158-162,211-214) and
        ``keyframe_id``.  The descriptors are packed once into pinned host memory and appended with ONE
        DMA, instead of one list->array->tensor->H2D conversion per descriptor (reference
        loop_closure_sparse_matching.py:64-65)."""
        msgs = list(msgs)
        if not msgs:
            return
        dim = len(msgs[0].descriptor) if self.dim is None else self.dim
        # pinned staging: cslam_nns_add waits for the copy of a pinned source before it returns (see the
        # contract in include/cslam_nns.h), so the block can go back to torch's pinned cache right after
        staging = torch.empty((len(msgs), dim), dtype=torch.float32,
                              pin_memory=torch.cuda.is_available())
        block = staging.numpy()
        for j, m in enumerate(msgs):
            block[j] = np.asarray(m.descriptor, dtype=np.float32)   # raises on a wrong length
        self.add_items(staging, [m.keyframe_id for m in msgs])

    def save(self, path, allow_pickle=False):
        """Write the database (rows 0..n-1 exactly as stored, and the row -> item map) to ``path`` (.npz).
        The reference has no persistence; this is what a restart needs.  Integer items (keyframe
        ids, the production case) are stored as a plain int64 array; arbitrary Python items need
        ``allow_pickle=True`` on both ``save`` and ``load`` (a pickled payload runs code when loaded)."""
        rows = self.data[:self.n].cpu().numpy() if self.n else np.zeros((0, self.dim or 0), np.float32)
        items = [self.items[i] for i in range(self.n)]
        if all(type(it) is int and -2 ** 63 <= it < 2 ** 63 for it in items):
            np.savez(path, data=rows, dim=np.int64(self.dim or 0), item_ids=np.asarray(items, dtype=np.int64))
            return
        if not allow_pickle:
            raise TypeError("items are not all Python ints: pass allow_pickle=True to store them pickled "
                            "(only load such a file from a source you trust)")
        import pickle
        np.savez(path, data=rows, dim=np.int64(self.dim or 0),
                 items=np.frombuffer(pickle.dumps(items), dtype=np.uint8))

    @classmethod
    def load(cls, path, allow_pickle=False):
        """Rebuild a matcher from ``save``'s file: one bulk append, identical rows and items.  A file whose
        items were pickled is refused unless ``allow_pickle=True``."""
        z = np.load(path if str(path).endswith(".npz") else str(path) + ".npz", allow_pickle=False)
        m = cls()
        if "item_ids" in z.files:
            items = z["item_ids"].tolist()
        elif not allow_pickle:
            raise ValueError("%s stores pickled items: pass allow_pickle=True if the file is trusted" % (path,))
        else:
            import pickle
            items = pickle.loads(z["items"].tobytes())
        if len(items):
            m.add_items(z["data"], items)
        elif int(z["dim"]) > 0:
            m.dim = int(z["dim"])
        return m

    def _append(self, h, rows):
        rows = rows.contiguous()
        on_dev = rows.is_cuda
        if on_dev and rows.device.index != h.lib.cslam_nns_device(h.h):
            rows = rows.cpu()
            on_dev = False
        rc = h.lib.cslam_nns_add(h.h, rows.data_ptr(), rows.shape[0], rows.shape[1],
                                 1 if on_dev else 0, _raw_stream(self._dev_index), None)
        if rc:
            _native.check(rc, h.h)

    # -- search (reference: nns_matching.py:44-63) --------------------------------------------------
    def search(self, query, k):
        """Search for nearest neighbors.

        Args:
            query (np.array): descriptor to match
            k (int): number of best matches to return

        Returns:
            list(int, np.array): best matches
        """
        if self._h is None:  # reference: len(self.data) == 0 -> [], []
            return [], []
        # Fast path for the production call (one host descriptor, k <= 128): same conversions as the
        # reference (:57-60), no torch objects in between, results land in a persistent block.
        dim = self.dim
        if not torch.is_tensor(query):
            q = query.astype(np.float32)                       # reference :57-58
            if q.ndim == 2 and not (q.shape[0] == 1 or dim == 1):
                return self.search_batch(torch.from_numpy(q), k)
        keep, ptr = self._host_query(query if torch.is_tensor(query) else q)
        k = int(k)
        if keep is not None and 0 <= k and min(k, self.n) <= 128:
            kk = min(k, self.n)
            if kk == 0:
                return [], np.empty(0, dtype=np.float32)
            h = self._h
            rc = h.lib.cslam_nns_search(h.h, ptr, 1, dim, 0, kk, self._res_ptr[0], self._res_ptr[1], 0,
                                        _raw_stream(self._dev_index), None)
            if rc:
                _native.check(rc, h.h)
            if kk == 1:
                return [self.items[int(self._res_idx[0])]], self._res_sim[:1].copy()
            return [self.items[i] for i in self._res_idx[:kk].tolist()], self._res_sim[:kk].copy()
        if not torch.is_tensor(query):
            query = torch.from_numpy(q)
        if query.dim() == 2 and not (query.shape[0] == 1 or self.dim == 1):
            return self.search_batch(query, k)
        out_dtype = np.float64 if query.dtype == torch.float64 else np.float32  # torch promotion
        q = query.reshape(1, self.dim).to(torch.float32)
        idx, sims = self._search_block(q, k)
        items = [self.items[i] for i in idx[0].tolist()]
        return items, sims[0].astype(out_dtype, copy=False)

    def _host_query(self, query):
        """``(keepalive, address)`` of ONE host fp32 query of ``dim`` contiguous values, or ``(None, 0)`` when
        the query needs the general conversion path (device tensor, other dtype, strided).  numpy input must
        already be float32 (the caller did ``astype``)."""
        dim = self.dim
        if not torch.is_tensor(query):
            if query.size != dim:                               # reference :60 (torch reshape) raises the same
                raise RuntimeError("shape '[1, %d]' is invalid for input of size %d" % (dim, query.size))
            if not query.flags.c_contiguous:
                query = np.ascontiguousarray(query)
            return query, _addr(query)
        if (not query.is_cuda and query.dtype == torch.float32 and query.is_contiguous()
                and query.numel() == dim and query.dim() <= 2):
            return query, query.data_ptr()
        return None, 0

    def search_batch(self, queries, k):
        """Additive API: ``B`` queries at once -> ``(list of B item lists, sims[B, kk])``."""
        if self._h is None:
            return [], []
        if not torch.is_tensor(queries):
            queries = torch.from_numpy(np.asarray(queries).astype(np.float32))
        q = queries.reshape(-1, self.dim).to(torch.float32)
        idx, sims = self._search_block(q, k)
        return self._items_of(idx), sims

    def _items_of(self, idx):
        """(B, kk) rows -> list of B item lists: one C-level lookup of all B * kk rows (a Python-level loop over
        2560 dict lookups costs more than the search of a small shard)."""
        B, kk = idx.shape
        flat = idx.ravel().tolist()
        if len(flat) == 0:
            return [[] for _ in range(B)]
        if len(flat) == 1:
            return [[self.items[flat[0]]]]
        got = operator.itemgetter(*flat)(self.items)
        return [list(got[i * kk:(i + 1) * kk]) for i in range(B)]

    def search_batches(self, batches, k, depth=2):
        """Additive API: a stream of host query blocks -> a generator of ``(items, sims)`` per block, in order,
        with block i+1's upload and block i-1's result read overlapping block i's search (``pipeline.py``)."""
        from .pipeline import pipelined_search
        if self._h is None:
            return iter([([], []) for _ in batches])

        def search_device(dq, kk, out_idx, out_sim):
            h = self._h
            rc = h.lib.cslam_nns_search(h.h, dq.data_ptr(), int(dq.shape[0]), int(dq.shape[1]), 1, kk,
                                        out_idx.data_ptr(), out_sim.data_ptr(), 1, _raw_stream(self._dev_index), None)
            if rc:
                _native.check(rc, h.h)

        def items_of(rows):
            return self._items_of(rows)

        return pipelined_search(self, batches, int(k), depth, True, search_device, items_of, lambda: None)

    def _search_block(self, q, k):
        """q: (B, dim) fp32 torch tensor, host or device -> (idx[B,kk] int64, sims[B,kk] f32)."""
        h = self._h
        k = int(k)
        if k < 0:  # python slice semantics of the reference's [:k] for a negative k
            k = max(self.n + k, 0)
        kk = min(k, self.n)
        B = int(q.shape[0])
        idx = np.empty((B, kk), dtype=np.int64)
        sims = np.empty((B, kk), dtype=np.float32)
        if kk == 0 or B == 0:
            return idx, sims
        q = q.contiguous()
        on_dev = q.is_cuda
        if on_dev and q.device.index != h.lib.cslam_nns_device(h.h):
            q = q.cpu()
            on_dev = False
        rc = h.lib.cslam_nns_search(h.h, q.data_ptr(), B, q.shape[1], 1 if on_dev else 0, kk,
                                    _addr(idx), _addr(sims), 0, _raw_stream(self._dev_index), None)
        if rc:
            _native.check(rc, h.h)
        return idx, sims

    # -- device-side candidate filter (additive) ---------------------------------------
    def _note_ids(self, items):
        if not self._ids_ok:
            return
        for it in items:
            if isinstance(it, (int, np.integer)) and not isinstance(it, (bool, np.bool_)) \
                    and -2 ** 63 <= int(it) < 2 ** 63:
                self._ids_pending.append(int(it))
            else:  # arbitrary Python objects stay host-only; the filter is then the caller's job
                self._ids_ok = False
                self._ids_pending = []
                return

    @property
    def ids_on_device(self):
        """True while every item added so far is an int, i.e. ``search_candidates`` can filter by id."""
        return self._ids_ok and self._h is not None

    def _flush_ids(self):
        if self._ids_pending:
            ids = np.asarray(self._ids_pending, dtype=np.int64)
            rc = self._h.lib.cslam_nns_set_row_ids(self._h.h, self._ids_first, len(ids),
                                                   ctypes.c_void_p(ids.ctypes.data))
            _native.check(rc, self._h.h)
            self._ids_first += len(ids)
            self._ids_pending = []

    def search_candidates(self, query, k, query_id=None, min_gap=0, threshold=-np.inf,
                          mode=_native.FILTER_INTRA, batched=None):
        """Top-``k`` search plus the candidate filter of the reference's callers, evaluated on the device.

        ``mode=FILTER_INTRA`` (``loop_closure_sparse_matching.py:83-97``): a leading hit whose id equals
        ``query_id`` is dropped, candidates with ``|id - query_id| < min_gap`` or ``sim < threshold`` are
        skipped, the first survivor is the match.  ``mode=FILTER_BEST`` (``:51,:71``): the best hit is the
        match iff ``sim >= threshold``.  ``query`` may be ``(dim,)`` or ``(B, dim)`` (then ``query_id`` is a
        sequence of B ids or None); ``batched=True`` forces the batch form of the result for B == 1.

        Returns ``(items, sims, pos)``: the UNFILTERED top-k exactly as ``search`` / ``search_batch`` return
        it, and the position of the accepted candidate in it (-1 = none; an int, or an int32 array of B).
        """
        if self._h is None:
            raise RuntimeError("search_candidates on an empty matcher")
        if not torch.is_tensor(query):
            query = torch.from_numpy(np.asarray(query).astype(np.float32))
        single = not (query.dim() == 2 and not (query.shape[0] == 1 or self.dim == 1))
        if batched is not None:
            single = not batched
        if single and 0 < int(k) and min(int(k), self.n) <= 128:
            # the per-keyframe call (one host descriptor): no torch objects, persistent result block
            keep, ptr = self._host_query(query)
            if keep is not None:
                kk = min(int(k), self.n)
                qid = None
                if query_id is not None:
                    if not self._ids_ok:
                        raise TypeError("search_candidates with query_id needs integer items (keyframe ids)")
                    self._flush_ids()
                    qid = self._qid_buf
                    qid[0] = query_id
                h = self._h
                rc = h.lib.cslam_nns_search_filtered(
                    h.h, ptr, 1, self.dim, 0, kk, int(mode), self._qid_ptr if qid is not None else None,
                    int(min_gap), float(np.float32(threshold)), self._res_ptr[0], self._res_ptr[1],
                    self._pos_ptr, _raw_stream(self._dev_index), None)
                if rc:
                    _native.check(rc, h.h)
                return ([self.items[i] for i in self._res_idx[:kk].tolist()], self._res_sim[:kk].copy(),
                        int(self._pos_buf[0]))
        q = (query.reshape(1, self.dim) if single else query.reshape(-1, self.dim))
        q = q.to(torch.float32).contiguous()
        B = int(q.shape[0])
        h = self._h
        on_dev = q.is_cuda
        if on_dev and q.device.index != h.lib.cslam_nns_device(h.h):
            q, on_dev = q.cpu(), False
        k = int(k)
        if k < 0:
            k = max(self.n + k, 0)
        kk = min(k, self.n)
        qids = None
        if query_id is not None:
            if not self._ids_ok:
                raise TypeError("search_candidates with query_id needs integer items (keyframe ids)")
            self._flush_ids()
            qids = np.ascontiguousarray(np.asarray(query_id, dtype=np.int64).reshape(-1))
            if len(qids) != B:
                raise ValueError("need one query_id per query")
        idx = np.empty((B, max(kk, 1)), dtype=np.int64)
        sims = np.empty((B, max(kk, 1)), dtype=np.float32)
        pos = np.full((B,), -1, dtype=np.int32)
        if kk > 0 and B > 0:
            thr = np.float32(threshold)   # numpy compares an np.float32 similarity with a python float in fp32
            rc = h.lib.cslam_nns_search_filtered(
                h.h, ctypes.c_void_p(q.data_ptr()), B, q.shape[1], 1 if on_dev else 0, kk, int(mode),
                ctypes.c_void_p(qids.ctypes.data) if qids is not None else None, int(min_gap),
                ctypes.c_float(float(thr)), ctypes.c_void_p(idx.ctypes.data),
                ctypes.c_void_p(sims.ctypes.data), ctypes.c_void_p(pos.ctypes.data), self._stream(), None)
            _native.check(rc, h.h)
        idx, sims = idx[:, :kk], sims[:, :kk]
        if single:
            return [self.items[i] for i in idx[0].tolist()], sims[0], int(pos[0])
        return [[self.items[i] for i in row] for row in idx.tolist()], sims, pos

    # -- one query, several databases (additive) -----------------------------------------
    @staticmethod
    def search_best_multi(matchers, query):
        """``[m.search_best(query) for m in matchers]`` with the query staged once, the kernels of all
        databases queued back to back and ONE host synchronisation (reference: the per-robot loop of
        ``loop_closure_sparse_matching.py:47-55``).  Empty matchers give ``(None, None)``."""
        matchers = list(matchers)
        live = [m for m in matchers if m._h is not None and m.n > 0]
        if not live:                        # keep the reference's quirks for empty-but-dimensioned DBs
            return [m.search_best(query) for m in matchers]
        m0 = live[0]
        lib = m0._h.lib
        if not torch.is_tensor(query):
            query = query.astype(np.float32)                    # reference :57-58
        keep, ptr = m0._host_query(query)
        on_dev = 0
        if keep is None:
            if not torch.is_tensor(query):
                query = torch.from_numpy(query)
            keep = query.reshape(1, m0.dim).to(torch.float32).contiguous()
            if keep.is_cuda and keep.device.index != lib.cslam_nns_device(m0._h.h):
                keep = keep.cpu()
            on_dev, ptr = (1 if keep.is_cuda else 0), keep.data_ptr()
        key = tuple(map(id, live))
        cache = m0._multi_cache
        if cache is None or cache[0] != key:   # the handle array and result block of this set of databases
            idx = np.empty((len(live),), dtype=np.int64)
            sims = np.empty((len(live),), dtype=np.float32)
            cache = m0._multi_cache = (key, (ctypes.c_void_p * len(live))(*[m._h.h for m in live]),
                                       idx, sims, _addr(idx), _addr(sims))
        _, hs, idx, sims, p_idx, p_sims = cache
        rc = lib.cslam_nns_search_multi(hs, len(live), ptr, m0.dim, on_dev, 1, p_idx, p_sims, None,
                                        _raw_stream(m0._dev_index))
        if rc:
            _native.check(rc, m0._h.h)
        out, j = [], 0
        for m in matchers:
            if j < len(live) and live[j] is m:
                out.append((m.items[int(idx[j])], sims[j]))   # sims[j]: a fresh np.float32 scalar
                j += 1
            else:
                out.append(m.search_best(query))
        return out

    # -- search_best (reference: nns_matching.py:65-78) ----------------------------------------------
    def search_best(self, query):
        """Search for the nearest neighbor.

        Returns:
            int, np.array: best match
        """
        if self._h is None:
            return None, None
        items, similarities = self.search(query, 1)
        return items[0], similarities[0]
