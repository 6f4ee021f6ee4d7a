"""Row-sharded matcher: one process per GPU, one exchange step.

The reference has no multi-GPU path for the matcher; this is the additive scaling layer named by
BASELINE.json's north_star.  Global row ``i`` lives on rank ``i % world`` as local row ``i // world``
(striped, so incremental ``add_item`` stays balanced and nothing is ever re-balanced).  A search is

    local top-k on every rank  ->  exchange of B*k packed 64-bit keys per rank  ->  merge

On GPUs the exchange and the merge are ONE kernel (K5, ``csrc/k5_exchange.cu``): every rank stores its keys
into every rank's exchange buffer over NVLink peer memory (opened once with CUDA IPC), publishes them with
release stores and merges what it received — no NCCL call on the search path.  The NCCL route (one
``all_gather_into_tensor`` + merge kernel K4) remains for k > 32 or more than 4096 queries per call, and
when ``CSLAM_NNS_EXCHANGE=nccl`` is set.

The keys carry the global row index and a total order (similarity desc, row asc), so the merged
result is bit-identical to a single-GPU search of the concatenated database for any world size.
``torch.distributed`` is the plumbing (NCCL over NVLink on GPUs; the host-side routing logic is
exercised with gloo in the CPU tests through ``plan_rows`` / ``merge_keys_host``).
"""
import ctypes
import os
import warnings

import numpy as np
import torch
import torch.distributed as dist

from . import _native
from .nns_matching import NearestNeighborsMatching


def plan_rows(first_global_row, count, rank, world):
    """Which of the ``count`` rows starting at global row ``first_global_row`` belong to ``rank``.

    Returns (positions within the block, their local row numbers)."""
    g = np.arange(first_global_row, first_global_row + count, dtype=np.int64)
    mine = np.nonzero(g % world == rank)[0]
    return mine, g[mine] // world


def merge_keys_host(keys, k):
    """Host restatement of the K4 merge on packed keys (used by the gloo tests): keys is an integer
    array [parts, B, k] of uint64 keys; returns (global_rows[B, kk], sims[B, kk])."""
    keys = np.asarray(keys, dtype=np.uint64)
    parts, B, kin = keys.shape
    flat = np.transpose(keys, (1, 0, 2)).reshape(B, parts * kin)
    order = np.argsort(flat, axis=1, kind="stable")[:, ::-1][:, :k]
    top = np.take_along_axis(flat, order, axis=1)
    valid = top != 0
    rows = (~top & np.uint64(0xFFFFFFFF)).astype(np.int64)
    ordk = (top >> np.uint64(32)).astype(np.uint32)
    neg = (ordk & np.uint32(0x80000000)) == 0
    bits = np.where(neg, ~ordk, ordk & np.uint32(0x7FFFFFFF)).astype(np.uint32)
    sims = bits.view(np.float32)
    rows = np.where(valid, rows, -1)
    sims = np.where(valid, sims, -np.inf).astype(np.float32)
    return rows, sims


class RowIdentity(object):
    """``items`` of a bulk-loaded shard set (``add_local_shard``): the item of global row ``i`` is ``i``.
    A mapping without storage, so that a 10M-row benchmark database does not need a 10M-entry dict."""

    def __getitem__(self, i):
        return i

    def __len__(self):
        return 0


class ShardedNearestNeighborsMatching(object):
    """Same three methods as the reference matcher; every rank calls them with the same arguments."""

    def __init__(self, dim=None, group=None, exchange=None):
        """exchange: "peer" (K5 over NVLink peer memory, the default on GPUs) or "nccl" (all-gather + K4);
        None reads CSLAM_NNS_EXCHANGE."""
        self.group = group
        self.exchange = exchange or os.environ.get("CSLAM_NNS_EXCHANGE", "peer")
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.local = NearestNeighborsMatching(dim)
        self.n = 0            # global row count
        self.items = dict()   # global row -> item (kept on every rank; tiny)
        self.dim = dim
        self._xchg = None     # peer-memory exchange (K5), opened lazily on the first device search
        self._xchg_tried = False
        self._q_stage = None  # device staging of the query block for search_batch (slice upload + all-gather)

    # -- peer-memory exchange ----------------------------------------------------------------------
    def _exchange(self, device):
        """Open the K5 exchange once: allocate this rank's buffers, all_gather the CUDA IPC handles, map
        the peers.  Collective: every rank takes the same branch (the outcome is agreed by an all_reduce)."""
        if self._xchg_tried:
            return self._xchg
        self._xchg_tried = True
        if self.world == 1 or self.exchange == "nccl" or device.type != "cuda":
            return None
        lib = _native.load()
        x = ctypes.c_void_p()
        rc = lib.cslam_nns_xchg_create(device.index, self.world, self.rank, ctypes.byref(x))
        mine = torch.zeros(128, dtype=torch.uint8)
        if rc == _native.OK:
            rc = lib.cslam_nns_xchg_export(x, ctypes.c_void_p(mine.data_ptr()))
        every = torch.empty((self.world, 128), dtype=torch.uint8, device=device)
        dist.all_gather_into_tensor(every, mine.to(device), group=self.group)
        every = every.cpu().contiguous()
        if rc == _native.OK:
            rc = lib.cslam_nns_xchg_connect(x, ctypes.c_void_p(every.data_ptr()))
        ok = torch.tensor([1 if rc == _native.OK else 0], dtype=torch.int32, device=device)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=self.group)
        if int(ok.item()) == 1:
            self._xchg = x
        else:
            why = lib.cslam_nns_xchg_last_error(x if x.value else None).decode("utf-8", "replace")
            warnings.warn("peer-memory exchange unavailable (%s); using the NCCL all-gather route" % why)
            if x.value:
                lib.cslam_nns_xchg_destroy(x)
        return self._xchg

    def __del__(self):
        try:
            if self._xchg is not None:
                _native.load().cslam_nns_xchg_destroy(self._xchg)
                self._xchg = None
        except Exception:
            pass

    def add_item(self, vector, item):
        if self.dim is None:
            self.dim = len(vector)
        if self.n % self.world == self.rank:
            self.local.add_item(vector, self.n)
        self.items[self.n] = item
        self.n += 1

    def add_items(self, matrix, items):
        if self.dim is None:
            self.dim = int(matrix.shape[1])
        mine, _local = plan_rows(self.n, len(items), self.rank, self.world)
        if len(mine):
            sel = matrix[torch.as_tensor(mine, device=matrix.device)] if torch.is_tensor(matrix) \
                else np.asarray(matrix)[mine]
            self.local.add_items(sel, [self.n + int(j) for j in mine])
        for j, item in enumerate(items):
            self.items[self.n + j] = item
        self.n += len(items)

    def add_local_shard(self, matrix_dev):
        """Bulk-load rows that are ALREADY this rank's shard (benchmarks build shards in place).
        Local row j is global row j*world + rank.  No per-row python bookkeeping: results of the
        device-level search are global row numbers."""
        if self.local.dim is None:
            self.local.dim = int(matrix_dev.shape[1])
            self.dim = self.local.dim
        h = self.local._open()
        self.local._append(h, matrix_dev.to(torch.float32))
        self.local.n += int(matrix_dev.shape[0])

    def finalize_bulk(self):
        """After add_local_shard on every rank: agree on the global row count."""
        t = torch.tensor([self.local.n], dtype=torch.int64, device="cuda")
        if self.world > 1:
            dist.all_reduce(t, group=self.group)
        self.n = int(t.item())
        if not self.items:
            self.items = RowIdentity()

    def _check_exchange(self):
        """After a host synchronisation: a K5 wait that timed out (dead or out-of-step peer) left -1 / -inf
        in the affected queries; report it for THIS call instead of on the next one."""
        if self._xchg is not None:
            lib = _native.load()
            if lib.cslam_nns_xchg_status(self._xchg) != _native.OK:
                why = lib.cslam_nns_xchg_last_error(self._xchg).decode("utf-8", "replace")
                raise _native.NativeError(_native.ECUDA, "peer-memory exchange failed: %s" % why)

    # -- search ------------------------------------------------------------------------------------
    def search_keys_device(self, q_dev, k):
        """Local top-k as packed keys (device int64 tensor [B, k]); stream-ordered, no sync."""
        if self.local.dim is None and self.dim is not None:
            self.local.dim = self.dim  # this rank owns no row yet
        h = self.local._open() if self.local.dim is not None else None
        B = int(q_dev.shape[0])
        keys = torch.zeros((B, k), dtype=torch.int64, device=q_dev.device)
        if h is None or self.local.n == 0:
            return keys
        rc = h.lib.cslam_nns_search_keys(h.h, ctypes.c_void_p(q_dev.data_ptr()), B, q_dev.shape[1], 1,
                                         k, self.world, self.rank, ctypes.c_void_p(keys.data_ptr()),
                                         self.local._stream())
        _native.check(rc, h.h)
        return keys

    def search_batch_device(self, q_dev, k, out_idx=None, out_sim=None, exchange=None):
        """Device in, device out: (global rows int64 [B,k], sims f32 [B,k]); -1/-inf past the end.
        Stream-ordered, no host synchronisation; callers that synchronise themselves should call
        ``check_exchange()`` afterwards.  ``exchange="nccl"`` forces the all-gather + K4 route for this call."""
        q_dev = q_dev.contiguous()
        B = int(q_dev.shape[0])
        if self.world == 1 and self.local._h is not None and self.local.n >= k:
            # one shard: nothing to exchange or merge — the kernels write (row, sim) themselves
            if out_idx is None:
                out_idx = torch.empty((B, k), dtype=torch.int64, device=q_dev.device)
                out_sim = torch.empty((B, k), dtype=torch.float32, device=q_dev.device)
            h = self.local._h
            rc = h.lib.cslam_nns_search(h.h, ctypes.c_void_p(q_dev.data_ptr()), B, q_dev.shape[1], 1, k,
                                        ctypes.c_void_p(out_idx.data_ptr()), ctypes.c_void_p(out_sim.data_ptr()),
                                        1, self.local._stream(), None)
            _native.check(rc, h.h)
            return out_idx, out_sim
        if out_idx is None:
            out_idx = torch.empty((B, k), dtype=torch.int64, device=q_dev.device)
            out_sim = torch.empty((B, k), dtype=torch.float32, device=q_dev.device)
        xchg = self._exchange(q_dev.device) if (k <= 32 and B <= 4096 and B * k <= 131072
                                                and exchange != "nccl") else None
        if xchg is not None:
            # local search + exchange + merge in ONE C-ABI call (cslam_nns_search_sharded)
            if self.local.dim is None and self.dim is not None:
                self.local.dim = self.dim  # this rank owns no row yet
            h = self.local._open()
            rc = h.lib.cslam_nns_search_sharded(h.h, xchg, q_dev.data_ptr(), B, q_dev.shape[1], 1, k, self.rank,
                                                self.world, 3, out_idx.data_ptr(), out_sim.data_ptr(),
                                                self.local._stream())
            _native.check(rc, h.h)
            return out_idx, out_sim
        keys = self.search_keys_device(q_dev, k)
        if self.world > 1:
            gathered = torch.empty((self.world, B, k), dtype=torch.int64, device=q_dev.device)
            dist.all_gather_into_tensor(gathered, keys, group=self.group)
        else:
            gathered = keys.view(1, B, k)
        lib = _native.load()
        rc = lib.cslam_nns_merge_keys(q_dev.device.index, ctypes.c_void_p(gathered.data_ptr()),
                                      self.world, B, k, ctypes.c_void_p(out_idx.data_ptr()),
                                      ctypes.c_void_p(out_sim.data_ptr()), 1,
                                      ctypes.c_void_p(torch.cuda.current_stream().cuda_stream), None)
        _native.check(rc, None)
        return out_idx, out_sim

    def plan_search_device(self, q_dev, k, out_idx, out_sim):
        """A device-in / device-out search of fixed buffers as a zero-argument callable: the argument conversion
        of ``search_batch_device`` (contiguity check, pointers, shapes) is done once, every call is one C-ABI call
        on torch's current stream.  For loops that repeat a search of the same buffers faster than Python can
        rebuild the arguments (a 32 us single-query kernel).  The tensors must stay alive and unchanged in
        shape; their contents may change between calls."""
        if not (self.world == 1 and self.local._h is not None and self.local.n >= k and q_dev.is_contiguous()
                and out_idx.is_contiguous() and out_sim.is_contiguous()):
            return lambda: self.search_batch_device(q_dev, k, out_idx, out_sim)
        h = self.local._h
        fn, hh = h.lib.cslam_nns_search, h.h
        qp, oi, os_ = q_dev.data_ptr(), out_idx.data_ptr(), out_sim.data_ptr()
        B, ld, dev_index = int(q_dev.shape[0]), int(q_dev.shape[1]), self.local._dev_index
        raw_stream = torch._C._cuda_getCurrentRawStream if hasattr(torch._C, "_cuda_getCurrentRawStream") else \
            (lambda i: torch.cuda.current_stream(i).cuda_stream)

        def call():
            rc = fn(hh, qp, B, ld, 1, k, oi, os_, 1, raw_stream(dev_index), None)
            if rc:
                _native.check(rc, hh)
        return call

    def check_exchange(self):
        """Raise if a peer-memory exchange of an earlier ``search_batch_device`` call timed out (call after
        synchronising the stream)."""
        self._check_exchange()

    def _upload_queries(self, q_host):
        """Host query block (identical on every rank) -> device.  With several ranks each one uploads only ITS
        1/world slice over its own PCIe link and one all-gather over NVLink completes the block on every GPU
        (world identical 4 MB uploads cost more than the kernel they feed at 8 GPUs)."""
        B, dim = int(q_host.shape[0]), int(q_host.shape[1])
        dev = torch.device("cuda", torch.cuda.current_device())
        if self.world == 1 or not dist.is_initialized() or B < 4 * self.world:
            return q_host.to(dev, non_blocking=True)
        per = (B + self.world - 1) // self.world
        if self._q_stage is None or self._q_stage.shape[0] != self.world * per or self._q_stage.shape[1] != dim:
            self._q_stage = torch.empty((self.world * per, dim), dtype=torch.float32, device=dev)
        lo, hi = min(self.rank * per, B), min((self.rank + 1) * per, B)
        mine = self._q_stage[self.rank * per:(self.rank + 1) * per]
        if hi > lo:
            mine[:hi - lo].copy_(q_host[lo:hi], non_blocking=True)
        dist.all_gather_into_tensor(self._q_stage, mine, group=self.group)
        return self._q_stage[:B]

    def search_batch(self, queries, k, root_only=False):
        """Host queries in, host results out.  ``root_only=True``: only rank 0 reads the results back
        (every rank holds identical results on the device; the others return ``(None, None)``)."""
        if not torch.is_tensor(queries):
            queries = torch.from_numpy(np.asarray(queries).astype(np.float32))
        k = int(k)
        kk = min(k, self.n)
        B = int(queries.reshape(-1, self.dim).shape[0]) if self.dim else 0
        if kk == 0 or B == 0:
            return [[] for _ in range(B)], np.empty((B, 0), np.float32)
        q = queries.reshape(-1, self.dim).to(torch.float32)
        q = q if q.is_cuda else self._upload_queries(q)
        idx, sims = self.search_batch_device(q, kk)
        if root_only and self.rank != 0:
            torch.cuda.current_stream().synchronize()
            self._check_exchange()
            return None, None
        idx, sims = idx.cpu().numpy(), sims.cpu().numpy()
        self._check_exchange()
        return self._items_of(idx), sims

    def _items_of(self, idx):
        """(B, kk) global rows -> list of B item lists (the identity mapping needs no lookups)."""
        if isinstance(self.items, RowIdentity):
            return idx.tolist()
        return [[self.items[i] for i in row] for row in idx.tolist()]

    def search_batches(self, batches, k, root_only=False, depth=2):
        """A stream of host query blocks -> a generator of ``(items, sims)``, one per block, in order: the
        pipelined form of ``search_batch`` for callers that have the next block ready while the current one is
        searched (message batches queued by the subscriber, a serving loop).  Every block still travels host ->
        device and its results device -> host; block i+1's upload (own stream, own copy engine) and block i-1's
        result read run while block i is searched, and up to ``depth`` blocks are in flight.  With several ranks
        every rank passes the same blocks and uploads each whole block over its own PCIe link (no collective and
        no SM use on the upload path: it has to run next to a kernel that occupies every SM).
        ``root_only=True``: ranks other than 0 yield ``(None, None)`` and read nothing back."""
        from .pipeline import pipelined_search
        reads = not (root_only and self.rank != 0)
        return pipelined_search(self, batches, int(k), depth, reads, self.search_batch_device, self._items_of,
                                self._check_exchange)

    def search(self, query, k):
        if self.n == 0 and self.dim is None:
            return [], []
        items, sims = self.search_batch(query, k)
        return items[0], sims[0]

    def search_best(self, query):
        if self.n == 0 and self.dim is None:
            return None, None
        items, sims = self.search(query, 1)
        return items[0], sims[0]
