"""Where does the host-side time of one single-query search go?  (development tool, B200 only)
Times the same 100k x 512 top-1 search through successively thinner layers of the stack."""
import ctypes
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from cslam_b200 import NearestNeighborsMatching  # noqa: E402

n, d, reps = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000, 512, 2000
m = NearestNeighborsMatching()
x = torch.rand((n, d), device="cuda")
x /= x.norm(dim=1, keepdim=True)
m.add_items(x, list(range(n)))
lib, h = m._h.lib, m._h.h
q_np = np.random.default_rng(0).random(d).astype(np.float32)
q_pin = torch.from_numpy(q_np).reshape(1, d).pin_memory()
q_page = torch.from_numpy(q_np).reshape(1, d)
q_dev = q_pin.cuda()
idx = np.empty((1, 1), np.int64)
sim = np.empty((1, 1), np.float32)
idx_d = torch.empty((1, 1), dtype=torch.int64, device="cuda")
sim_d = torch.empty((1, 1), dtype=torch.float32, device="cuda")
stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def raw(qt, on_dev, dst_dev):
    oi = ctypes.c_void_p(idx_d.data_ptr() if dst_dev else idx.ctypes.data)
    os_ = ctypes.c_void_p(sim_d.data_ptr() if dst_dev else sim.ctypes.data)
    lib.cslam_nns_search(h, ctypes.c_void_p(qt.data_ptr()), 1, d, on_dev, 1, oi, os_, dst_dev, stream, None)


def timeit(name, fn, sync_each=False):
    for _ in range(50):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
        if sync_each:
            torch.cuda.synchronize()
    torch.cuda.synchronize()
    print("%-58s %7.2f us" % (name, (time.perf_counter() - t0) / reps * 1e6), flush=True)


timeit("search_best(numpy query)                [public API]", lambda: m.search_best(q_np))
timeit("search_best(cpu torch query)", lambda: m.search_best(q_page[0]))
timeit("_search_block(pinned host q) -> host", lambda: m._search_block(q_pin, 1))
timeit("_search_block(pageable host q) -> host", lambda: m._search_block(q_page, 1))
timeit("_search_block(device q) -> host", lambda: m._search_block(q_dev, 1))
timeit("C-ABI: pinned host q -> host results", lambda: raw(q_pin, 0, 0))
timeit("C-ABI: pageable host q -> host results", lambda: raw(q_page, 0, 0))
timeit("C-ABI: device q -> host results", lambda: raw(q_dev, 1, 0))
timeit("C-ABI: device q -> device results + sync each", lambda: raw(q_dev, 1, 1), sync_each=True)
timeit("C-ABI: device q -> device results, no sync (back to back)", lambda: raw(q_dev, 1, 1))
timeit("C-ABI: pinned host q -> device results, no sync", lambda: raw(q_pin, 0, 1))
timeit("C-ABI: pageable host q -> device results, no sync", lambda: raw(q_page, 0, 1))
