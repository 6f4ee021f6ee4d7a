"""Timeline of the single-query kernel K1 (development tool, B200 only).  Needs the instrumented build:
    make -C cslam_b200/csrc K1_PROF=1 OUT=../_lib/libcslam_nns_prof.so
Prints, for the LAST of a few back-to-back searches, the per-CTA %globaltimer stamps relative to the first
CTA's start: CTA start, first bytes landed, end of the streaming loop (warp 0), ticket drawn."""
import ctypes
import os
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from cslam_b200 import _native  # noqa: E402

_native.LIB_PATH = os.path.join(os.path.dirname(_native.LIB_PATH), "libcslam_nns_prof.so")
from cslam_b200 import NearestNeighborsMatching  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000
d = int(sys.argv[2]) if len(sys.argv) > 2 else 512
m = NearestNeighborsMatching()
x = torch.rand((n, d), device="cuda")
x /= x.norm(dim=1, keepdim=True)
m.add_items(x, list(range(n)))
lib = m._h.lib
lib.cslam_k1_prof_read.argtypes = [ctypes.c_void_p, ctypes.c_int]
q = torch.rand((1, d), device="cuda")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for sync_each in (True, False):
    for _ in range(20):
        m._search_block(q, 1)
    torch.cuda.synchronize()
    idx_d = torch.empty((1, 1), dtype=torch.int64, device="cuda")
    sim_d = torch.empty((1, 1), dtype=torch.float32, device="cuda")
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    e0.record()
    for _ in range(10):
        lib.cslam_nns_search(m._h.h, q.data_ptr(), 1, d, 1, 1, idx_d.data_ptr(), sim_d.data_ptr(), 1, st, None)
        if sync_each:
            torch.cuda.synchronize()
    e1.record()
    torch.cuda.synchronize()
    buf = np.zeros(8 * 148, dtype=np.uint64)
    assert lib.cslam_k1_prof_read(buf.ctypes.data, 148) == 0
    t = buf.reshape(148, 8).astype(np.int64)
    t0 = t[:, 0].min()
    r = (t - t0) / 1e3
    print("n=%d d=%d %s: events %.2f us per search" % (n, d, "sync after each" if sync_each else "back to back",
                                                      e0.elapsed_time(e1) * 100))
    for j, name in ((0, "CTA start"), (7, "reached the issue block"), (4, "ring primed (TMA issued)"), (6, "plain ld.global.cv returned"),
                    (5, "query in registers"), (1, "first bytes landed (TMA)"),
                    (2, "stream loop done (warp 0)"), (3, "ticket drawn")):
        c = r[:, j]
        print("  %-28s min %7.2f  median %7.2f  max %7.2f us" % (name, c.min(), np.median(c), c.max()))
    print("  loop-done per CTA (us, CTA 0..147):", " ".join("%.1f" % v for v in r[:, 2]))
    print("  first-byte latency (per CTA)   median %.2f us;  loop duration median %.2f us;  tail median %.2f us"
          % (np.median(r[:, 1] - r[:, 0]), np.median(r[:, 2] - r[:, 1]), np.median(r[:, 3] - r[:, 2])))
