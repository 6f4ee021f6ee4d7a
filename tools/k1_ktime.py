"""Back-to-back kernel time of the single-query search for several k (development tool, B200 only)."""
import ctypes
import sys

import torch

sys.path.insert(0, ".")
from cslam_b200 import NearestNeighborsMatching  # noqa: E402

for n, d in ((100_000, 512), (250_000, 256), (2_000_000, 256)):
    m = NearestNeighborsMatching()
    x = torch.rand((n, d), device="cuda")
    x /= x.norm(dim=1, keepdim=True)
    m.add_items(x, list(range(n)))
    del x
    q = torch.rand((1, d), device="cuda")
    lib, h = m._h.lib, m._h.h
    out = []
    for k in (1, 2, 10, 32, 33, 128):
        ms = ctypes.c_float()
        rc = lib.cslam_nns_time_search(h, q.data_ptr(), 1, d, k, 200,
                                       ctypes.c_void_p(torch.cuda.current_stream().cuda_stream), ctypes.byref(ms))
        assert rc == 0
        out.append("k=%d %.1f us" % (k, ms.value * 1e3))
    print("%d x %d: %s" % (n, d, "  ".join(out)), flush=True)
    del m
