"""Development: where the re-scoring kernel (K6) spends its time.  Runs one screened batch with the K6 trace on
(CSLAM_NNS_K6_TRACE=1: %globaltimer stamps of thread 0 of every query's CTA) and prints the phase durations.

    python tools/k6_phase_probe.py [rows] [dim]
"""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ["CSLAM_NNS_K6_TRACE"] = "1"
import numpy as np
import torch

from cslam_b200 import NearestNeighborsMatching

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 250_000
dim = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
B, k = 256, 10
dev = torch.device("cuda", 0)
gen = torch.Generator(device=dev)
gen.manual_seed(1)
m = NearestNeighborsMatching(dim=dim)
h = m._open()
done = 0
while done < rows:
    c = min(rows - done, (1 << 30) // (dim * 4))
    x = torch.rand((c, dim), generator=gen, device=dev)
    x /= x.norm(dim=1, keepdim=True)
    m._append(h, x)
    m.n += c
    done += c
q = torch.rand((B, dim), generator=gen, device=dev)
lib, hh = m._h.lib, m._h.h
for it in range(3):
    idx, sims = m._search_block(q, k)
dbg = np.zeros((B, 64), np.uint64)
print("debug_copy rc", lib.cslam_nns_debug_copy(hh, 2, dbg.ctypes.data, dbg.nbytes))
a, b = ctypes.c_int64(), ctypes.c_int64()
lib.cslam_nns_screen_stats(hh, ctypes.byref(a), ctypes.byref(b))
print("rows %d dim %d screened %d fallback %d" % (rows, dim, a.value, b.value))
ts = dbg[:, 59:64].astype(np.int64)
t0 = ts[:, 0].min()
names = ["load union", "radix select", "re-score >= T (+ select)", "band re-score"]
for j in range(4):
    d = (ts[:, j + 1] - ts[:, j]) / 1e3
    print("%-28s median %.1f us  p90 %.1f  max %.1f" % (names[j], np.median(d), np.percentile(d, 90), d.max()))
print("CTA start (from first): median %.1f us max %.1f;  CTA end: median %.1f us max %.1f" % (
    np.median(ts[:, 0] - t0) / 1e3, (ts[:, 0] - t0).max() / 1e3, np.median(ts[:, 4] - t0) / 1e3,
    (ts[:, 4] - t0).max() / 1e3))
print("ntop median %d, band total median %d max %d" % (np.median(dbg[:, 57]), np.median(dbg[:, 49]), dbg[:, 49].max()))
