#!/usr/bin/env python
"""Sweep the single-query kernel's ring geometry on one GPU (device-resident timing, CUDA events).

    python tools/k1_sweep.py [--rows 100000 --dim 512 --k 1]
Prints one line per (stage_bytes, stages, ctas): ms per search, GB/s, fraction of the measured peak.
"""
import argparse
import ctypes
import itertools
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from cslam_b200 import NearestNeighborsMatching, _native  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--rows", type=int, default=100000)
ap.add_argument("--dim", type=int, default=512)
ap.add_argument("--k", type=int, default=1)
ap.add_argument("--iters", type=int, default=200)
ap.add_argument("--stage-bytes", default="4096,8192,16384")
ap.add_argument("--stages", default="2,3,4,6")
ap.add_argument("--ctas", default="148,296")
args = ap.parse_args()
try:
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    peak = 6650.0
lib = _native.load()
x = torch.rand((args.rows, args.dim), device="cuda")
x /= x.norm(dim=1, keepdim=True)
q = torch.rand((1, args.dim), device="cuda")
ld = (args.dim + 3) // 4 * 4
nbytes = args.rows * ld * 4
best = None
for sb, st, ct in itertools.product([int(v) for v in args.stage_bytes.split(",")],
                                    [int(v) for v in args.stages.split(",")],
                                    [int(v) for v in args.ctas.split(",")]):
    os.environ["CSLAM_K1_STAGE_BYTES"], os.environ["CSLAM_K1_STAGES"], os.environ["CSLAM_K1_CTAS"] = \
        str(sb), str(st), str(ct)
    m = NearestNeighborsMatching()
    m.add_items(x, list(range(8)) + [0] * (args.rows - 8)) if False else None
    m.dim = args.dim
    h = m._open()
    m._append(h, x)
    m.n = args.rows
    ms = ctypes.c_float()
    rc = lib.cslam_nns_time_search(h.h, ctypes.c_void_p(q.data_ptr()), 1, args.dim, args.k, args.iters,
                                   ctypes.c_void_p(torch.cuda.current_stream().cuda_stream), ctypes.byref(ms))
    if rc != 0:
        print("stage_bytes=%d stages=%d ctas=%d -> error %s" % (sb, st, ct, lib.cslam_nns_last_error(h.h)))
        continue
    gbs = nbytes / (ms.value / 1e3) / 1e9
    print("stage_bytes=%6d stages=%d ctas=%3d  %.4f ms  %7.1f GB/s  %.3f of peak" % (sb, st, ct, ms.value, gbs, gbs / peak), flush=True)
    if best is None or gbs > best[0]:
        best = (gbs, sb, st, ct)
    del m, h
print("best:", best)
