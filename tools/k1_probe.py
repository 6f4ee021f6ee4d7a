#!/usr/bin/env python
"""Small fixed workload for ncu: a few single-query searches on a 100k x 512 DB (config 2 shape)."""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from cslam_b200 import NearestNeighborsMatching  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--rows", type=int, default=100000)
ap.add_argument("--dim", type=int, default=512)
ap.add_argument("--k", type=int, default=1)
ap.add_argument("--batch", type=int, default=1)
ap.add_argument("--iters", type=int, default=6)
a = ap.parse_args()
x = torch.rand((a.rows, a.dim), device="cuda")
x /= x.norm(dim=1, keepdim=True)
m = NearestNeighborsMatching()
m.dim = a.dim
h = m._open()
m._append(h, x)
m.n = a.rows
q = torch.rand((a.batch, a.dim))
for _ in range(a.iters):
    idx, sims = m._search_block(q, a.k)
torch.cuda.synchronize()
print("ok", idx[0, :3], sims[0, :3], "variant", h.lib.cslam_nns_last_variant(h.h))
