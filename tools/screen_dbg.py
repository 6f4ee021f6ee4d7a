"""Development: dump the screening pass's partial key lists and look for keys that do not decode to a row."""
import sys, os, ctypes
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'tests'))
import numpy as np, torch, synth
from cslam_b200 import NearestNeighborsMatching
os.environ["CSLAM_NNS_SCREEN_NOFB"] = "1"
os.environ["CSLAM_NNS_K6_TRACE"] = "1"
n, d, B, k = 20000, 256, 256, 10
kl = int(os.environ.get("CSLAM_NNS_SCREEN_LIST", "0")) or max(12, min(32, 2 * k + 4))
db = synth.unit_rows(900 + n, n, d); qs = synth.queries_for(901 + n, db, B)
m = NearestNeighborsMatching(); m.add_items(db, list(range(n)))
lib, h = m._h.lib, m._h.h
lib.cslam_nns_set_screen(h, 1, 0)
idx, sims = m._search_block(torch.from_numpy(qs), k)
a, b = ctypes.c_int64(), ctypes.c_int64()
lib.cslam_nns_screen_stats(h, ctypes.byref(a), ctypes.byref(b))
parts = 2 * min((n + 255) // 256, 74)
keys = np.zeros((parts, B, kl), np.uint64)
rc = lib.cslam_nns_debug_copy(h, 0, keys.ctypes.data, keys.nbytes)
print("rc", rc, "parts", parts, "kl", kl, "screened", a.value, "fallback", b.value)
rows = (~keys & np.uint64(0xFFFFFFFF)).astype(np.int64)
nz = keys != 0
bad = nz & (rows >= n)
print("nonzero keys", int(nz.sum()), "bad-row keys", int(bad.sum()), "per query", bad.sum(axis=(0, 2))[:8])
pp, bb, jj = np.nonzero(bad)
print("bad by part (first 20):", np.bincount(pp, minlength=parts)[:20], "... by j:", np.bincount(jj, minlength=kl))
for t in range(min(8, len(pp))):
    key = int(keys[pp[t], bb[t], jj[t]])
    print("  part %d q %d j %d key %016x row %d" % (pp[t], bb[t], jj[t], key, rows[pp[t], bb[t], jj[t]]))
q0 = keys[:, 0, :].reshape(-1); q0 = q0[q0 != 0]
print("query 0: %d keys, top 12:" % len(q0), ["%016x" % int(x) for x in np.sort(q0)[::-1][:12]])
par = np.zeros(5 * B + 16, np.uint32)
lib.cslam_nns_debug_copy(h, 1, par.ctypes.data, par.nbytes)
print("gthr[0:4]", par[2 * B:2 * B + 4], "ebound[0:4]", par[4 * B:4 * B + 4].view(np.float32), "tail", par[5 * B:5 * B + 9])

dbg = np.zeros((B, 64), np.uint64)
print("dbg rc", lib.cslam_nns_debug_copy(h, 2, dbg.ctypes.data, dbg.nbytes))
f = lambda v: float(np.array([int(v) & 0xffffffff], np.uint32).view(np.float32)[0])
for b_ in (0, 1, 2):
    d_ = dbg[b_]
    print("q%d band:" % b_, ["%016x" % int(x) for x in d_[:8]])
    print("   exact:", ["%016x" % int(x) for x in d_[24:32]])
    print("   n_keys %d nband %d T %08x lim %.6f g %d E %.3g bad %d n %d total %d" % (
        int(d_[48]), int(d_[49]), int(d_[50]), f(d_[51]), int(d_[52]), f(d_[53]), int(d_[54]), int(d_[55]), int(d_[56])))
print("nband histogram:", np.bincount(np.minimum(dbg[:, 49].astype(np.int64), 300) // 10)[:31])
