#!/usr/bin/env python
"""Headline benchmark: NN-search queries/s on the configuration BASELINE.json's metric is quoted on.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--workload c3|c2|c2_k10|c4|c5|swarm|sc] [--no-legs]

Primary workload ``c3`` = BASELINE.json configs[2]: 1M x 4096 fp32 descriptors, batches of 256 queries,
top-10 (the "1M x 4096 DB" the metric names).  One *step* = one batch of 256 queries searched against the
whole database.  With N > 1 ranks (torchrun) the SAME database is row-striped over the ranks (strong
scaling), every step ends with one exchange of the B*k packed keys fused with the merge, and timing is the
max over ranks of a CUDA-event interval.

Without ``--workload`` the line also carries ``legs``: short runs of the other BASELINE configurations, each
with its own ``roofline``, ``e2e``, ``clocks``, ``parity`` and ``gpu_reference`` —

    N == 1 : c2 (100k x 512, search_best), c2_k10 (same database, search k=10), c4 (2M x 256, add_item between
             searches), swarm (configs[3] as 8 robots through LoopClosureSparseMatching), c5 (10M x 256, B=256)
    N  > 1 : c5 (10M x 256 row-striped over the N ranks: the scaling sweep BASELINE configs[4] names)

Every workload ends with a PARITY GATE outside any timing (``parity``): a subsample of the queries of the last
steps is searched again by the reference's own three torch calls (cslam/nns_matching.py:61-63, restated in
tests/refops.py) on the SAME resident database view, every rank scoring its own shard, and the library's
rows / similarities must satisfy BASELINE.md §5's rule; a violation makes the process exit non-zero.

``--impl reference`` times the reference's CPU path for the same metric: the reference is Python + torch
(cslam/nns_matching.py) and /root/reference does not exist on the benchmark box, so this arm runs the
reference's own three torch calls on CPU tensors (all host threads) and, beside it, the C oracle port
(oracle/nns_oracle.c, one pthread per core), both on a bounded row sample — see DESIGN.md §7.
"""
import argparse
import ctypes
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

WORKLOADS = {
    # name: (rows, dim, batch, k, description)
    "c3": (1_000_000, 4096, 256, 10, "BASELINE configs[2]: NetVLAD-shaped 1M x 4096 fp32, batch 256, top-10"),
    "c2": (100_000, 512, 1, 1, "BASELINE configs[1]: CosPlace-shaped 100k x 512 fp32, single query, top-1"),
    "c2_k10": (100_000, 512, 1, 10, "BASELINE configs[1]'s database, single query, top-10 (the intra-robot "
                                   "search of loop_closure_sparse_matching.py:80-81)"),
    "c4": (2_000_000, 256, 1, 1, "BASELINE configs[3]: 2M x 256 fp32, add_item between single-query top-1 searches"),
    "c5": (10_000_000, 256, 256, 10, "BASELINE configs[4]: 10M x 256 fp32 row-sharded, batch 256, top-10"),
}
METRIC = "NN-search queries/s at 1M×4096 DB; achieved HBM GB/s and fraction of roofline"
HBM_NOMINAL_GBS = 8000.0   # the "roughly 8 TB/s peak" BASELINE.json's north_star states its target against


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=None, choices=sorted(WORKLOADS) + ["sc", "swarm"])
    ap.add_argument("--no-legs", action="store_true", help="primary workload only")
    ap.add_argument("--robots", type=int, default=8, help="swarm workload: number of robots R")
    ap.add_argument("--rows", type=int, default=0, help="override the row count (debugging only)")
    ap.add_argument("--dim", type=int, default=0, help="override the descriptor size (debugging only)")
    ap.add_argument("--batch", type=int, default=0, help="override the batch size (debugging only)")
    ap.add_argument("--k", type=int, default=0, help="override k (debugging only)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-torch-bar", action="store_true",
                    help="skip the stock torch-on-CUDA restatement of the reference's search (gpu_reference)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the parity gate (debugging only)")
    ap.add_argument("--parity-queries", type=int, default=32)
    return ap.parse_args()


# -------------------------------------------------------------------------------------------------
# CPU arms — used by cpu_baseline and by --impl reference
# -------------------------------------------------------------------------------------------------

def _cpu_sample(rows, dim, nq, seed=3):
    import numpy as np
    sample_rows = min(rows, max(1024, (256 << 20) // (dim * 4)))      # <= 256 MB of descriptors
    rng = np.random.default_rng(seed)
    db = rng.random((sample_rows, dim), dtype=np.float32)
    db /= np.linalg.norm(db, axis=1, keepdims=True)
    qs = rng.random((nq, dim), dtype=np.float32)
    return sample_rows, db, qs


def cpu_port_qps(rows, dim, batch, k, budget_s=12.0):
    """Time the C oracle on a bounded row sample of the workload; returns (queries/s scaled to the
    full row count, cores, description of the sample)."""
    import numpy as np
    import subprocess
    lib_path = os.path.join(ROOT, "oracle", "_build", "liboracle.so")
    if not os.path.exists(lib_path):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle")])
    lib = ctypes.CDLL(lib_path)
    lib.oracle_search.restype = ctypes.c_int
    lib.oracle_search.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_int64, ctypes.c_int64,
                                  ctypes.c_void_p, ctypes.c_int64, ctypes.c_int, ctypes.c_void_p,
                                  ctypes.c_void_p]
    lib.oracle_num_threads.restype = ctypes.c_int
    cores = lib.oracle_num_threads()
    nq = batch
    sample_rows, db, qs = _cpu_sample(rows, dim, nq)
    idx = np.empty((nq, k), np.int64)
    sims = np.empty((nq, k), np.float32)

    def run(nquery):
        t0 = time.perf_counter()
        lib.oracle_search(db.ctypes.data, sample_rows, dim, dim, qs.ctypes.data, nquery, k,
                          idx.ctypes.data, sims.ctypes.data)
        return time.perf_counter() - t0

    run(min(nq, 2))                                   # warm
    t1 = run(min(nq, 4)) / min(nq, 4)                 # seconds per query on the sample
    nrun = int(max(1, min(nq, budget_s / max(t1, 1e-9))))
    t = run(nrun)
    qps_sample = nrun / t
    scale = rows / sample_rows
    desc = ("%d of %d queries over a %d-row sample (1/%.0f of the rows, cost scaled linearly in rows); "
            "C port of nns_matching.py:57-63, %d pthreads" % (nrun, nq, sample_rows, scale, cores))
    return qps_sample / scale, cores, desc


def cpu_torch_qps(rows, dim, batch, k, budget_s=12.0):
    """The reference's OWN arithmetic on the host: the three torch calls of cslam/nns_matching.py:61-63 on CPU
    tensors (what the reference class executes with its device forced to CPU), one query at a time like the
    reference's API, all intra-op threads torch has.  Bounded row sample, cost scaled linearly in rows."""
    import torch
    if torch.get_num_threads() < (os.cpu_count() or 1):   # torchrun exports OMP_NUM_THREADS=1
        torch.set_num_threads(os.cpu_count() or 1)
    sample_rows, db, qs = _cpu_sample(rows, dim, min(batch, 64))
    view = torch.from_numpy(db)
    tq = torch.from_numpy(qs)

    def one(j):
        q = tq[j % len(tq)].reshape(1, dim)
        sim = torch.nn.functional.cosine_similarity(view, q)                 # :61
        ns = torch.argsort(sim, descending=True)[:k]                         # :62
        return ns.tolist(), torch.gather(sim, 0, ns).cpu().numpy()           # :63

    one(0)
    t0 = time.perf_counter()
    one(1)
    t1 = max(time.perf_counter() - t0, 1e-6)
    nrun = int(max(2, min(256, budget_s / t1)))
    t0 = time.perf_counter()
    for j in range(nrun):
        one(j)
    t = time.perf_counter() - t0
    scale = rows / sample_rows
    threads = torch.get_num_threads()
    desc = ("%d single-query searches over a %d-row sample (1/%.0f of the rows, cost scaled linearly in rows); the "
            "reference's own torch calls (cosine_similarity + argsort + gather, nns_matching.py:61-63) on CPU "
            "tensors, torch %s, %d intra-op threads, os.cpu_count()=%s"
            % (nrun, sample_rows, scale, torch.__version__, threads, os.cpu_count()))
    return nrun / t / scale, threads, desc


def cpu_baseline_block(rows, dim, batch, k, budget_s):
    """cpu_baseline object: the reference's torch-CPU arithmetic (value) with the C port beside it."""
    v, threads, desc = cpu_torch_qps(rows, dim, batch, k, budget_s * 0.5)
    pv, cores, pdesc = cpu_port_qps(rows, dim, batch, k, budget_s * 0.5)
    return {"value": v, "unit": "queries/s", "cores": threads, "kind": "port", "sample": desc,
            "extrapolated": True,
            "c_port": {"value": pv, "unit": "queries/s", "cores": cores, "sample": pdesc}}


# -------------------------------------------------------------------------------------------------
# clocks sampler
# -------------------------------------------------------------------------------------------------

class ClockSampler(threading.Thread):
    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop_evt = threading.Event()

    def run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            names = {
                getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
                getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
                getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
                getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
            }
            get_reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
                getattr(nv, "nvmlDeviceGetCurrentClocksThrottleReasons")
            while not self._stop_evt.is_set():
                self.samples.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                mask = get_reasons(h)
                for bit, name in names.items():
                    if mask & bit:
                        self.reasons.add(name)
                time.sleep(0.05)
        except Exception as e:  # pragma: no cover
            self.reasons.add("sampler_error:%s" % type(e).__name__)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons)}


def gpu_reference_bar(view, q_block, k, nq=20):
    """The reference's search arithmetic (cslam/nns_matching.py:57-63: cosine_similarity over the stored
    rows, full argsort, gather, two device->host reads) restated with stock torch ops on the SAME device and
    the same resident database (the zero-copy `.data` view), one query at a time as the reference's API is.
    /root/reference itself cannot travel to the benchmark box; this is PyTorch library code used as a bar,
    never part of the product path."""
    import torch

    def one(q):
        sim = torch.nn.functional.cosine_similarity(view, q.view(1, -1))
        order = torch.argsort(sim, descending=True)[:k]
        return order.tolist(), torch.gather(sim, 0, order).cpu().numpy()

    nb_ = int(q_block.shape[0])
    try:
        one(q_block[0])
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for j in range(nq):
            one(q_block[j % nb_])
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / nq
    except RuntimeError as e:  # e.g. out of memory for the similarity temporaries
        return {"value": None, "unit": "queries/s", "kind": "port", "sample": "failed: %s" % str(e)[:80]}
    finally:
        torch.cuda.empty_cache()
    return {"value": 1.0 / dt, "unit": "queries/s", "kind": "port",
            "sample": "%d single-query searches (torch %s: cosine_similarity + argsort + gather + 2 D2H reads, the "
                      "reference's nns_matching.py:61-63), same B200, same resident database"
                      % (nq, torch.__version__)}


def measured_peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return p["hbm_gbs"], p["bf16_tflops"], p.get("bf16_tflops_sustained", p["bf16_tflops"]), "measured"
    except Exception:
        return 6650.0, 1590.0, 1400.0, "fallback"


_TF32_PEAK = None


def measure_tf32_peak(dev):
    """cuBLAS TF32 dense matmul on this GPU in this run (torch.matmul 8192^3 with allow_tf32, outside every
    timed region): best of 10 (burst) and a ~1.5 s back-to-back loop (sustained, under the power cap) — the
    denominator of the batched kernel's tensor roofline, instead of assuming bf16 / 2."""
    global _TF32_PEAK
    if _TF32_PEAK is not None:
        return _TF32_PEAK
    import torch
    try:
        old = torch.backends.cuda.matmul.allow_tf32
        torch.backends.cuda.matmul.allow_tf32 = True
        n = 8192
        a = torch.randn((n, n), device=dev, dtype=torch.float32)
        b = torch.randn((n, n), device=dev, dtype=torch.float32)
        c = torch.empty((n, n), device=dev, dtype=torch.float32)
        for _ in range(3):
            torch.matmul(a, b, out=c)
        torch.cuda.synchronize()
        best = 1e9
        for _ in range(10):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            torch.matmul(a, b, out=c)
            e1.record()
            e1.synchronize()
            best = min(best, e0.elapsed_time(e1))
        reps = int(max(10, 1500.0 / best))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            torch.matmul(a, b, out=c)
        e1.record()
        e1.synchronize()
        sust = e0.elapsed_time(e1) / reps
        torch.backends.cuda.matmul.allow_tf32 = old
        del a, b, c
        torch.cuda.empty_cache()
        fl = 2.0 * n ** 3 / 1e12
        _TF32_PEAK = {"burst": fl / (best / 1e3), "sustained": fl / (sust / 1e3),
                      "how": "torch.matmul fp32 8192^3 with allow_tf32 (cuBLAS tf32 tensor-core GEMM) on this GPU in "
                             "this run: best of 10 and %d back to back" % reps}
    except Exception as e:  # pragma: no cover
        _TF32_PEAK = {"burst": None, "sustained": None, "how": "failed: %s" % str(e)[:80]}
    return _TF32_PEAK


# -------------------------------------------------------------------------------------------------
def reference_arm(args, name, rows, dim, batch, k):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    vals, pvals = [], []
    desc, pdesc, threads, cores = "", "", 1, 1
    per = max(2.0, 60.0 / max(1, args.steps + args.warmup))
    for i in range(args.warmup + args.steps):
        qps, threads, desc = cpu_torch_qps(rows, dim, batch, k, budget_s=per * 0.6)
        pq, cores, pdesc = cpu_port_qps(rows, dim, batch, k, budget_s=per * 0.4)
        if i >= args.warmup:
            vals.append(qps)
            pvals.append(pq)
    v = sum(vals) / len(vals)
    pv = sum(pvals) / len(pvals)
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "queries/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": batch / v * 1e3,
            "extrapolated": True,
            "extrapolation": "each step times a bounded sample of the workload on the host (rows and queries as "
                             "stated in cpu_baseline.sample) and scales linearly in rows; ms_per_step is the "
                             "extrapolated cost of one full %d-query step, not the time this process ran" % batch,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": workload_config(name, args.gpus, rows, dim, batch, k),
            "cpu_baseline": {"value": v, "unit": "queries/s", "cores": threads, "kind": "port", "sample": desc,
                             "extrapolated": True,
                             "c_port": {"value": pv, "unit": "queries/s", "cores": cores, "sample": pdesc}},
            "e2e": {"value": v, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def workload_config(name, gpus, rows, dim, batch, k):
    return {"workload": WORKLOADS[name][4], "rows": rows, "dim": dim, "batch": batch, "k": k,
            "sharding": "row-striped over %d rank(s), one exchange of B*k keys per rank per step" % gpus
            if gpus > 1 else "single GPU",
            "l2": "database (%.2f GB) is far larger than the 126 MB L2; no flush needed" % (rows * dim * 4 / 1e9)}


def bench_scancontext(args):
    """ScanContextMatching.search through the public Python API (host query in, host result
    out), next to the oracle port of the reference on the host (numpy; the reference itself additionally
    rebuilds a KD-tree per search).  One step = one search."""
    import numpy as np
    import torch
    import synth
    from cslam_b200 import _native
    from cslam_b200.lidar_pr import ScanContextMatching
    from oracle import sc_oracle
    n = args.rows or 50_000
    torch.cuda.set_device(0)
    lib = _native.load()
    base = synth.scancontexts(61, 2000)
    m = ScanContextMatching()
    for i in range(n):   # distinct descriptors: a base scan, rolled and scaled
        m.add_item(np.roll(base[i % 2000].reshape(20, 60), i // 2000, axis=1) * (1.0 + 1e-3 * (i // 2000)), i)
    qs = synth.sc_queries_for(1061, base, 64)
    for j in range(args.warmup):
        m.search(qs[j % 64], 1)
    torch.cuda.synchronize()
    l0 = lib.cslam_nns_launch_count()
    t0 = time.perf_counter()
    for j in range(args.steps):
        m.search(qs[j % 64], 1)
    dt = time.perf_counter() - t0
    launches = lib.cslam_nns_launch_count() - l0
    # CPU: the oracle on a bounded sample of the same database (cost is linear in rows for step 1 and
    # constant for step 2, so it is NOT scaled: the sample is stated)
    ns = min(n, 20000)
    db = np.stack([np.roll(base[i % 2000].reshape(20, 60), i // 2000, axis=1).reshape(-1) for i in range(ns)])
    items = {i: i for i in range(ns)}
    t1 = time.perf_counter()
    reps = 3
    for j in range(reps):
        sc_oracle.search(db, items, qs[j])
    cpu = reps / (time.perf_counter() - t1)
    hbm_peak, _, _, which = measured_peaks()
    bytes_ = n * 20 * 8 + 10 * 1200 * 8
    line = {"metric": "ScanContext searches/s (ring-key shortlist of 10 + column-shift distance)",
            "value": args.steps / dt, "unit": "searches/s", "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": "SURVEY 8 f4: ScanContextMatching.search, %d x (20 x 60) float64 database" % n,
                       "rows": n, "l2": "latency-bound at SLAM sizes: 2 kernels + 1 H2D + 2 D2H per search"},
            "roofline": {"bound": "hbm", "achieved": bytes_ / (dt / args.steps) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                         "frac": bytes_ / (dt / args.steps) / 1e9 / hbm_peak, "traffic": None,
                         "note": "algorithmic bytes = ring keys n*20*8 + 10 candidates * 9.6 KB; the call is "
                                 "latency-bound (host round trip), the fraction is reported for completeness"},
            "cpu_baseline": {"value": cpu, "unit": "searches/s", "cores": 1, "kind": "port",
                             "sample": "%d searches over the first %d rows with oracle/sc_oracle.py (numpy, 1 thread)" % (reps, ns)},
            "e2e": {"value": args.steps / dt, "unit": "searches/s", "h2d_bytes_per_step": 1220 * 8, "d2h_bytes_per_step": 200},
            "gpu_launches": int(launches)}
    print(json.dumps(line))


def run_swarm(args, steps, warmup, cpu_budget=12.0):
    """SURVEY §8 d1 C4 in its swarm form + f1/f2: R robots (default 8), the 2M x 256 database of BASELINE
    configs[3] split into R per-robot databases of 2M/R rows.  One step = one local keyframe through the
    drop-in ``LoopClosureSparseMatching.add_local_global_descriptor`` (1 add_item on the local database +
    R-1 search_best on the other robots' databases, threshold test, edge bookkeeping) — host descriptor in,
    host matches out, so the whole leg is end to end.  Timed three times on the same object: through
    ``search_best_multi`` (one staging, R-1 kernels back to back, ONE synchronisation; the default), the same
    without the launch overlap, and through the reference's per-robot loop (R-1 synchronised searches).
    Parity: the (keyframe, similarity) every robot's database returns for a subsample of the descriptors is
    checked against the reference's torch ops on that database's view."""
    import numpy as np
    import torch
    import refops
    from cslam_b200 import _native
    from cslam_b200.loop_closure_sparse_matching import CandidateEdges, LoopClosureSparseMatching
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    lib = _native.load()
    R, dim = args.robots, args.dim or 256
    rows = args.rows or 2_000_000
    per = rows // R
    params = {"robot_id": 0, "max_nb_robots": R, "frontend.sensor_type": "stereo",
              "frontend.similarity_threshold": 0.9, "frontend.nb_best_matches": 10,
              "frontend.intra_loop_min_inbetween_keyframes": 20}
    lcsm = LoopClosureSparseMatching(params, candidate_selector=CandidateEdges(0, R))
    gen = torch.Generator(device=dev)
    gen.manual_seed(4)
    probe = None
    for r in range(R):
        x = torch.rand((per, dim), generator=gen, device=dev, dtype=torch.float32)
        x /= x.norm(dim=1, keepdim=True)
        (lcsm.local_nnsm if r == 0 else lcsm.other_robots_nnsm[r]).add_items(x, list(range(per)))
        if r == 1:
            probe = x[:4096].cpu().numpy()
        del x
    # host-timed leg: let the worker threads of the previous leg's CPU baseline (OpenMP pools spin for a while after
    # a parallel region) go idle before anything is timed on the host
    time.sleep(1.0)
    nq = warmup + steps
    rng = np.random.default_rng(5)
    qs = rng.random((nq, dim)).astype(np.float32)
    qs[::4] = probe[: len(qs[::4])] + 0.02 * rng.standard_normal((len(qs[::4]), dim)).astype(np.float32)
    qs /= np.linalg.norm(qs, axis=1, keepdims=True)          # every 4th keyframe revisits a place of robot 1

    def run(fused, lo, hi):
        lcsm._dense = fused       # False -> the reference's per-robot loop over search_best
        n_match = 0
        for i in range(lo, hi):
            n_match += len(lcsm.add_local_global_descriptor(qs[i], per + i))
        return n_match

    out = {}
    for name, fused in (("loop", False), ("multi_serial", True), ("multi", True)):
        # multi_serial: the same single call, but every kernel waits for the previous one to drain
        os.environ["CSLAM_NNS_MULTI_OVERLAP"] = "0" if name == "multi_serial" else "1"
        run(fused, 0, warmup)
        torch.cuda.synchronize()
        sampler = ClockSampler(0)
        sampler.start()
        l0 = lib.cslam_nns_launch_count()
        t0 = time.perf_counter()
        nm = run(fused, warmup, nq)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        out[name] = {"s_per_step": dt / steps, "matches": nm, "launches": lib.cslam_nns_launch_count() - l0,
                     "clocks": sampler.stop()}
    assert out["loop"]["matches"] == out["multi"]["matches"]
    lcsm._dense = True
    # parity gate: search_best_multi's answers for a subsample of the descriptors, per robot database
    parity = None
    if not args.no_parity:
        from cslam_b200 import NearestNeighborsMatching
        robots = [i for i in range(R) if i != 0]
        sub = list(range(0, nq, max(1, nq // min(args.parity_queries, nq))))[:args.parity_queries]
        res = [NearestNeighborsMatching.search_best_multi([lcsm.other_robots_nnsm[i] for i in robots], qs[j])
               for j in sub]
        qd = torch.from_numpy(qs[sub]).to(dev)
        rep = None
        for a, i in enumerate(robots):
            mch = lcsm.other_robots_nnsm[i]
            rows_ = np.asarray([[res[b][a][0]] for b in range(len(sub))], np.int64)
            sims_ = np.asarray([[res[b][a][1]] for b in range(len(sub))], np.float32)
            r_ = refops.check_against_reference(mch.data[:mch.n], qd, 1, rows_, sims_)
            if rep is None:
                rep = r_
            else:
                for f in ("queries", "non_tie", "ties", "bad_sims", "bad_fp64"):
                    setattr(rep, f, getattr(rep, f) + getattr(r_, f))
                rep.max_rel, rep.max_abs64 = max(rep.max_rel, r_.max_rel), max(rep.max_abs64, r_.max_abs64)
                rep.first = rep.first or r_.first
        parity = rep.as_dict()
        parity["queries"] = "%d descriptors x %d robot databases" % (len(sub), len(robots))
    hbm_peak, _, _, which = measured_peaks()
    dt = out["multi"]["s_per_step"]
    bytes_ = float(R - 1) * per * dim * 4
    cpu = None
    if not args.no_cpu_baseline:
        cpu = cpu_baseline_block(per, dim, 1, 1, cpu_budget)
    qps = (R - 1) / dt
    line = {"metric": METRIC, "value": qps, "unit": "queries/s", "n_gpus": 1, "steps": steps,
            "warmup": warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "BASELINE configs[3] as a swarm: %d robots, %d x %d fp32 per robot database, one "
                                   "step = add_local_global_descriptor (1 add_item + %d search_best + threshold)"
                                   % (R, per, dim, R - 1),
                       "rows": per * R, "dim": dim, "batch": 1, "k": 1, "robots": R,
                       "l2": "the %d databases searched per step total %.2f GB, far larger than the 126 MB L2; no "
                             "flush needed" % (R - 1, bytes_ / 1e9)},
            "roofline": {"bound": "hbm", "achieved": bytes_ / dt / 1e9, "peak": hbm_peak, "unit": "GB/s",
                         "frac": bytes_ / dt / 1e9 / hbm_peak, "frac_nominal_8tbs": bytes_ / dt / 1e9 / HBM_NOMINAL_GBS,
                         "traffic": None, "kernel": "k1_single_query",
                         "note": "END-TO-END figure: (R-1) * rows * ld * 4 algorithmic bytes per step / wall time of "
                                 "the whole Python call (H2D of the row and the query, R-1 kernels, one sync, edge "
                                 "bookkeeping); peak = %s copy bandwidth" % which},
            "cpu_baseline": cpu,
            "e2e": {"value": qps, "unit": "queries/s", "h2d_bytes_per_step": 2 * dim * 4,
                    "d2h_bytes_per_step": (R - 1) * 12},
            "per_robot_loop": {"value": (R - 1) / out["loop"]["s_per_step"], "unit": "queries/s",
                               "ms_per_step": out["loop"]["s_per_step"] * 1e3,
                               "note": "same kernels through the reference's loop of R-1 synchronised search_best calls"},
            "multi_without_overlap": {"value": (R - 1) / out["multi_serial"]["s_per_step"], "unit": "queries/s",
                                      "ms_per_step": out["multi_serial"]["s_per_step"] * 1e3,
                                      "note": "one call, one synchronisation, but ordinary launches: each database's "
                                              "kernel starts only after the previous one has drained"},
            "clocks": out["multi"]["clocks"], "gpu_launches": int(out["multi"]["launches"]),
            "matches": out["multi"]["matches"], "parity": parity}
    del lcsm
    torch.cuda.empty_cache()
    return line


class Ctx(object):
    """Process-wide state of one bench run (device, ranks, library)."""

    def __init__(self, args):
        import torch
        import torch.distributed as dist
        from cslam_b200 import _native
        self.args = args
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(self.local_rank)
        self.dev = torch.device("cuda", self.local_rank)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=self.dev)
        assert self.world == args.gpus, "launch with torchrun --nproc-per-node %d" % args.gpus
        self.lib = _native.load()

    def barrier(self):
        import torch
        import torch.distributed as dist
        if self.world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(self, v):
        import torch
        import torch.distributed as dist
        t = torch.tensor([v], dtype=torch.float64, device=self.dev)
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())


def run_dense(ctx, name, rows, dim, batch, k, steps, warmup, cpu_budget=12.0, overridden=False):
    """One dense-matcher workload on ctx.world ranks; returns the line dict on rank 0 (None elsewhere) and
    whether the parity gate passed (identical on every rank)."""
    import numpy as np
    import torch
    import torch.distributed as dist
    import refops
    from cslam_b200.sharded import ShardedNearestNeighborsMatching
    args, world, rank, dev, lib = ctx.args, ctx.world, ctx.rank, ctx.dev, ctx.lib
    single = batch == 1

    # ---- synthetic database: this rank's stripe, generated on the device in chunks ------------------
    m = ShardedNearestNeighborsMatching(dim=dim)
    my_rows = (rows - rank + world - 1) // world
    gen = torch.Generator(device=dev)
    gen.manual_seed(3 + 1000 * rank)
    chunk = max(1, min(my_rows, (1 << 30) // (dim * 4)))
    done = 0
    probe = None
    while done < my_rows:
        c = min(chunk, my_rows - done)
        x = torch.rand((c, dim), generator=gen, device=dev, dtype=torch.float32)
        x /= x.norm(dim=1, keepdim=True)
        if probe is None:
            probe = x[: max(4 * batch, 64)].clone()
        m.add_local_shard(x)
        done += c
        del x
    m.finalize_bulk()
    assert m.n == rows
    # queries: same on every rank; every 10th is a noisy copy of a rank-0 row (true neighbours exist)
    gq = torch.Generator(device=dev)
    gq.manual_seed(4)
    nb = 4 if not single else 64
    queries = torch.rand((nb * batch, dim), generator=gq, device=dev, dtype=torch.float32)
    if probe.shape[0] >= queries.shape[0]:
        queries[::10] = probe[: queries.shape[0]][::10] + 0.02 * torch.randn(
            (queries[::10].shape[0], dim), generator=gq, device=dev)
    queries /= queries.norm(dim=1, keepdim=True)
    if world > 1:
        dist.broadcast(queries, 0)
    qb = [queries[i * batch:(i + 1) * batch].contiguous() for i in range(nb)]
    out_idx = torch.empty((batch, k), dtype=torch.int64, device=dev)
    out_sim = torch.empty((batch, k), dtype=torch.float32, device=dev)
    h = m.local._h

    # BASELINE configs[3] (c4): "incremental add_item between queries" — every step first appends one new
    # descriptor (device-resident leg: a stream-ordered device-to-device row copy; end-to-end leg: the
    # public add_item with a host row), then searches.  The database grows by one row per step.
    incremental = name == "c4" and world == 1
    extra = None
    if incremental:
        extra = torch.rand((64, dim), generator=gq, device=dev, dtype=torch.float32)
        extra /= extra.norm(dim=1, keepdim=True)

    last_block = [0]   # query block of the most recent device_step: its results are what out_idx / out_sim hold

    # one prepared call per query block (ShardedNearestNeighborsMatching.plan_search_device: pointers and shapes
    # converted once), so that the Python loop can keep ahead of a 32 us kernel
    plans = [m.plan_search_device(qb[j], k, out_idx, out_sim) for j in range(nb)] if world == 1 else None

    def device_step(i):
        if incremental:
            m.add_local_shard(extra[i % 64:i % 64 + 1])
        if plans is not None:
            plans[i % nb]()
        else:
            m.search_batch_device(qb[i % nb], k, out_idx, out_sim)
        last_block[0] = i % nb

    # ---- device-resident leg: `value` (CUDA events around the K steps, nothing between the launches) ----
    for i in range(warmup):
        device_step(i)
    ctx.barrier()
    sampler = ClockSampler(ctx.local_rank)
    sampler.start()
    launches0 = lib.cslam_nns_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        device_step(i)
    e1.record()
    ctx.barrier()
    ms = e0.elapsed_time(e1)
    launches = lib.cslam_nns_launch_count() - launches0
    clocks = sampler.stop()
    ms = ctx.max_over_ranks(ms)
    # ---- roofline pass: the same steps again with the dominant kernel of every step bracketed by CUDA events on
    # its launch stream inside the library (the events keep consecutive searches from overlapping, which is why
    # this is a separate pass and not the timed one) -----------------------------------------------------------
    psteps = min(steps, 200)
    lib.cslam_nns_profile(h.h, 1)
    pe0, pe1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    pe0.record()
    for i in range(psteps):
        device_step(i)
    pe1.record()
    ctx.barrier()
    pass_ms = pe0.elapsed_time(pe1) / psteps      # a whole step of this pass (dominant kernel + everything else)
    kms, kl = ctypes.c_float(), ctypes.c_int64()
    lib.cslam_nns_profile_read(h.h, ctypes.byref(kms), ctypes.byref(kl))
    lib.cslam_nns_profile(h.h, 0)
    b2b = None
    if single and rank == 0:
        # the same search queued 200x inside the library (no Python between the launches)
        b2b = ctypes.c_float()
        lib.cslam_nns_time_search(h.h, ctypes.c_void_p(qb[0].data_ptr()), batch, dim, k, 200,
                                  ctypes.c_void_p(torch.cuda.current_stream().cuda_stream), ctypes.byref(b2b))
    if clocks["sm_mhz"] is None:
        # the timed region was shorter than one NVML poll (50 ms): repeat the same step for ~0.5 s right
        # after it, outside any timing, only to read the clocks under this load (same count on every rank)
        reps = int(min(200000, max(8, 0.5 / max(ms / 1e3 / steps, 1e-6))))
        sampler = ClockSampler(ctx.local_rank)
        sampler.start()
        time.sleep(0.1)
        for i in range(reps):
            device_step(i)
        ctx.barrier()
        clocks = sampler.stop()
        clocks["sampled"] = "%d untimed repeats of the step right after the timed region" % reps
    else:
        clocks["sampled"] = "during the timed region"
    value = steps * batch / (ms / 1e3)
    variant = lib.cslam_nns_last_variant(h.h)
    screen = None
    if hasattr(lib, "cslam_nns_screen_stats"):
        a_, b_ = ctypes.c_int64(), ctypes.c_int64()
        lib.cslam_nns_screen_stats(h.h, ctypes.byref(a_), ctypes.byref(b_))
        if a_.value:
            screen = {"queries_screened": int(a_.value), "queries_fallback": int(b_.value),
                      "fallback_rate": b_.value / max(1, a_.value)}

    # ---- results of the last timed step, kept for the parity gate ------------------------------------
    if single:
        par_q = queries[: min(args.parity_queries, queries.shape[0])].contiguous()
        par_idx = torch.empty((par_q.shape[0], k), dtype=torch.int64, device=dev)
        par_sim = torch.empty((par_q.shape[0], k), dtype=torch.float32, device=dev)
        for j in range(par_q.shape[0]):
            m.search_batch_device(par_q[j:j + 1], k, par_idx[j:j + 1], par_sim[j:j + 1])
    else:
        last = last_block[0]
        npar = min(args.parity_queries, batch)
        par_q = qb[last][:npar].contiguous()
        par_idx, par_sim = out_idx[:npar].clone(), out_sim[:npar].clone()
    torch.cuda.synchronize()

    # ---- end-to-end leg: public Python API, host buffers --------------------------------------------
    e2e = None
    if not args.no_e2e:
        if world == 1:   # the drop-in's items dict: row -> keyframe id (an int, as in production); the 10M-row
            # database gets the storage-free identity mapping instead of a 10M-entry dict
            from cslam_b200.sharded import RowIdentity
            m.local.items = dict(zip(range(rows + steps + 8), range(rows + steps + 8))) \
                if rows <= 4_000_000 else RowIdentity()
        hq = [torch.from_numpy(q.cpu().numpy()).pin_memory() for q in qb]
        hq_np = [q.numpy() for q in hq]
        extra_host = extra.cpu() if incremental else None

        def e2e_step(i):
            if world > 1:     # every rank uploads 1/world of the block, one all-gather, results to rank 0 only
                return m.search_batch(hq[i % nb], k, root_only=True)
            if incremental:   # the reference's add_item: one host row, item = its row number
                m.local.add_item(extra_host[m.local.n % 64], m.local.n)
            if not single:    # host queries in, host (items, sims) out through the public batched call
                return m.local.search_batch(hq[i % nb], k)
            if k == 1:        # the reference's signature: numpy descriptor in, (item, np.float32) out
                return m.local.search_best(hq_np[i % nb][0])
            return m.local.search(hq_np[i % nb][0], k)

        for i in range(min(warmup, 3)):
            e2e_step(i)
        ctx.barrier()
        t0 = time.perf_counter()
        for i in range(steps):
            e2e_step(i)
        torch.cuda.synchronize()
        dt = ctx.max_over_ranks(time.perf_counter() - t0)
        sync_e2e = None
        if not single:
            # the batched workloads' headline end-to-end figure goes through the pipelined public call
            # (search_batches: a generator over host blocks; block i+1's upload and block i-1's result read overlap
            # block i's search).  Every step's H2D and D2H happen inside the timed region; the one-call-at-a-time
            # figure measured above is kept beside it as `sync`.
            sync_e2e = {"value": steps * batch / dt, "ms_per_step": dt / steps * 1e3,
                        "api": "search_batch(host queries, k), one synchronous call per step"}

            def run_stream(nsteps):
                blocks = (hq[i % nb] for i in range(nsteps))
                gen = m.search_batches(blocks, k, root_only=True) if world > 1 else \
                    m.local.search_batches(blocks, k)
                got = 0
                for _res in gen:
                    got += 1
                assert got == nsteps

            run_stream(min(warmup, 3) + 2)
            torch.cuda.synchronize()
            ctx.barrier()
            t0 = time.perf_counter()
            run_stream(steps)
            torch.cuda.synchronize()
            dt = ctx.max_over_ranks(time.perf_counter() - t0)
        e2e = {"value": steps * batch / dt, "unit": "queries/s", "ms_per_step": dt / steps * 1e3,
               "h2d_bytes_per_step": batch * dim * 4 + (dim * 4 if incremental else 0),
               "d2h_bytes_per_step": batch * k * 12,
               "api": ("ShardedNearestNeighborsMatching.search_batches(stream of host blocks, k, root_only=True): "
                       "every rank uploads each block over its own PCIe link on a copy stream, two blocks in flight, "
                       "rank 0 alone reads the results (pinned, own stream)") if world > 1 else
                      ("NearestNeighborsMatching.search_batches(stream of host blocks, k) -> (items, sims) per "
                       "block: uploads, searches and result reads of consecutive blocks overlap (two in flight)"
                       if not single else
                       ("add_item(host row) + " if incremental else "") +
                       ("search_best(numpy descriptor)" if k == 1 else "search(numpy descriptor, k)"))}
        if sync_e2e is not None:
            e2e["pipelined"] = True
            e2e["sync"] = sync_e2e

    # ---- parity gate (outside any timing) ------------------------------------------------------------
    parity, ok = None, True
    if not args.no_parity:
        view = m.local.data[:m.local.n]
        rep = refops.check_against_reference(view, par_q, k, par_idx, par_sim, rank, world)
        parity = rep.as_dict()
        ok = rep.ok
        if world > 1:
            # the NCCL all-gather + K4 route must return the same bits as the peer-memory kernel, and a repeat of
            # the timed step's batch the same bits as the timed step
            qfull = qb[last_block[0]] if not single else par_q
            i2, s2 = m.search_batch_device(qfull, k, exchange="nccl")
            i1, s1 = m.search_batch_device(qfull, k)
            torch.cuda.synchronize()
            m.check_exchange()
            npar_ = par_idx.shape[0]
            same = bool(torch.equal(i1, i2) and torch.equal(s1, s2) and torch.equal(i1[:npar_], par_idx)
                        and torch.equal(s1[:npar_], par_sim))
            flag = torch.tensor([1 if same else 0], device=dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            parity["peer_route_equals_nccl_route_bitwise"] = bool(int(flag.item()))
            parity["ok"] = parity["ok"] and parity["peer_route_equals_nccl_route_bitwise"]
            ok = parity["ok"]

    # ---- the reference's own GPU arithmetic on the same box (SURVEY §8 d3 iii) -----------------------
    gpu_ref = None
    if world == 1 and not args.no_torch_bar:
        gpu_ref = gpu_reference_bar(m.local.data[:m.local.n], qb[0] if not single else queries, k)

    tf32 = measure_tf32_peak(dev) if (variant == 2 and rank == 0) else None
    xchg_peer = getattr(m, "_xchg", None) is not None
    del m, h
    torch.cuda.empty_cache()
    if rank != 0:
        return None, ok

    # ---- roofline of the dominant kernel --------------------------------------------------------------
    hbm_peak, tf_burst, tf_sust, which = measured_peaks()
    k_ms = kms.value / max(1, kl.value)
    local_rows = my_rows
    if variant == 2:
        per_launch_q = min(batch, 256)
        flops = 2.0 * local_rows * dim * per_launch_q
        achieved = flops / (k_ms / 1e3) / 1e12
        if tf32 and tf32["sustained"]:
            peak, peak_note = tf32["sustained"], "cuBLAS tf32 GEMM measured in this run, sustained (%s)" % tf32["how"]
        else:
            peak, peak_note = tf_sust / 2.0, "%s bf16 sustained %.0f TF/s / 2" % (which, tf_sust)
        bytes_ = float(local_rows) * ((dim + 3) // 4 * 4) * 4
        roof = {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                "frac": achieved / peak, "traffic": None, "kernel": "k2_tcgen05",
                "kernel_ms": k_ms, "launches": kl.value,
                "peak_tf32_burst": tf32["burst"] if tf32 else None, "peak_bf16_sustained_half": tf_sust / 2.0,
                "step_minus_kernel_ms": pass_ms - k_ms * (kl.value / max(1, psteps)),
                "hbm_GBps": bytes_ / (k_ms / 1e3) / 1e9, "hbm_frac": bytes_ / (k_ms / 1e3) / 1e9 / hbm_peak,
                "hbm_frac_nominal_8tbs": bytes_ / (k_ms / 1e3) / 1e9 / HBM_NOMINAL_GBS,
                "note": "achieved = algorithmic 2*N*D*B flops per launch of the dominant kernel / its CUDA-event "
                        "duration; step_minus_kernel_ms = what a step costs besides that kernel (query preparation, "
                        "re-scoring, merges, exchange, launch gaps), same pass; peak = %s; hbm_* = N*ld*4 database bytes / the same duration against the %s copy "
                        "peak %.0f GB/s" % (peak_note, which, hbm_peak)}
        if screen:
            roof["screening"] = screen
    else:
        bytes_ = float(local_rows) * ((dim + 3) // 4 * 4) * 4
        # A step of these workloads IS one launch of the single-query kernel (c4: plus a 1 KB row append), so the
        # kernel's average launch duration over the timed region is the region's CUDA-event time / K: no events
        # between the launches, consecutive searches chained as in production.  The per-launch figure of the
        # separate pass with events around every launch (which serialises them) is kept as bracketed_*.
        step_ms = ms / steps
        achieved = bytes_ / (step_ms / 1e3) / 1e9
        bracketed = bytes_ / (k_ms / 1e3) / 1e9
        b2b_ms = b2b.value / max(1, batch)
        roof = {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                "frac": achieved / hbm_peak, "frac_nominal_8tbs": achieved / HBM_NOMINAL_GBS,
                "traffic": None, "kernel": "k1_single_query",
                "kernel_ms": step_ms, "launches": int(steps),
                "bracketed_kernel_ms": k_ms, "bracketed_launches": kl.value, "bracketed_GBps": bracketed,
                "bracketed_frac": bracketed / hbm_peak,
                "kernel_ms_back_to_back": b2b_ms,
                "back_to_back_GBps": bytes_ / (b2b_ms / 1e3) / 1e9,
                "back_to_back_frac": bytes_ / (b2b_ms / 1e3) / 1e9 / hbm_peak,
                "back_to_back_frac_nominal_8tbs": bytes_ / (b2b_ms / 1e3) / 1e9 / HBM_NOMINAL_GBS,
                "stream_ms_per_search": step_ms,
                "stream_GBps": achieved,
                "stream_frac": achieved / hbm_peak,
                "stream_frac_nominal_8tbs": achieved / HBM_NOMINAL_GBS,
                "note": "achieved = N*ld*4 algorithmic bytes per launch / the kernel's average launch duration over "
                        "the timed region (CUDA events at both ends of the K steps / K: a step is one launch%s; "
                        "Python-driven stream, consecutive searches of one database chained by programmatic "
                        "dependent launch%s); bracketed_* = the same with CUDA events around EVERY launch in a "
                        "separate pass, which serialises the launches; back_to_back = the same search queued 200x "
                        "inside the library; peak = %s copy bandwidth"
                        % (" plus a 1 KB row append" if incremental else "",
                           " — here the append sits between the searches, so nothing is chained" if incremental
                           else "", which)}

    # traffic: DRAM bytes per launch of the dominant kernel from the committed ncu --set full capture of this
    # workload (profiles/ncu_traffic.json), null when none exists for it
    try:
        tr = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json"))).get(name)
        if tr and world == 1 and not overridden:
            roof["traffic"] = tr["dram_bytes_read"] + tr["dram_bytes_write"]
            roof["traffic_source"] = tr["source"]
            if "tensor_pipe_active_pct" in tr:   # sm__pipe_tensor_cycles_active, same capture
                roof["tensor_pipe_active_pct_ncu"] = tr["tensor_pipe_active_pct"]
    except Exception:
        pass

    cpu = None
    if not args.no_cpu_baseline:
        cpu = cpu_baseline_block(rows, dim, batch, k, cpu_budget)

    line = {"metric": METRIC, "value": value, "unit": "queries/s", "n_gpus": world, "steps": steps,
            "warmup": warmup, "ms_per_step": ms / steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None,
            "dtype": "f32 (tf32 tensor-core products, fp32 accumulate; every returned similarity is an fp32-exact "
                     "3xTF32 or fp32 re-scored value)" if variant == 2 else "f32", "data": "synthetic",
            "config": dict(workload_config(name, world, rows, dim, batch, k), search_variant=variant,
                           **({"step": "1 add_item (one new row) + 1 search; the database grows by one row per step"}
                              if incremental else {}),
                           **({"exchange": "K5 peer-memory kernel (csrc/k5_exchange.cu)"
                               if xchg_peer else "NCCL all-gather + K4"} if world > 1 else {})),
            "roofline": roof, "cpu_baseline": cpu, "e2e": e2e, "clocks": clocks, "gpu_launches": int(launches),
            "parity": parity}
    if gpu_ref is not None:
        line["gpu_reference"] = gpu_ref
    return line, ok


def leg_steps(name, steps):
    """Short legs: enough steps for a stable figure, bounded so that the default run stays within minutes."""
    return {"c2": max(steps, 1000), "c2_k10": max(steps, 1000), "c4": max(steps, 300), "c5": max(steps, 10),
            "swarm": max(steps, 200)}.get(name, steps)


def main():
    args = parse()
    if args.workload == "sc":
        return bench_scancontext(args)
    name = args.workload or "c3"
    if name == "swarm":
        if args.impl == "reference":   # the CPU arm of one per-robot database: BASELINE configs[3]'s shape
            return reference_arm(args, "c4", (args.rows or 2_000_000) // args.robots, args.dim or 256, 1, 1)
        line = run_swarm(args, args.steps, args.warmup)
        print(json.dumps(line))
        return 0 if (line["parity"] is None or line["parity"]["ok"]) else 3
    rows, dim, batch, k, _ = WORKLOADS[name]
    overridden = bool(args.rows or args.dim or args.batch or args.k)
    rows, dim, batch, k = args.rows or rows, args.dim or dim, args.batch or batch, args.k or k
    if args.impl == "reference":
        return reference_arm(args, name, rows, dim, batch, k)

    import torch.distributed as dist
    ctx = Ctx(args)
    line, ok = run_dense(ctx, name, rows, dim, batch, k, args.steps, args.warmup, overridden=overridden)
    all_ok = ok
    if args.workload is None and not args.no_legs and not overridden:
        legs = {}
        names = ["c2", "c2_k10", "c4", "swarm", "c5"] if ctx.world == 1 else ["c5"]
        for leg in names:
            if leg == "swarm":
                l_ = run_swarm(args, leg_steps(leg, args.steps), max(args.warmup, 20), cpu_budget=3.0)
                ok_ = l_["parity"] is None or l_["parity"]["ok"]
            else:
                r_, d_, b_, k_, _ = WORKLOADS[leg]
                l_, ok_ = run_dense(ctx, leg, r_, d_, b_, k_, leg_steps(leg, args.steps), max(args.warmup, 3),
                                    cpu_budget=3.0)
            all_ok = all_ok and ok_
            if ctx.rank == 0:
                l_.pop("metric", None)
                legs[leg] = l_
        if ctx.rank == 0:
            line["legs"] = legs
    if ctx.rank == 0:
        line["parity_all_ok"] = bool(all_ok)
        print(json.dumps(line))
        if not all_ok:
            sys.stderr.write("PARITY GATE FAILED: see the `parity` blocks of the line above\n")
    if ctx.world > 1:
        dist.destroy_process_group()
    return 0 if all_ok else 3


if __name__ == "__main__":
    sys.exit(main() or 0)
