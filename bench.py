#!/usr/bin/env python
"""Headline benchmark: NN-search queries/s on the configuration BASELINE.json's metric is quoted on.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c3|c2|c4|c5|swarm|sc]

Default workload ``c3`` = BASELINE.json configs[2]: 1M x 4096 fp32 descriptors, batches of 256
queries, top-10 (the "1M x 4096 DB" the metric names).  One *step* = one batch of 256 queries
searched against the whole database.  With N > 1 ranks (torchrun) the SAME database is row-striped
over the ranks (strong scaling), every step ends with one all-gather of the B*k packed keys and
the merge kernel, and timing is the max over ranks of a CUDA-event interval.

The line printed by rank 0 carries, besides the base contract: ``roofline`` (dominant kernel,
timed live with CUDA events around that kernel only), ``cpu_baseline`` (the C oracle port on the
host cores, bounded sample), ``e2e`` (same metric through the public Python API with HOST buffers),
``clocks`` and ``gpu_launches``.

``--impl reference`` times the reference's CPU path for the same metric: the reference is Python +
torch (cslam/nns_matching.py) and /root/reference does not exist on the benchmark box, so this arm
runs the oracle port (oracle/nns_oracle.c, one pthread per core) — see DESIGN.md.
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
    "c4": (2_000_000, 256, 1, 1, "BASELINE configs[3]: 2M x 256 fp32, add_item between single-query top-1 searches"),
    "c5": (10_000_000, 256, 256, 10, "BASELINE configs[4]: 10M x 256 fp32 row-sharded, batch 256, top-10"),
}
METRIC = "NN-search queries/s at 1M×4096 DB; achieved HBM GB/s and fraction of roofline"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS) + ["sc", "swarm"])
    ap.add_argument("--robots", type=int, default=8, help="swarm workload: number of robots R")
    ap.add_argument("--rows", type=int, default=0, help="override the row count (debugging only)")
    ap.add_argument("--dim", type=int, default=0, help="override the descriptor size (debugging only)")
    ap.add_argument("--batch", type=int, default=0, help="override the batch size (debugging only)")
    ap.add_argument("--k", type=int, default=0, help="override k (debugging only)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-torch-bar", action="store_true",
                    help="skip the stock torch-on-CUDA restatement of the reference's search (N=1 only)")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


# -------------------------------------------------------------------------------------------------
# CPU arm (oracle port) — used by cpu_baseline and by --impl reference
# -------------------------------------------------------------------------------------------------

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
    sample_rows = min(rows, max(1024, (256 << 20) // (dim * 4)))      # <= 256 MB of descriptors
    rng = np.random.default_rng(3)
    db = rng.random((sample_rows, dim), dtype=np.float32)
    db /= np.linalg.norm(db, axis=1, keepdims=True)
    nq = batch
    qs = rng.random((nq, dim), dtype=np.float32)
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


def torch_cuda_bar(view, q_block, k):
    """The reference's search arithmetic (cslam/nns_matching.py:57-63: cosine_similarity over the stored
    rows, full argsort, gather, two device->host reads) restated with stock torch ops on the SAME device and
    the same resident database (the zero-copy `.data` view), one query at a time as the reference's API is.
    /root/reference itself cannot travel to the benchmark box; this is PyTorch library code used as a bar,
    never part of the product path.  A few queries are timed and the rate is per query."""
    import torch

    def one(q):
        sim = torch.nn.functional.cosine_similarity(view, q.view(1, -1))
        order = torch.argsort(sim, descending=True)[:k]
        return order.cpu().numpy(), sim[order].cpu().numpy()

    nq, nb_ = 4, int(q_block.shape[0])
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
            "sample": "%d single-query searches (torch %s: cosine_similarity + argsort + gather + 2 D2H reads), "
                      "same device, same resident database" % (nq, torch.__version__)}


def measured_peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return p["hbm_gbs"], p["bf16_tflops"], p.get("bf16_tflops_sustained", p["bf16_tflops"]), "measured"
    except Exception:
        return 6650.0, 1590.0, 1400.0, "fallback"


# -------------------------------------------------------------------------------------------------
def reference_arm(args, rows, dim, batch, k):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    vals = []
    desc, cores = "", 1
    for i in range(args.warmup + args.steps):
        qps, cores, desc = cpu_port_qps(rows, dim, batch, k, budget_s=max(2.0, 60.0 / max(1, args.steps + args.warmup)))
        if i >= args.warmup:
            vals.append(qps)
    v = sum(vals) / len(vals)
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "queries/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": batch / v * 1e3,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": workload_config(args, rows, dim, batch, k),
            "cpu_baseline": {"value": v, "unit": "queries/s", "cores": cores, "kind": "port", "sample": desc},
            "e2e": {"value": v, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def workload_config(args, rows, dim, batch, k):
    return {"workload": WORKLOADS[args.workload][4], "rows": rows, "dim": dim, "batch": batch, "k": k,
            "sharding": "row-striped over %d rank(s), one exchange of B*k keys per rank per step" % args.gpus
            if args.gpus > 1 else "single GPU",
            "l2": "database (%.1f GB) is far larger than the 126 MB L2; no flush needed" % (rows * dim * 4 / 1e9)}


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
    rng = np.random.default_rng(7)
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


def bench_swarm(args):
    """SURVEY §8 d1 C4 in its swarm form + f1/f2: R robots (default 8), the 2M x 256 database of BASELINE
    configs[3] split into R per-robot databases of 2M/R rows.  One step = one local keyframe through the
    drop-in ``LoopClosureSparseMatching.add_local_global_descriptor`` (1 add_item on the local database +
    R-1 search_best on the other robots' databases, threshold test, edge bookkeeping) — host descriptor in,
    host matches out, so the whole leg is end to end.  Timed twice on the same object: through
    ``search_best_multi`` (one staging, R-1 kernels back to back, ONE synchronisation; the default) and
    through the reference's per-robot loop (R-1 synchronised searches)."""
    import numpy as np
    import torch
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
    nq = args.warmup + args.steps
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
        run(fused, 0, args.warmup)
        torch.cuda.synchronize()
        sampler = ClockSampler(0)
        sampler.start()
        l0 = lib.cslam_nns_launch_count()
        t0 = time.perf_counter()
        nm = run(fused, args.warmup, nq)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        out[name] = {"s_per_step": dt / args.steps, "matches": nm, "launches": lib.cslam_nns_launch_count() - l0,
                     "clocks": sampler.stop()}
    assert out["loop"]["matches"] == out["multi"]["matches"]
    lcsm._dense = True
    hbm_peak, _, _, which = measured_peaks()
    dt = out["multi"]["s_per_step"]
    bytes_ = float(R - 1) * per * dim * 4
    cpu = None
    if not args.no_cpu_baseline:
        v, cores, desc = cpu_port_qps(per, dim, 1, 1)
        cpu = {"value": v, "unit": "queries/s", "cores": cores, "kind": "port", "sample": desc}
    qps = (R - 1) / dt
    line = {"metric": METRIC, "value": qps, "unit": "queries/s", "n_gpus": 1, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "BASELINE configs[3] as a swarm: %d robots, %d x %d fp32 per robot database, one "
                                   "step = add_local_global_descriptor (1 add_item + %d search_best + threshold)"
                                   % (R, per, dim, R - 1),
                       "rows": per * R, "dim": dim, "batch": 1, "k": 1, "robots": R,
                       "l2": "the %d databases searched per step total %.2f GB, far larger than the 126 MB L2; no "
                             "flush needed" % (R - 1, bytes_ / 1e9)},
            "roofline": {"bound": "hbm", "achieved": bytes_ / dt / 1e9, "peak": hbm_peak, "unit": "GB/s",
                         "frac": bytes_ / dt / 1e9 / hbm_peak, "traffic": None, "kernel": "k1_single_query",
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
            "matches": out["multi"]["matches"]}
    print(json.dumps(line))


def main():
    args = parse()
    if args.workload == "sc":
        return bench_scancontext(args)
    if args.workload == "swarm":
        if args.impl == "reference":   # the CPU arm of one per-robot database: BASELINE configs[3]'s shape
            args.workload = "c4"
            return reference_arm(args, (args.rows or 2_000_000) // args.robots, args.dim or 256, 1, 1)
        return bench_swarm(args)
    rows, dim, batch, k, _ = WORKLOADS[args.workload]
    if args.rows:
        rows = args.rows
    if args.dim:
        dim = args.dim
    if args.batch:
        batch = args.batch
    if args.k:
        k = args.k
    if args.impl == "reference":
        return reference_arm(args, rows, dim, batch, k)

    import numpy as np
    import torch
    import torch.distributed as dist
    from cslam_b200 import _native
    from cslam_b200.sharded import ShardedNearestNeighborsMatching

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    assert world == args.gpus, "launch with torchrun --nproc-per-node %d" % args.gpus
    lib = _native.load()

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
            probe = x[: 4 * batch].clone()
        m.add_local_shard(x)
        done += c
        del x
    m.finalize_bulk()
    assert m.n == rows
    # queries: same on every rank; every 10th is a noisy copy of a rank-0 row (true neighbours exist)
    gq = torch.Generator(device=dev)
    gq.manual_seed(4)
    nb = 4
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

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # BASELINE configs[3] (c4): "incremental add_item between queries" — every step first appends one new
    # descriptor (device-resident leg: a stream-ordered device-to-device row copy; end-to-end leg: the
    # public add_item with a host row), then searches.  The database grows by one row per step.
    incremental = args.workload == "c4" and world == 1
    extra = None
    if incremental:
        extra = torch.rand((64, dim), generator=gq, device=dev, dtype=torch.float32)
        extra /= extra.norm(dim=1, keepdim=True)

    def device_step(i):
        if incremental:
            m.add_local_shard(extra[i % 64:i % 64 + 1])
        m.search_batch_device(qb[i % nb], k, out_idx, out_sim)

    # ---- device-resident leg: `value` + roofline -----------------------------------------------------
    for i in range(args.warmup):
        device_step(i)
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    lib.cslam_nns_profile(h.h, 1)
    launches0 = lib.cslam_nns_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        device_step(i)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    launches = lib.cslam_nns_launch_count() - launches0
    kms, kl = ctypes.c_float(), ctypes.c_int64()
    lib.cslam_nns_profile_read(h.h, ctypes.byref(kms), ctypes.byref(kl))
    lib.cslam_nns_profile(h.h, 0)
    clocks = sampler.stop()
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    if clocks["sm_mhz"] is None:
        # the timed region was shorter than one NVML poll (50 ms): repeat the same step for ~0.5 s right
        # after it, outside any timing, only to read the clocks under this load (same count on every rank)
        reps = int(min(200000, max(8, 0.5 / max(ms / 1e3 / args.steps, 1e-6))))
        sampler = ClockSampler(local_rank)
        sampler.start()
        time.sleep(0.1)
        for i in range(reps):
            device_step(i)
        barrier()
        clocks = sampler.stop()
        clocks["sampled"] = "%d untimed repeats of the step right after the timed region" % reps
    else:
        clocks["sampled"] = "during the timed region"
    value = args.steps * batch / (ms / 1e3)
    variant = lib.cslam_nns_last_variant(h.h)

    # ---- end-to-end leg: public Python API, host buffers --------------------------------------------
    e2e = None
    if not args.no_e2e:
        hq = [torch.from_numpy(q.cpu().numpy()).pin_memory() for q in qb]

        extra_host = extra.cpu() if incremental else None

        def e2e_step(q_host):
            if incremental:  # the reference's add_item: one host row, item = its row number
                m.local.add_item(extra_host[m.local.n % 64], m.local.n)
            if world == 1:   # host queries in, host (rows, sims) out through cslam_nns_search
                return m.local._search_block(q_host, k)
            q_dev = q_host.to(dev, non_blocking=True)
            i_dev, s_dev = m.search_batch_device(q_dev, k)
            return i_dev.cpu(), s_dev.cpu()

        for i in range(min(args.warmup, 2)):
            e2e_step(hq[i % nb])
        barrier()
        t0 = time.perf_counter()
        for i in range(args.steps):
            e2e_step(hq[i % nb])
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        t = torch.tensor([dt], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e = {"value": args.steps * batch / float(t.item()), "unit": "queries/s",
               "h2d_bytes_per_step": batch * dim * 4 + (dim * 4 if incremental else 0),
               "d2h_bytes_per_step": batch * k * 12}

    # ---- the reference's own GPU arithmetic on the same box (informational bar, SURVEY §8 d3 iii) ------
    torch_bar = None
    if world == 1 and not args.no_torch_bar:
        torch_bar = torch_cuda_bar(m.local.data[:m.local.n], qb[0], k)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel --------------------------------------------------------------
    hbm_peak, tf_burst, tf_sust, which = measured_peaks()
    k_ms = kms.value / max(1, kl.value)
    local_rows = my_rows
    if variant == 2:
        per_launch_q = min(batch, 256)
        flops = 2.0 * local_rows * dim * per_launch_q
        achieved = flops / (k_ms / 1e3) / 1e12
        peak = tf_sust / 2.0
        roof = {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                "frac": achieved / peak, "traffic": None, "kernel": "k2_tcgen05_3xtf32",
                "issued_frac": 3.0 * achieved / peak, "kernel_ms": k_ms, "launches": kl.value,
                "note": "achieved = algorithmic 2*N*D*B flops per launch / CUDA-event duration; peak = "
                        "%s bf16 sustained %.0f TF/s / 2 (tf32 dense is half the bf16 rate); issued_frac "
                        "counts the 3 tf32 products of the 3xTF32 split" % (which, tf_sust)}
    else:
        # the same kernel timed as 200 back-to-back launches inside the library (no Python between launches)
        b2b = ctypes.c_float()
        lib.cslam_nns_time_search(h.h, ctypes.c_void_p(qb[0].data_ptr()), batch, dim, k, 200,
                                  ctypes.c_void_p(torch.cuda.current_stream().cuda_stream), ctypes.byref(b2b))
        bytes_ = float(local_rows) * ((dim + 3) // 4 * 4) * 4
        achieved = bytes_ / (k_ms / 1e3) / 1e9
        roof = {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                "frac": achieved / hbm_peak, "traffic": None, "kernel": "k1_single_query",
                "kernel_ms": k_ms, "launches": kl.value, "kernel_ms_back_to_back": b2b.value / max(1, batch),
                "note": "achieved = N*ld*4 algorithmic bytes per launch / CUDA-event duration; peak = %s "
                        "copy bandwidth" % which}

    # traffic: DRAM bytes per launch of the dominant kernel from the committed ncu --set full capture of this
    # workload (profiles/ncu_traffic.json), null when none exists for it
    try:
        tr = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json"))).get(args.workload)
        if tr and world == 1 and not (args.rows or args.dim or args.batch or args.k):
            roof["traffic"] = tr["dram_bytes_read"] + tr["dram_bytes_write"]
            roof["traffic_source"] = tr["source"]
            if "tensor_pipe_active_pct" in tr:   # sm__pipe_tensor_cycles_active, same capture
                roof["tensor_pipe_active_pct_ncu"] = tr["tensor_pipe_active_pct"]
    except Exception:
        pass

    cpu = None
    if not args.no_cpu_baseline:
        v, cores, desc = cpu_port_qps(rows, dim, batch, k)
        cpu = {"value": v, "unit": "queries/s", "cores": cores, "kind": "port", "sample": desc}

    line = {"metric": METRIC, "value": value, "unit": "queries/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32 (3xTF32 tensor-core products, fp32 accumulate)"
            if variant == 2 else "f32", "data": "synthetic",
            "config": dict(workload_config(args, rows, dim, batch, k), search_variant=variant,
                           **({"step": "1 add_item (one new row) + 1 search; the database grows by one row per step"}
                              if incremental else {}),
                           **({"exchange": "K5 peer-memory kernel (csrc/k5_exchange.cu)"
                               if getattr(m, "_xchg", None) is not None else "NCCL all-gather + K4"} if world > 1 else {})),
            "roofline": roof, "cpu_baseline": cpu, "e2e": e2e, "clocks": clocks, "gpu_launches": int(launches)}
    if torch_bar is not None:
        line["torch_cuda_reference"] = torch_bar
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
