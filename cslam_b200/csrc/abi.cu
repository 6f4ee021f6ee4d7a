// C-ABI of the matcher (include/cslam_nns.h).  Host-side logic only: storage, staging, variant
// dispatch.  The arithmetic is in k1_single.cu / k2_batched.cu / k3_sort.cu / k4_merge.cu.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <cmath>
#include <limits>
#include <string>
#include <vector>

#include "../../include/cslam_nns.h"
#include "arena.h"
#include "kernels.h"

namespace cslam {
static std::atomic<long long> g_launches{0};
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
}  // namespace cslam

using namespace cslam;

namespace {

std::string g_create_error;

int env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return (v && *v) ? atoi(v) : dflt;
}

struct DevBuf {
  void* p = nullptr;
  size_t bytes = 0;
  bool ensure(size_t want) {
    if (want <= bytes) return true;
    if (p) cudaFree(p);
    p = nullptr;
    bytes = 0;
    size_t sz = want < 4096 ? 4096 : want;
    if (cudaMalloc(&p, sz) != cudaSuccess) {
      cudaGetLastError();
      return false;
    }
    bytes = sz;
    return true;
  }
  void free() {
    if (p) cudaFree(p);
    p = nullptr;
    bytes = 0;
  }
};

struct PinBuf {
  void* p = nullptr;
  size_t bytes = 0;
  bool ensure(size_t want) {
    if (want <= bytes) return true;
    if (p) cudaFreeHost(p);
    p = nullptr;
    bytes = 0;
    size_t sz = want < 4096 ? 4096 : want;
    if (cudaHostAlloc(&p, sz, cudaHostAllocMapped | cudaHostAllocPortable) != cudaSuccess) {
      cudaGetLastError();
      return false;
    }
    bytes = sz;
    return true;
  }
  void free() {
    if (p) cudaFreeHost(p);
    p = nullptr;
    bytes = 0;
  }
};

}  // namespace

struct cslam_nns_s {
  int device = 0;
  int num_sms = 148;
  long long dim = 0, ld = 0;
  long long n = 0;
  long long capacity = 0;   // logical rows, reference sequence initial * 2^j
  long long initial = 1000;
  Arena arena;
  std::string err;
  int variant = 0, last_variant = 0;
  K1Tuning k1;
  // device scratch
  DevBuf q, q_lo, inv_qn, scratch, ticket, part_keys, res_dev, all_sims, sort_tmp, sort_kv, qids, shard_keys;
  DevBuf screen_stats;             // [2] u64: queries screened / handed to the 3xTF32 fallback
  bool screen = true;              // screening route enabled (CSLAM_NNS_SCREEN=0 or cslam_nns_set_screen disable it)
  long long screen_min_rows = 8192;  // below this the second and third launch cost more than the 3xTF32 pass saves
  int seed_stride = 0;               // the screening pass is seeded from every seed_stride-th row (0: sized by seed_rows)
  long long seed_rows = 0;           // rows of the seeding sample (0: sized from the database, see search_device)
  long long seed_min_rows = 200000;  // ... when the database has at least this many rows
  long long seed_max_dim = 1024;     // ... and rows this short (the warm-up hides under the products of longer rows)
  PinBuf res_host;
  // integer row identities for the device-side candidate filter (cslam_nns_search_filtered)
  Arena ids_arena;
  std::vector<long long> ids_host;
  long long ids_synced = 0;  // rows [0, ids_synced) of ids_host are in ids_arena
  // optional per-kernel timing (cslam_nns_profile)
  cudaEvent_t add_ev = nullptr;  // completion of an append whose host source is pinned (cslam_nns_add)
  // chained single-query searches of this database (programmatic dependent launch, k1_single.cu)
  unsigned k1_seq = 0;             // ticket / scratch slot of the next launch = k1_seq % kK1Slots
  bool k1_chain = false;           // the last thing this handle queued on k1_stream was a single-query search
  cudaStream_t k1_stream = nullptr;
  bool chain_enabled = true;       // CSLAM_NNS_CHAIN=0 switches the chaining off
  bool data_exposed = false;       // cslam_nns_data_ptr was called: rows may change behind the library's back
  // host-destination fast path: the kernel publishes a sequence number in mapped memory, the host spins on it
  PinBuf done_host;
  unsigned done_seq = 0;
  bool want_done = false;
  bool profile = false;
  std::vector<cudaEvent_t> ev;   // pairs
  size_t ev_used = 0;
};

namespace {

struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    if (prev != dev) cudaSetDevice(dev);
    else prev = -1;
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

int fail(cslam_nns_t h, int code, const std::string& msg) {
  if (h) h->err = msg;
  else g_create_error = msg;
  return code;
}
int fail_cuda(cslam_nns_t h, const char* what, cudaError_t e) {
  cudaGetLastError();
  return fail(h, CSLAM_NNS_ECUDA, std::string(what) + ": " + cudaGetErrorString(e));
}
#define CK(h, call)                                   \
  do {                                                \
    cudaError_t e__ = (call);                         \
    if (e__ != cudaSuccess) return fail_cuda(h, #call, e__); \
  } while (0)

// CSLAM_NNS_DEBUG_SYNC=1: synchronise after every launch of the batched route and name the kernel that faulted
// (compute-sanitizer is not available on every pool; this is the poor man's bisect)
int debug_sync(cslam_nns_t h, cudaStream_t stream, const char* what) {
  static const bool on = env_int("CSLAM_NNS_DEBUG_SYNC", 0) != 0;
  if (!on) return 0;
  cudaError_t e = cudaStreamSynchronize(stream);
  if (e != cudaSuccess) return fail_cuda(h, what, e);
  return 0;
}
#define DBG_SYNC(h, stream, what)                  \
  do {                                             \
    int rc__ = debug_sync(h, stream, what);        \
    if (rc__) return rc__;                         \
  } while (0)

int grow_to(cslam_nns_t h, long long rows, cudaStream_t stream) {
  if (h->dim <= 0) return fail(h, CSLAM_NNS_ESTATE, "descriptor size not set");
  long long cap = h->capacity > 0 ? h->capacity : h->initial;
  while (cap < rows) cap *= 2;  // nns_matching.py:38 doubles
  if (cap == h->capacity) return 0;
  const size_t bytes = (size_t)cap * h->ld * 4;
  std::string err;
  if (!h->arena.ensure(h->device, bytes, stream, &err)) return fail(h, CSLAM_NNS_ENOMEM, err);
  h->capacity = cap;
  return 0;
}

struct KernelTimer {
  cslam_nns_t h;
  cudaStream_t stream;
  cudaEvent_t stop = nullptr;
  KernelTimer(cslam_nns_t h_, cudaStream_t s) : h(h_), stream(s) {
    if (!h->profile) return;
    if (h->ev_used + 2 > h->ev.size()) {
      cudaEvent_t a, b;
      if (cudaEventCreate(&a) != cudaSuccess || cudaEventCreate(&b) != cudaSuccess) return;
      h->ev.push_back(a);
      h->ev.push_back(b);
    }
    cudaEventRecord(h->ev[h->ev_used], stream);
    stop = h->ev[h->ev_used + 1];
    h->ev_used += 2;
  }
  ~KernelTimer() {
    if (stop) cudaEventRecord(stop, stream);
  }
};

__global__ void fill_kernel(long long* idx, float* sim, long long B, int k, int from) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int w = k - from;
  if (i >= B * w) return;
  const long long b = i / w;
  const int j = from + (int)(i - b * w);
  if (idx) idx[b * k + j] = -1;
  if (sim) sim[b * k + j] = -INFINITY;
}

// Candidate filter of the reference's callers, one thread per query.  Reads the top-k
// (rows, sims) the search left in device memory, maps rows to keyframe ids through the device id table
// and applies loop_closure_sparse_matching.py:51,71 (mode 0) or :83-97 (mode 1).  Results go to the
// caller's (mapped pinned) arrays with pitch k; entries past k_run are filled.
__global__ void filter_kernel(const long long* __restrict__ idx, const float* __restrict__ sim,
                              long long B, int k_run, int k, const long long* __restrict__ row_ids,
                              const long long* __restrict__ qids, long long qid0, int has_qid, int mode,
                              long long min_gap, float thr, long long* o_idx, float* o_sim, int* o_pos) {
  const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  const long long qid = qids ? qids[b] : qid0;
  int pos = -1;
  for (int j = 0; j < k_run; ++j) {
    const long long r = idx[b * k_run + j];
    const float s = sim[b * k_run + j];
    o_idx[b * k + j] = r;
    o_sim[b * k + j] = s;
    if (pos >= 0 || r < 0) continue;
    if (mode == CSLAM_NNS_FILTER_BEST) {
      if (j == 0 && s >= thr) pos = 0;  // :51,:71 (a NaN similarity does not pass `>=`)
      continue;
    }
    if (has_qid) {
      const long long id = row_ids[r];
      if (j == 0 && id == qid) continue;  // :83-84 kfs[0] == kf_id -> dropped
      long long gap = id - qid;
      if (gap < 0) gap = -gap;
      if (gap < min_gap) continue;        // :89-91
    }
    if (s < thr) continue;                // :93-94 (a NaN similarity is not skipped by `<`)
    pos = j;
  }
  for (int j = k_run; j < k; ++j) {
    o_idx[b * k + j] = -1;
    o_sim[b * k + j] = -INFINITY;
  }
  o_pos[b] = pos;
}

// Bring the device id table up to date with ids_host (rows [ids_synced, n)).
int sync_row_ids(cslam_nns_t h, cudaStream_t stream) {
  if ((long long)h->ids_host.size() < h->n)
    return fail(h, CSLAM_NNS_ESTATE, "row ids missing: cslam_nns_set_row_ids must cover every row");
  if (h->ids_synced >= h->n) return 0;
  std::string err;
  if (!h->ids_arena.ensure(h->device, (size_t)h->capacity * 8, stream, &err))
    return fail(h, CSLAM_NNS_ENOMEM, err);
  long long* dst = (long long*)h->ids_arena.base();
  CK(h, cudaMemcpyAsync(dst + h->ids_synced, h->ids_host.data() + h->ids_synced,
                        (size_t)(h->n - h->ids_synced) * 8, cudaMemcpyHostToDevice, stream));
  h->ids_synced = h->n;
  return 0;
}

// Run the search for B device-resident queries (pitch q_ld) and write k_run entries per query to
// device-accessible out pointers (pitch k_run).  Any of the three outputs may be null.
int search_device(cslam_nns_t h, const float* dq, long long B, long long q_ld, int k_run,
                  u64* o_keys, long long* o_idx, float* o_sim, long long idx_stride,
                  long long idx_offset, cudaStream_t stream, bool overlap_prev = false) {
  int variant = h->variant;
  if (variant == 0) {
    if (k_run > kK1MaxK) variant = 3;
    else if (B >= 2 && k2_supported((int)(B > kK2MaxBatch ? kK2MaxBatch : B), k_run, h->dim)) variant = 2;
    else variant = 1;
  }
  if (variant == 2 && !k2_supported((int)(B > kK2MaxBatch ? kK2MaxBatch : B), k_run, h->dim))
    return fail(h, CSLAM_NNS_EINVAL, "k does not fit the batched kernel's fused select");
  if (variant == 1 && k_run > kK1MaxK)
    return fail(h, CSLAM_NNS_EINVAL, "k exceeds the fused single-query select (128)");
  h->last_variant = variant;

  if (!h->ticket.p) {
    if (!h->ticket.ensure(64)) return fail(h, CSLAM_NNS_ENOMEM, "scratch allocation failed");
    CK(h, cudaMemsetAsync(h->ticket.p, 0, 64, stream));  // the kernels leave it at zero again
  }
  const size_t scratch_slot = (size_t)k1_max_grid(h->num_sms) * kK1MaxK;   // keys per ticket / scratch slot
  if (!h->scratch.ensure(scratch_slot * kK1Slots * 8))
    return fail(h, CSLAM_NNS_ENOMEM, "scratch allocation failed");
  if (variant != 1) h->k1_chain = false;

  if (variant == 1) {
    for (long long b = 0; b < B; ++b) {
      // Every launch takes the next ticket / scratch slot of a ring of kK1Slots, so that a search chained behind
      // the previous search of this database (same stream, nothing of ours queued in between, rows not exposed
      // through cslam_nns_data_ptr) may be scheduled while that one is still finishing.
      const unsigned slot = h->k1_seq++ % kK1Slots;
      const bool same_db = h->k1_chain && h->k1_stream == stream && h->chain_enabled && !h->data_exposed &&
                           !h->profile;
      K1Args a;
      a.db = (const float*)h->arena.base();
      a.n = h->n;
      a.ld = (int)h->ld;
      a.dim = (int)h->dim;
      a.q = dq + b * q_ld;
      a.k = k_run;
      a.scratch = (u64*)h->scratch.p + scratch_slot * slot;
      a.ticket = (unsigned*)h->ticket.p + 4 * slot;   // 16 bytes per slot: ticket word, pad, best key
      a.out_keys = o_keys ? o_keys + b * k_run : nullptr;
      a.out_idx = o_idx ? o_idx + b * k_run : nullptr;
      a.out_sim = o_sim ? o_sim + b * k_run : nullptr;
      a.idx_stride = idx_stride;
      a.idx_offset = idx_offset;
      a.all_sims = nullptr;
      const bool other_db = overlap_prev && B == 1 && !h->profile;   // cslam_nns_search_multi: independent databases
      a.overlap_prev = other_db || same_db;
      a.dep_wait = !other_db && same_db;
      if (h->want_done && b == B - 1 && k_run <= 32) {
        a.done_flag = (unsigned*)h->done_host.p;
        a.done_value = ++h->done_seq;
      }
      KernelTimer kt(h, stream);
      CK(h, k1_launch(a, h->k1, h->num_sms, stream));
      h->k1_chain = true;
      h->k1_stream = stream;
    }
    return 0;
  }
  if (variant == 3) {
    const long long n = h->n;
    if (!h->all_sims.ensure((size_t)n * 4) || !h->sort_kv.ensure((size_t)n * 16))
      return fail(h, CSLAM_NNS_ENOMEM, "two-pass buffers");
    const size_t tb = sort_temp_bytes(n);
    if (!h->sort_tmp.ensure(tb)) return fail(h, CSLAM_NNS_ENOMEM, "sort temp");
    uint32_t* kv = (uint32_t*)h->sort_kv.p;
    for (long long b = 0; b < B; ++b) {
      K1Args a;
      a.db = (const float*)h->arena.base();
      a.n = n;
      a.ld = (int)h->ld;
      a.dim = (int)h->dim;
      a.q = dq + b * q_ld;
      a.k = 1;
      a.scratch = (u64*)h->scratch.p;
      a.ticket = (unsigned*)h->ticket.p;
      a.out_keys = nullptr;
      a.out_idx = nullptr;
      a.out_sim = nullptr;
      a.idx_stride = 1;
      a.idx_offset = 0;
      a.all_sims = (float*)h->all_sims.p;
      CK(h, k1_launch(a, h->k1, h->num_sms, stream));
      CK(h, sort_select((const float*)h->all_sims.p, n, k_run, h->sort_tmp.p, tb, kv, kv + n,
                        kv + 2 * n, kv + 3 * n, o_keys ? o_keys + b * k_run : nullptr,
                        o_idx ? o_idx + b * k_run : nullptr, o_sim ? o_sim + b * k_run : nullptr,
                        idx_stride, idx_offset, stream));
    }
    return 0;
  }
  // variant 2: tcgen05 batched kernel, <= 256 queries per pass
  const DriverApi& d = driver_api();
  if (!d.tensorMapEncodeTiled) return fail(h, CSLAM_NNS_ECUDA, "cuTensorMapEncodeTiled unavailable");
  for (long long b0 = 0; b0 < B; b0 += kK2MaxBatch) {
    const int Bc = (int)((B - b0) > kK2MaxBatch ? kK2MaxBatch : (B - b0));
    const int grid = k2_parts(h->n, Bc, h->dim, h->num_sms);
    const bool screen = h->screen && h->n >= h->screen_min_rows && k2_screen_supported(Bc, k_run, (int)h->dim);
    const int kl = screen ? k2_screen_list_len(k_run) : k_run;   // list capacity of the first pass
    if (!h->q_lo.ensure((size_t)2 * Bc * q_ld * 4) || !h->inv_qn.ensure(k2_qpar_bytes(Bc)) ||
        !h->part_keys.ensure((size_t)grid * Bc * kl * 8))
      return fail(h, CSLAM_NNS_ENOMEM, "batched scratch");
    const float* qc = dq + b0 * q_ld;
    CK(h, k2_prepare(qc, Bc, (int)h->dim, (int)q_ld, (float*)h->q_lo.p, (float*)h->inv_qn.p, stream));
    DBG_SYNC(h, stream, "after k2_prepare");
    K2Args a;
    a.db = (const float*)h->arena.base();
    a.n = h->n;
    a.ld = (int)h->ld;
    a.dim = (int)h->dim;
    a.q_hi = qc;
    a.q_lo = (const float*)h->q_lo.p;
    a.inv_qn = (const float*)h->inv_qn.p;
    a.q_ld = (int)q_ld;
    a.B = Bc;
    a.k = k_run;
    a.part_keys = (u64*)h->part_keys.p;
    a.encode_fn = d.tensorMapEncodeTiled;
    u64* ok_ = o_keys ? o_keys + b0 * k_run : nullptr;
    long long* oi_ = o_idx ? o_idx + b0 * k_run : nullptr;
    float* os_ = o_sim ? o_sim + b0 * k_run : nullptr;
    const char* kerr = nullptr;
    cudaError_t e;
    // Screening route (large batches, big enough databases): ONE tf32 product over the whole database, exact
    // fp32 re-scoring of the few candidates per query with a completeness proof (k6_rescore.cu), and the 3xTF32
    // pass only for the queries the proof does not cover — queued unconditionally, it exits at once when there
    // are none, so the call stays stream-ordered without a host round trip.
    if (screen) {
      if (!h->screen_stats.p) {
        if (!h->screen_stats.ensure(64)) return fail(h, CSLAM_NNS_ENOMEM, "screening counters");
        CK(h, cudaMemsetAsync(h->screen_stats.p, 0, 64, stream));
      }
      a.screen = true;
      a.k = kl;
      // Seed the pass with a bound from a SAMPLE: the same kernel over every S-th row (a tensor map with S times
      // the row pitch) and a select of the kl-th largest screened similarity per query.  Without it every one of
      // the 2 * 74 lists of a query warms up on its own, ~kl * ln(rows per list / kl) insertions, each a lane-
      // divergent round in the epilogue: ~0.3 ms per launch whatever the database size, i.e. most of a shard's
      // kernel on 8 GPUs.  With the bound of a 1/S sample only the ~S * kl best rows of a query ever enter a list.
      // The sample is small (seed_rows rows in all) so that its own warm-up stays cheap.  Short rows only: with
      // long rows (dim > 1024) a tile's products take far longer than its epilogue and the warm-up is hidden.
      // Sample size: 1-8 256-row tiles per CTA pair (18 944 rows each), one per 312 500 rows of 256 floats — a
      // tile of the sample costs ~2.6 us at 256 floats (its epilogue is a running maximum) and a 4x larger sample
      // leaves a quarter of the rows above the seed for the full pass's insertion rounds (1.25M x 256: kernel
      // 0.304 -> 0.265 ms with four tiles); longer rows get proportionally fewer tiles.
      long long tiles_per_pair = h->n / (312500ll * ((h->dim + 255) / 256));
      tiles_per_pair = tiles_per_pair < 1 ? 1 : (tiles_per_pair > 8 ? 8 : tiles_per_pair);
      const long long sample_rows = h->seed_rows > 0 ? h->seed_rows : 18944 * tiles_per_pair;
      long long stride = h->seed_stride > 0 ? h->seed_stride : h->n / sample_rows;
      if (stride > 1 && h->n >= h->seed_min_rows && h->dim <= h->seed_max_dim) {
        // The sample pass keeps ONE entry per list — the best row of the 128 columns a thread sees — and shares no
        // bound between lists: with nothing to start from, every column of a list's first tile is a candidate until
        // the list is full, and one convergent insertion round per column made the sample pass cost 52 us of
        // epilogue for 2.6 us of products (lists of 8; role profile, 1.25M x 256).  The 2 * pairs maxima belong to
        // distinct rows, so their kl-th largest is reached by kl real rows: a valid seed, and nearly as tight as the
        // kl-th largest of the whole sample (its best kl rows mostly fall into different lists).
        const int kls = 1;
        K2Args sa = a;
        sa.k = kls;
        sa.n = (h->n + stride - 1) / stride;
        sa.ld = a.ld * (int)stride;
        sa.gthr_plane = -1;              // lists of the sample pass do not publish or read a shared bound
        e = k2_launch(sa, h->num_sms, stream, &kerr);
        if (e != cudaSuccess) return fail_cuda(h, kerr ? kerr : "k2_launch (sample)", e);
        DBG_SYNC(h, stream, "after the sample pass");
        K6Args sr;
        sr.db = a.db;
        sr.n = sa.n;
        sr.ld = sa.ld;
        sr.dim = a.dim;
        sr.q = qc;
        sr.q_ld = (int)q_ld;
        sr.qpar = a.inv_qn;
        sr.part_keys = a.part_keys;
        sr.parts = k2_parts(sa.n, Bc, h->dim, h->num_sms);
        sr.B = Bc;
        sr.k = kl;        // rank of the seed: the kl-th largest screened similarity of the sample
        sr.kl = kls;      // capacity of the sample pass's lists
        sr.out_keys = nullptr;
        sr.out_idx = nullptr;
        sr.out_sim = nullptr;
        sr.idx_stride = 1;
        sr.idx_offset = 0;
        sr.stats = nullptr;
        sr.seed = true;
        CK(h, k6_rescore(sr, stream));
        DBG_SYNC(h, stream, "after the seed kernel");
      }
      {
        KernelTimer kt(h, stream);
        e = k2_launch(a, h->num_sms, stream, &kerr);
      }
      if (e != cudaSuccess) return fail_cuda(h, kerr ? kerr : "k2_launch (screening)", e);
      DBG_SYNC(h, stream, "after the screening kernel");
      K6Args r;
      r.db = a.db;
      r.n = a.n;
      r.ld = a.ld;
      r.dim = a.dim;
      r.q = qc;
      r.q_ld = (int)q_ld;
      r.qpar = a.inv_qn;
      r.part_keys = a.part_keys;
      r.parts = grid;
      r.B = Bc;
      r.k = k_run;
      r.kl = kl;
      r.out_keys = ok_;
      r.out_idx = oi_;
      r.out_sim = os_;
      r.idx_stride = idx_stride;
      r.idx_offset = idx_offset;
      r.stats = (unsigned long long*)h->screen_stats.p;
      if (env_int("CSLAM_NNS_K6_TRACE", 0) && h->sort_kv.ensure((size_t)Bc * 64 * 8)) r.debug = (u64*)h->sort_kv.p;
      CK(h, k6_rescore(r, stream));
      DBG_SYNC(h, stream, "after k6_rescore");
      const unsigned* tail = (const unsigned*)h->inv_qn.p + (size_t)kQParPlanes * Bc;
      if (env_int("CSLAM_NNS_SCREEN_NOFB", 0)) continue;   // development: leave the screening pass's lists in place
      a.screen = false;
      a.k = k_run;
      a.gthr_plane = kQParGthrFull;
      a.active_mask = tail;
      a.active_count = tail + 8;
      e = k2_launch(a, h->num_sms, stream, &kerr);
      if (e != cudaSuccess) return fail_cuda(h, kerr ? kerr : "k2_launch (fallback)", e);
      DBG_SYNC(h, stream, "after the fallback kernel");
      CK(h, k4_merge((const u64*)h->part_keys.p, grid, Bc, k_run, k_run, ok_, oi_, os_, idx_stride, idx_offset,
                     stream, (const unsigned*)h->inv_qn.p + (size_t)kQParGthrFull * Bc, tail));
      DBG_SYNC(h, stream, "after the fallback merge");
      continue;
    }
    {
      KernelTimer kt(h, stream);
      e = k2_launch(a, h->num_sms, stream, &kerr);
    }
    if (e != cudaSuccess) return fail_cuda(h, kerr ? kerr : "k2_launch", e);
    CK(h, k4_merge((const u64*)h->part_keys.p, grid, Bc, k_run, k_run, ok_, oi_, os_, idx_stride, idx_offset, stream,
                   k2_floor_plane((const float*)h->inv_qn.p, Bc, h->dim)));
  }
  return 0;
}

// Host side of the sequence-word handshake (k1_single.cu: done_flag): spin on the mapped word the last CTA writes
// after the results, instead of a stream synchronisation (~8 us cheaper per synchronous single-query search).
// The stream is polled now and then so that a faulted kernel ends the wait with its error.
cudaError_t wait_done(cslam_nns_t h, unsigned want, cudaStream_t stream) {
  volatile unsigned* flag = (volatile unsigned*)h->done_host.p;
  for (unsigned it = 1;; ++it) {
    if (*flag == want) return cudaSuccess;
    if ((it & 0x3fff) == 0) {
      const cudaError_t e = cudaStreamQuery(stream);
      if (e != cudaErrorNotReady) return e;   // finished (flag or not), or failed
      cudaGetLastError();
    }
#if defined(__x86_64__)
    __builtin_ia32_pause();
#endif
  }
}

// Bring the queries to device memory with a 16-byte aligned pitch (multiple of 4 floats).
int stage_queries(cslam_nns_t h, const float* queries, long long B, long long q_ld, int src_is_device,
                  cudaStream_t stream, const float** dq, long long* dq_ld) {
  const long long dim = h->dim;
  if (q_ld <= 0) q_ld = dim;
  if (q_ld < dim) return fail(h, CSLAM_NNS_EINVAL, "query pitch smaller than dim");
  if (src_is_device && (q_ld % 4 == 0) && ((uintptr_t)queries % 16 == 0)) {
    *dq = queries;
    *dq_ld = q_ld;
    return 0;
  }
  const long long pld = h->ld;
  if (!h->q.ensure((size_t)B * pld * 4)) return fail(h, CSLAM_NNS_ENOMEM, "query staging");
  if (pld != dim) CK(h, cudaMemsetAsync(h->q.p, 0, (size_t)B * pld * 4, stream));
  CK(h, cudaMemcpy2DAsync(h->q.p, (size_t)pld * 4, queries, (size_t)q_ld * 4, (size_t)dim * 4,
                          (size_t)B, src_is_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice,
                          stream));
  h->k1_chain = false;  // the staging copy is an ordinary stream operation: the search starts after it, unchained
  *dq = (const float*)h->q.p;
  *dq_ld = pld;
  return 0;
}

}  // namespace

#pragma GCC visibility push(default)
extern "C" {

int cslam_nns_abi_version(void) { return 1; }
int64_t cslam_nns_launch_count(void) { return g_launches.load(); }

const char* cslam_nns_last_error(cslam_nns_t h) { return h ? h->err.c_str() : g_create_error.c_str(); }

int cslam_nns_create(int device, int64_t dim, int64_t initial_capacity, cslam_nns_t* out) {
  if (!out) return fail(nullptr, CSLAM_NNS_EINVAL, "out is NULL");
  *out = nullptr;
  if (dim < 0) return fail(nullptr, CSLAM_NNS_EINVAL, "dim must be >= 0");
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0) {
    cudaGetLastError();
    return fail(nullptr, CSLAM_NNS_ENODEV,
                std::string("no CUDA device available (") +
                    (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0") +
                    "); this library has no CPU fallback");
  }
  if (device < 0 || device >= count) return fail(nullptr, CSLAM_NNS_EINVAL, "bad device ordinal");
  int major = 0, sms = 0;
  cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
  if (major != 10)
    return fail(nullptr, CSLAM_NNS_ENODEV,
                "device is not compute capability 10.x; kernels are built for sm_100a only");
  DeviceGuard g(device);
  cudaFree(0);
  cslam_nns_t h = new cslam_nns_s();
  h->device = device;
  h->num_sms = sms;
  h->initial = initial_capacity > 0 ? initial_capacity : 1000;
  h->k1.stage_bytes = env_int("CSLAM_K1_STAGE_BYTES", 8192);
  h->k1.stages = env_int("CSLAM_K1_STAGES", 2);  // 8 KB x 2 per warp: best of the B200 sweep (profiles/r1_k1_sweep_v2.log)
  h->k1.ctas = env_int("CSLAM_K1_CTAS", 0);
  h->variant = env_int("CSLAM_NNS_VARIANT", 0);
  h->chain_enabled = env_int("CSLAM_NNS_CHAIN", 1) != 0;
  h->screen = env_int("CSLAM_NNS_SCREEN", 1) != 0;
  h->screen_min_rows = env_int("CSLAM_NNS_SCREEN_MIN_ROWS", 8192);
  h->seed_stride = env_int("CSLAM_NNS_SEED_STRIDE", 0);
  h->seed_rows = env_int("CSLAM_NNS_SEED_ROWS", 0);
  h->seed_max_dim = env_int("CSLAM_NNS_SEED_MAX_DIM", 1024);
  h->seed_min_rows = env_int("CSLAM_NNS_SEED_MIN_ROWS", 200000);
  *out = h;
  if (dim > 0) {
    int rc = cslam_nns_set_dim(h, dim);
    if (rc == 0) rc = grow_to(h, h->initial, 0);
    // the zero-fill of the new chunk was queued on the legacy stream: finish it here, so that an add or a
    // search on any (possibly non-blocking) stream can never be overtaken by it
    if (rc == 0 && cudaStreamSynchronize(0) != cudaSuccess) rc = fail_cuda(h, "cudaStreamSynchronize", cudaGetLastError());
    if (rc != 0) {
      g_create_error = h->err;
      delete h;
      *out = nullptr;
      return rc;
    }
  }
  return 0;
}

int cslam_nns_destroy(cslam_nns_t h) {
  if (!h) return 0;
  {
    DeviceGuard g(h->device);
    cudaDeviceSynchronize();
    h->arena.release();
    h->ids_arena.release();
    DevBuf* bufs[] = {&h->q,         &h->q_lo,    &h->inv_qn,   &h->scratch,  &h->ticket, &h->qids, &h->shard_keys, &h->screen_stats,
                      &h->part_keys, &h->res_dev, &h->all_sims, &h->sort_tmp, &h->sort_kv};
    for (DevBuf* b : bufs) b->free();
    h->res_host.free();
    h->done_host.free();
    for (cudaEvent_t e : h->ev) cudaEventDestroy(e);
    if (h->add_ev) cudaEventDestroy(h->add_ev);
  }
  delete h;
  return 0;
}

int cslam_nns_set_dim(cslam_nns_t h, int64_t dim) {
  if (!h) return CSLAM_NNS_EINVAL;
  if (dim <= 0) return fail(h, CSLAM_NNS_EINVAL, "dim must be positive");
  if (h->dim > 0) return h->dim == dim ? 0 : fail(h, CSLAM_NNS_ESTATE, "dim already set");
  h->dim = dim;
  h->ld = (dim + 3) / 4 * 4;
  return 0;
}

int cslam_nns_reserve(cslam_nns_t h, int64_t rows) {
  if (!h) return CSLAM_NNS_EINVAL;
  DeviceGuard g(h->device);
  int rc = grow_to(h, rows, 0);
  if (rc == 0) CK(h, cudaStreamSynchronize(0));  // see cslam_nns_create: the zero-fill must not trail behind
  return rc;
}

int64_t cslam_nns_size(cslam_nns_t h) { return h ? h->n : -1; }
int64_t cslam_nns_capacity(cslam_nns_t h) { return h ? h->capacity : -1; }
int64_t cslam_nns_dim(cslam_nns_t h) { return h ? h->dim : -1; }
int64_t cslam_nns_ld(cslam_nns_t h) { return h ? h->ld : -1; }
void* cslam_nns_data_ptr(cslam_nns_t h) {
  if (!h) return nullptr;
  h->data_exposed = true;  // rows may now change without the library knowing: searches are no longer chained
  return h->arena.base();
}
int cslam_nns_device(cslam_nns_t h) { return h ? h->device : -1; }

int cslam_nns_add(cslam_nns_t h, const float* rows, int64_t count, int64_t src_ld, int src_is_device,
                  void* stream_, int64_t* new_n) {
  if (!h) return CSLAM_NNS_EINVAL;
  if (count < 0 || (count > 0 && !rows)) return fail(h, CSLAM_NNS_EINVAL, "bad rows/count");
  if (h->dim <= 0) return fail(h, CSLAM_NNS_ESTATE, "descriptor size not set");
  if (src_ld <= 0) src_ld = h->dim;
  if (src_ld < h->dim) return fail(h, CSLAM_NNS_EINVAL, "source pitch smaller than dim");
  cudaStream_t stream = (cudaStream_t)stream_;
  DeviceGuard g(h->device);
  if (count > 0) {
    h->k1_chain = false;  // the copy is an ordinary stream operation: the next search starts after it
    if (h->n + count > 0xffffffffll) return fail(h, CSLAM_NNS_EINVAL, "more than 2^32 rows");
    int rc = grow_to(h, h->n + count, stream);
    if (rc) return rc;
    float* dst = (float*)h->arena.base() + (size_t)h->n * h->ld;
    // plain DMA: the pad lanes [dim, ld) were zeroed when the chunk was mapped and are never written
    CK(h, cudaMemcpy2DAsync(dst, (size_t)h->ld * 4, rows, (size_t)src_ld * 4, (size_t)h->dim * 4,
                            (size_t)count,
                            src_is_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, stream));
    h->n += count;
    if (!src_is_device) {
      // Contract (include/cslam_nns.h): a HOST source may be reused as soon as the call returns, like the
      // reference's synchronous `.to(device)` (nns_matching.py:41).  The driver stages a pageable source
      // before the copy call returns, but a PINNED (page-locked / registered) source is read by the copy
      // engine later, asynchronously: wait for that copy — and only for it — before returning.
      cudaPointerAttributes at;
      const cudaError_t pe = cudaPointerGetAttributes(&at, rows);
      if (pe != cudaSuccess) cudaGetLastError();
      if (pe != cudaSuccess || at.type != cudaMemoryTypeUnregistered) {
        if (!h->add_ev) CK(h, cudaEventCreateWithFlags(&h->add_ev, cudaEventDisableTiming));
        CK(h, cudaEventRecord(h->add_ev, stream));
        CK(h, cudaEventSynchronize(h->add_ev));
      }
    }
  }
  if (new_n) *new_n = h->n;
  return 0;
}

int cslam_nns_search(cslam_nns_t h, const float* queries, int64_t B, int64_t q_ld, int src_is_device,
                     int k, int64_t* out_idx, float* out_sim, int dst_is_device, void* stream_,
                     int64_t* out_kk) {
  if (!h) return CSLAM_NNS_EINVAL;
  if (B < 0 || k < 0 || (B > 0 && !queries)) return fail(h, CSLAM_NNS_EINVAL, "bad queries/B/k");
  if (h->dim <= 0) return fail(h, CSLAM_NNS_ESTATE, "descriptor size not set");
  cudaStream_t stream = (cudaStream_t)stream_;
  DeviceGuard g(h->device);
  const long long kk = k < h->n ? k : h->n;
  if (out_kk) *out_kk = kk;
  if (B == 0 || k == 0) return 0;

  if (kk == 0) {  // empty DB: nothing but fill
    if (dst_is_device) {
      const long long tot = B * k;
      fill_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, stream>>>((long long*)out_idx, out_sim, B, k, 0);
      count_launch();
      CK(h, cudaGetLastError());
    } else {
      for (long long i = 0; i < B * k; ++i) {
        if (out_idx) out_idx[i] = -1;
        if (out_sim) out_sim[i] = -std::numeric_limits<float>::infinity();
      }
    }
    return 0;
  }
  const int k_run = (int)kk;
  const float* dq;
  long long dq_ld;
  int rc = stage_queries(h, queries, B, q_ld, src_is_device, stream, &dq, &dq_ld);
  if (rc) return rc;

  if (dst_is_device) {
    if (k_run == k) return search_device(h, dq, B, dq_ld, k_run, nullptr, (long long*)out_idx, out_sim, 1, 0, stream);
    if (!h->res_dev.ensure((size_t)B * k_run * 12)) return fail(h, CSLAM_NNS_ENOMEM, "result buffer");
    long long* di = (long long*)h->res_dev.p;
    float* ds = (float*)(di + B * k_run);
    rc = search_device(h, dq, B, dq_ld, k_run, nullptr, di, ds, 1, 0, stream);
    if (rc) return rc;
    if (out_idx)
      CK(h, cudaMemcpy2DAsync(out_idx, (size_t)k * 8, di, (size_t)k_run * 8, (size_t)k_run * 8, B,
                              cudaMemcpyDeviceToDevice, stream));
    if (out_sim)
      CK(h, cudaMemcpy2DAsync(out_sim, (size_t)k * 4, ds, (size_t)k_run * 4, (size_t)k_run * 4, B,
                              cudaMemcpyDeviceToDevice, stream));
    const long long tot = B * (k - k_run);
    fill_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, stream>>>((long long*)out_idx, out_sim, B, k, k_run);
    count_launch();
    CK(h, cudaGetLastError());
    return 0;
  }

  // host destination: kernels write into mapped pinned memory, one stream sync, then a host copy
  if (!h->res_host.ensure((size_t)B * k_run * 12)) return fail(h, CSLAM_NNS_ENOMEM, "pinned result buffer");
  long long* hi = (long long*)h->res_host.p;
  float* hs = (float*)(hi + B * k_run);
  // one query, fused select: the kernel's last CTA publishes a sequence number behind the results and the host
  // spins on it (mapped pinned memory) instead of synchronising the stream
  const unsigned seq0 = h->done_seq;
  if (B == 1 && k_run <= 32 && !h->done_host.p && h->done_host.ensure(64))
    memset(h->done_host.p, 0, 64);   // a recycled pinned block may still hold another handle's last sequence number
  h->want_done = B == 1 && k_run <= 32 && h->done_host.p != nullptr;
  rc = search_device(h, dq, B, dq_ld, k_run, nullptr, hi, hs, 1, 0, stream);
  h->want_done = false;
  if (rc) return rc;
  if (h->done_seq != seq0) CK(h, wait_done(h, h->done_seq, stream));
  else CK(h, cudaStreamSynchronize(stream));
  for (long long b = 0; b < B; ++b) {
    if (out_idx) {
      memcpy(out_idx + b * k, hi + b * k_run, (size_t)k_run * 8);
      for (int j = k_run; j < k; ++j) out_idx[b * k + j] = -1;
    }
    if (out_sim) {
      memcpy(out_sim + b * k, hs + b * k_run, (size_t)k_run * 4);
      for (int j = k_run; j < k; ++j) out_sim[b * k + j] = -std::numeric_limits<float>::infinity();
    }
  }
  return 0;
}

int cslam_nns_search_multi(const cslam_nns_t* hs, int count, const float* query, int64_t q_len,
                           int src_is_device, int k, int64_t* out_idx, float* out_sim, int64_t* out_kk,
                           void* stream_) {
  if (!hs || count <= 0) {
    g_create_error = "cslam_nns_search_multi: no handles";
    return CSLAM_NNS_EINVAL;
  }
  cslam_nns_t h0 = nullptr;  // first searchable database: owns the staging and result buffers
  for (int i = 0; i < count; ++i) {
    cslam_nns_t h = hs[i];
    if (!h || h->dim <= 0 || h->n == 0) continue;
    if (!h0) h0 = h;
    else if (h->device != h0->device || h->dim != h0->dim)
      return fail(h0, CSLAM_NNS_EINVAL, "cslam_nns_search_multi: databases differ in device or dim");
  }
  const float ninf = -std::numeric_limits<float>::infinity();
  if (k < 0 || (h0 && !query)) return fail(h0, CSLAM_NNS_EINVAL, "bad query/k");
  for (long long i = 0; i < (long long)count * k; ++i) {
    if (out_idx) out_idx[i] = -1;
    if (out_sim) out_sim[i] = ninf;
  }
  if (out_kk) for (int i = 0; i < count; ++i) out_kk[i] = 0;
  if (!h0 || k == 0) return 0;
  if (q_len < h0->dim) return fail(h0, CSLAM_NNS_EINVAL, "query shorter than dim");
  cudaStream_t stream = (cudaStream_t)stream_;
  DeviceGuard g(h0->device);
  const float* dq;
  long long dq_ld;
  int rc = stage_queries(h0, query, 1, q_len, src_is_device, stream, &dq, &dq_ld);
  if (rc) return rc;
  if (!h0->res_host.ensure((size_t)count * k * 12)) return fail(h0, CSLAM_NNS_ENOMEM, "pinned result buffer");
  long long* hi = (long long*)h0->res_host.p;
  float* hsim = (float*)(hi + (size_t)count * k);
  std::vector<int> kks(count, 0);
  cslam_nns_t prev_h = nullptr;
  for (int i = 0; i < count; ++i) {
    cslam_nns_t h = hs[i];
    if (!h || h->dim <= 0 || h->n == 0) continue;
    kks[i] = (int)(k < h->n ? k : h->n);
    // every database after the first may start on the SMs the previous database's kernel has vacated
    // (programmatic stream serialisation): the searches share the query and nothing else.  Not for a
    // handle that appears twice (shared scratch) or one whose scratch is still to be set up (a memset
    // would sit between the two launches anyway).
    bool overlap = prev_h != nullptr && h->ticket.p != nullptr && env_int("CSLAM_NNS_MULTI_OVERLAP", 1) != 0;
    for (int j = 0; j < i && overlap; ++j) overlap = hs[j] != h;
    rc = search_device(h, dq, 1, dq_ld, kks[i], nullptr, hi + (size_t)i * k, hsim + (size_t)i * k, 1, 0, stream,
                       overlap);
    prev_h = h;
    if (rc) {
      if (h != h0) h0->err = h->err;
      cudaStreamSynchronize(stream);
      return rc;
    }
  }
  CK(h0, cudaStreamSynchronize(stream));
  for (int i = 0; i < count; ++i) {
    if (out_idx) memcpy(out_idx + (size_t)i * k, hi + (size_t)i * k, (size_t)kks[i] * 8);
    if (out_sim) memcpy(out_sim + (size_t)i * k, hsim + (size_t)i * k, (size_t)kks[i] * 4);
    if (out_kk) out_kk[i] = kks[i];
  }
  return 0;
}

int cslam_nns_set_row_ids(cslam_nns_t h, int64_t first_row, int64_t count, const int64_t* ids) {
  if (!h) return CSLAM_NNS_EINVAL;
  if (first_row < 0 || count < 0 || (count > 0 && !ids)) return fail(h, CSLAM_NNS_EINVAL, "bad ids/range");
  if (first_row > (long long)h->ids_host.size()) return fail(h, CSLAM_NNS_EINVAL, "row ids would leave a gap");
  if ((long long)h->ids_host.size() < first_row + count) h->ids_host.resize((size_t)(first_row + count));
  for (long long i = 0; i < count; ++i) h->ids_host[(size_t)(first_row + i)] = ids[i];
  if (count > 0 && first_row < h->ids_synced) h->ids_synced = first_row;  // overwritten: upload again
  return 0;
}

int cslam_nns_search_filtered(cslam_nns_t h, const float* queries, int64_t B, int64_t q_ld,
                              int src_is_device, int k, int mode, const int64_t* query_ids,
                              int64_t min_gap, float threshold, int64_t* out_idx, float* out_sim,
                              int32_t* out_match_pos, void* stream_, int64_t* out_kk) {
  if (!h) return CSLAM_NNS_EINVAL;
  if (B < 0 || k < 0 || (B > 0 && !queries)) return fail(h, CSLAM_NNS_EINVAL, "bad queries/B/k");
  if (mode != CSLAM_NNS_FILTER_BEST && mode != CSLAM_NNS_FILTER_INTRA)
    return fail(h, CSLAM_NNS_EINVAL, "unknown filter mode");
  if (B > 0 && (!out_match_pos || (k > 0 && (!out_idx || !out_sim))))
    return fail(h, CSLAM_NNS_EINVAL, "output arrays are required");
  if (k > kK1MaxK) return fail(h, CSLAM_NNS_EINVAL, "filtered search supports k <= 128");
  if (h->dim <= 0) return fail(h, CSLAM_NNS_ESTATE, "descriptor size not set");
  cudaStream_t stream = (cudaStream_t)stream_;
  DeviceGuard g(h->device);
  const long long kk = k < h->n ? k : h->n;
  if (out_kk) *out_kk = kk;
  if (B == 0) return 0;
  if (kk == 0) {
    for (long long i = 0; i < B * k; ++i) {
      out_idx[i] = -1;
      out_sim[i] = -std::numeric_limits<float>::infinity();
    }
    for (long long b = 0; b < B; ++b) out_match_pos[b] = -1;
    return 0;
  }
  const int k_run = (int)kk;
  const bool need_ids = mode == CSLAM_NNS_FILTER_INTRA && query_ids != nullptr;
  int rc = need_ids ? sync_row_ids(h, stream) : 0;
  if (rc) return rc;
  const float* dq;
  long long dq_ld;
  rc = stage_queries(h, queries, B, q_ld, src_is_device, stream, &dq, &dq_ld);
  if (rc) return rc;
  if (!h->res_dev.ensure((size_t)B * k_run * 12)) return fail(h, CSLAM_NNS_ENOMEM, "result buffer");
  if (!h->res_host.ensure((size_t)B * k * 12 + (size_t)B * 4)) return fail(h, CSLAM_NNS_ENOMEM, "pinned result buffer");
  long long* di = (long long*)h->res_dev.p;
  float* ds = (float*)(di + B * k_run);
  rc = search_device(h, dq, B, dq_ld, k_run, nullptr, di, ds, 1, 0, stream);
  if (rc) return rc;
  const long long* dqids = nullptr;
  if (need_ids && B > 1) {  // one query's id travels as a kernel argument
    if (!h->qids.ensure((size_t)B * 8)) return fail(h, CSLAM_NNS_ENOMEM, "query id buffer");
    CK(h, cudaMemcpyAsync(h->qids.p, query_ids, (size_t)B * 8, cudaMemcpyHostToDevice, stream));
    dqids = (const long long*)h->qids.p;
  }
  long long* hi = (long long*)h->res_host.p;
  float* hsim = (float*)(hi + B * k);
  int* hpos = (int*)(hsim + B * k);
  h->k1_chain = false;  // the filter kernel sits between this search and the next one
  filter_kernel<<<(unsigned)((B + 127) / 128), 128, 0, stream>>>(
      di, ds, B, k_run, k, (const long long*)h->ids_arena.base(), dqids,
      need_ids ? (long long)query_ids[0] : 0ll, need_ids ? 1 : 0, mode, (long long)min_gap, threshold,
      hi, hsim, hpos);
  count_launch();
  CK(h, cudaGetLastError());
  CK(h, cudaStreamSynchronize(stream));
  memcpy(out_idx, hi, (size_t)B * k * 8);
  memcpy(out_sim, hsim, (size_t)B * k * 4);
  memcpy(out_match_pos, hpos, (size_t)B * 4);
  return 0;
}

int cslam_nns_search_keys(cslam_nns_t h, const float* queries, int64_t B, int64_t q_ld,
                          int src_is_device, int k, int64_t row_stride, int64_t row_offset,
                          uint64_t* out_keys_dev, void* stream_) {
  if (!h) return CSLAM_NNS_EINVAL;
  if (B <= 0 || k <= 0 || !queries || !out_keys_dev) return fail(h, CSLAM_NNS_EINVAL, "bad arguments");
  if (k > kK1MaxK) return fail(h, CSLAM_NNS_EINVAL, "sharded search supports k <= 128");
  if (h->dim <= 0) return fail(h, CSLAM_NNS_ESTATE, "descriptor size not set");
  cudaStream_t stream = (cudaStream_t)stream_;
  DeviceGuard g(h->device);
  if (h->n == 0) {
    CK(h, cudaMemsetAsync(out_keys_dev, 0, (size_t)B * k * 8, stream));
    return 0;
  }
  const float* dq;
  long long dq_ld;
  int rc = stage_queries(h, queries, B, q_ld, src_is_device, stream, &dq, &dq_ld);
  if (rc) return rc;
  return search_device(h, dq, B, dq_ld, k, (u64*)out_keys_dev, nullptr, nullptr, row_stride,
                       row_offset, stream);
}

int cslam_nns_search_sharded(cslam_nns_t h, cslam_nns_xchg_t x, const float* queries, int64_t B, int64_t q_ld,
                             int src_is_device, int k, int rank, int world, int phase, int64_t* out_idx_dev,
                             float* out_sim_dev, void* stream_) {
  if (!h) return CSLAM_NNS_EINVAL;
  if (!x || B <= 0 || k <= 0 || world < 1 || rank < 0 || rank >= world || phase < 1 || phase > 3)
    return fail(h, CSLAM_NNS_EINVAL, "bad arguments to cslam_nns_search_sharded");
  cudaStream_t stream = (cudaStream_t)stream_;
  DeviceGuard g(h->device);
  if (!h->shard_keys.ensure((size_t)B * k * 8)) return fail(h, CSLAM_NNS_ENOMEM, "shard key buffer");
  uint64_t* keys = (uint64_t*)h->shard_keys.p;
  if (phase & 1) {
    if (h->dim <= 0 || h->n == 0) {  // this rank owns no row yet
      CK(h, cudaMemsetAsync(keys, 0, (size_t)B * k * 8, stream));
    } else {
      int rc = cslam_nns_search_keys(h, queries, B, q_ld, src_is_device, k, world, rank, keys, stream_);
      if (rc) return rc;
    }
  }
  int rc = cslam_nns_xchg_merge(x, keys, B, k, phase, out_idx_dev, out_sim_dev, stream_);
  if (rc) return fail(h, rc, std::string("exchange: ") + cslam_nns_xchg_last_error(x));
  return 0;
}

int cslam_nns_merge_keys(int device, const uint64_t* keys_dev, int parts, int64_t B, int k,
                         int64_t* out_idx, float* out_sim, int dst_is_device, void* stream_,
                         int64_t* out_kk) {
  if (!keys_dev || parts <= 0 || B <= 0 || k <= 0 || k > kK1MaxK) {
    g_create_error = "bad arguments to cslam_nns_merge_keys";
    return CSLAM_NNS_EINVAL;
  }
  cudaStream_t stream = (cudaStream_t)stream_;
  DeviceGuard g(device);
  if (dst_is_device) {
    cudaError_t e = k4_merge((const u64*)keys_dev, parts, B, k, k, nullptr, (long long*)out_idx,
                             out_sim, 1, 0, stream);
    if (e != cudaSuccess) return fail_cuda(nullptr, "k4_merge", e);
    return 0;
  }
  void* pin = nullptr;
  if (cudaHostAlloc(&pin, (size_t)B * k * 12, cudaHostAllocMapped) != cudaSuccess)
    return fail(nullptr, CSLAM_NNS_ENOMEM, "pinned result buffer");
  long long* hi = (long long*)pin;
  float* hs = (float*)(hi + B * k);
  cudaError_t e = k4_merge((const u64*)keys_dev, parts, B, k, k, nullptr, hi, hs, 1, 0, stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
  if (e != cudaSuccess) {
    cudaFreeHost(pin);
    return fail_cuda(nullptr, "k4_merge", e);
  }
  if (out_idx) memcpy(out_idx, hi, (size_t)B * k * 8);
  if (out_sim) memcpy(out_sim, hs, (size_t)B * k * 4);
  if (out_kk) {
    long long c = 0;
    while (c < k && hi[c] >= 0) ++c;
    *out_kk = c;
  }
  cudaFreeHost(pin);
  return 0;
}

int cslam_nns_profile(cslam_nns_t h, int enable) {
  if (!h) return CSLAM_NNS_EINVAL;
  h->profile = enable != 0;
  return 0;
}

int cslam_nns_profile_read(cslam_nns_t h, float* ms_total, int64_t* launches) {
  if (!h) return CSLAM_NNS_EINVAL;
  DeviceGuard g(h->device);
  float total = 0.f;
  for (size_t i = 0; i + 1 < h->ev_used; i += 2) {
    CK(h, cudaEventSynchronize(h->ev[i + 1]));
    float ms = 0.f;
    CK(h, cudaEventElapsedTime(&ms, h->ev[i], h->ev[i + 1]));
    total += ms;
  }
  if (ms_total) *ms_total = total;
  if (launches) *launches = (int64_t)(h->ev_used / 2);
  h->ev_used = 0;
  return 0;
}

int cslam_nns_set_variant(cslam_nns_t h, int variant) {
  if (!h || variant < 0 || variant > 3) return CSLAM_NNS_EINVAL;
  h->variant = variant;
  return 0;
}
int cslam_nns_last_variant(cslam_nns_t h) { return h ? h->last_variant : -1; }

// development: copy an internal device buffer to the host (0 = partial key lists, 1 = query parameter block)
int cslam_nns_debug_copy(cslam_nns_t h, int what, void* dst, int64_t bytes) {
  if (!h || !dst || bytes <= 0) return CSLAM_NNS_EINVAL;
  DeviceGuard g(h->device);
  const DevBuf& b = what == 0 ? h->part_keys : what == 2 ? h->sort_kv : h->inv_qn;
  if (!b.p || (size_t)bytes > b.bytes) return fail(h, CSLAM_NNS_EINVAL, "debug buffer smaller than requested");
  CK(h, cudaMemcpy(dst, b.p, (size_t)bytes, cudaMemcpyDeviceToHost));
  return 0;
}

int cslam_nns_set_screen(cslam_nns_t h, int enable, int64_t min_rows) {
  if (!h) return CSLAM_NNS_EINVAL;
  h->screen = enable != 0;
  if (min_rows >= 0) h->screen_min_rows = min_rows;
  return 0;
}

int cslam_nns_screen_stats(cslam_nns_t h, int64_t* screened, int64_t* fallback) {
  if (!h) return CSLAM_NNS_EINVAL;
  unsigned long long v[8] = {};
  if (h->screen_stats.p) {
    DeviceGuard g(h->device);
    CK(h, cudaMemcpy(v, h->screen_stats.p, sizeof v, cudaMemcpyDeviceToHost));
  }
  if (screened) *screened = (int64_t)v[0];
  if (fallback) *fallback = (int64_t)v[1];
  if (env_int("CSLAM_NNS_SCREEN_VERBOSE", 0))
    fprintf(stderr, "[screen] screened %llu fallback %llu: nan/bad %llu, < k candidates %llu, band overflow %llu, "
                    "outside bound too close %llu, nan bound %llu, bad rows %llu\n", v[0], v[1], v[2], v[3], v[4], v[5],
            v[6], v[7]);
  return 0;
}

int cslam_nns_time_search(cslam_nns_t h, const float* queries_dev, int64_t B, int64_t q_ld, int k,
                          int iters, void* stream_, float* ms_per_search) {
  if (!h || !queries_dev || B <= 0 || k <= 0 || iters <= 0 || !ms_per_search)
    return fail(h, CSLAM_NNS_EINVAL, "bad arguments");
  if (h->n == 0) return fail(h, CSLAM_NNS_ESTATE, "empty database");
  cudaStream_t stream = (cudaStream_t)stream_;
  DeviceGuard g(h->device);
  const int k_run = (int)(k < h->n ? k : h->n);
  const float* dq;
  long long dq_ld;
  int rc = stage_queries(h, queries_dev, B, q_ld, 1, stream, &dq, &dq_ld);
  if (rc) return rc;
  if (!h->res_dev.ensure((size_t)B * k_run * 12)) return fail(h, CSLAM_NNS_ENOMEM, "result buffer");
  long long* di = (long long*)h->res_dev.p;
  float* ds = (float*)(di + B * k_run);
  cudaEvent_t e0, e1;
  CK(h, cudaEventCreate(&e0));
  CK(h, cudaEventCreate(&e1));
  rc = search_device(h, dq, B, dq_ld, k_run, nullptr, di, ds, 1, 0, stream);  // warm
  if (rc == 0) {
    cudaEventRecord(e0, stream);
    for (int i = 0; i < iters && rc == 0; ++i)
      rc = search_device(h, dq, B, dq_ld, k_run, nullptr, di, ds, 1, 0, stream);
    cudaEventRecord(e1, stream);
    cudaError_t e = cudaEventSynchronize(e1);
    if (rc == 0 && e != cudaSuccess) rc = fail_cuda(h, "cudaEventSynchronize", e);
    float ms = 0.f;
    if (rc == 0) {
      cudaEventElapsedTime(&ms, e0, e1);
      *ms_per_search = ms / iters;
    }
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  return rc;
}

}  // extern "C"
#pragma GCC visibility pop
