// Shared device helpers for the cslam NNS kernels (sm_100a only).
//
// Key encoding.  Every candidate is one 64-bit key
//     key = orderable(sim) << 32 | ~row            (row < 2^32)
// so that "larger key" == "higher similarity, then lower row index".  orderable() is the usual
// monotone float->uint map; NaN is canonicalised to 0xFFFFFFFF so it sorts FIRST, which is what
// torch.argsort(descending=True) does at cslam/nns_matching.py:62.  key 0 can never be produced by
// a real candidate and marks an empty slot.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace cslam {

typedef unsigned long long u64;

static constexpr float kEps = 1e-8f;  // torch cosine_similarity eps (nns_matching.py:61)
static constexpr unsigned kFull = 0xffffffffu;

__host__ __device__ __forceinline__ uint32_t sim_to_ord(float s) {
#ifdef __CUDA_ARCH__
  uint32_t b = __float_as_uint(s);
#else
  uint32_t b;
  memcpy(&b, &s, 4);
#endif
  if ((b & 0x7fffffffu) > 0x7f800000u) return 0xffffffffu;  // NaN sorts first
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

__host__ __device__ __forceinline__ float ord_to_sim(uint32_t o) {
  uint32_t b;
  if (o == 0xffffffffu) b = 0x7fc00000u;
  else b = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o;
#ifdef __CUDA_ARCH__
  return __uint_as_float(b);
#else
  float f;
  memcpy(&f, &b, 4);
  return f;
#endif
}

__host__ __device__ __forceinline__ u64 make_key(float sim, uint32_t row) {
  return ((u64)sim_to_ord(sim) << 32) | (u64)(~row);
}
__host__ __device__ __forceinline__ uint32_t key_row(u64 key) { return ~(uint32_t)key; }
__host__ __device__ __forceinline__ float key_sim(u64 key) { return ord_to_sim((uint32_t)(key >> 32)); }

// ------------------------------------------------------------------------------------------------
// Warp-level top-k list kept in shared memory: `list[0..k)` sorted descending, empty slots = 0.
// All 32 lanes call these together with identical (warp-uniform) arguments.
// ------------------------------------------------------------------------------------------------

__device__ __forceinline__ void warp_list_clear(u64* list, int k, int lane) {
  for (int i = lane; i < k; i += 32) list[i] = 0ull;
  __syncwarp();
}

// Insert `key` (warp-uniform) if it beats the current k-th entry.  `thr` mirrors list[k-1].
__device__ __forceinline__ void warp_list_insert(u64* list, int k, u64 key, u64& thr, int lane) {
  if (key <= thr) return;
  int pos = 0;
  for (int base = 0; base < k; base += 32) {
    int i = base + lane;
    bool gt = (i < k) && (list[i] > key);
    pos += __popc(__ballot_sync(kFull, gt));
  }
  // shift [pos, k-2] -> [pos+1, k-1], highest chunk first so every read sees pre-shift data
  for (int base = ((k - 1) >> 5) << 5; base >= 0; base -= 32) {
    int i = base + lane;
    bool act = (i < k) && (i > pos);
    u64 v = 0ull;
    if (act) v = list[i - 1];
    __syncwarp();
    if (act) list[i] = v;
    __syncwarp();
  }
  if (lane == 0) list[pos] = key;
  __syncwarp();
  thr = list[k - 1];
}

// Offer one key per lane (`have` says whether this lane's key is real).
__device__ __forceinline__ void warp_list_offer(u64* list, int k, u64 key, bool have, u64& thr,
                                                int lane) {
  unsigned m = __ballot_sync(kFull, have && key > thr);
  while (m) {
    int src = __ffs(m) - 1;
    m &= m - 1;
    u64 cand = __shfl_sync(kFull, key, src);
    warp_list_insert(list, k, cand, thr, lane);
  }
}

// Stream `count` keys from memory (global or shared) into the list.
__device__ __forceinline__ void warp_list_stream(u64* list, int k, u64& thr, const u64* src,
                                                 int64_t count, int lane) {
  for (int64_t base = 0; base < count; base += 32) {
    int64_t i = base + lane;
    u64 key = 0ull;
    bool have = i < count;
    if (have) key = src[i];
    warp_list_offer(list, k, key, have && key != 0ull, thr, lane);
  }
}

// Per-warp running top-k.  k == 1: the best key is the register `thr`.  k <= 32: lane j holds the
// j-th best key in `mine` (sorted descending, 0 = empty).  k > 32: sorted list in shared memory.
// `floor` (optional): a proven lower bound of the final k-th key; anything below it is dropped even
// while the list is not full yet (used by the merge steps, where it removes most insertions).
__device__ __forceinline__ void offer(u64* list, int k, u64 key, bool have, u64& thr, u64& mine,
                                      int lane, u64 floor = 0ull) {
  unsigned m = __ballot_sync(kFull, have && key > thr && key >= floor);
  if (m == 0) return;  // the common case after warm-up
  if (k == 1) {
    u64 best = (have && key > thr) ? key : 0ull;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
      const u64 o = __shfl_xor_sync(kFull, best, off);
      best = o > best ? o : best;
    }
    thr = best;  // best > old thr by construction
  } else if (k <= 32) {
    while (m) {
      const int src = __ffs(m) - 1;
      m &= m - 1;
      const u64 cand = __shfl_sync(kFull, key, src);
      if (cand > thr) {
        const int pos = __popc(__ballot_sync(kFull, lane < k && mine > cand));
        const u64 up = __shfl_up_sync(kFull, mine, 1);
        if (lane < k) {
          if (lane > pos) mine = up;
          else if (lane == pos) mine = cand;
        }
        thr = __shfl_sync(kFull, mine, k - 1);
      }
    }
  } else {
    while (m) {
      const int src = __ffs(m) - 1;
      m &= m - 1;
      warp_list_insert(list, k, __shfl_sync(kFull, key, src), thr, lane);
    }
  }
}

// Top 32 of the union of two descending lists held one key per lane (0 = empty).  C[i] = max(A[i], B[31 - i])
// is a bitonic sequence that holds exactly the 32 largest keys of the union; five compare-exchange stages
// sort it descending.  ~12 shuffles whatever the lists hold, against ~200 clk for EVERY key the one-at-a-time
// insertion of offer() accepts: this is what the merge steps of the k <= 32 tails use.
__device__ __forceinline__ u64 warp_merge_desc(u64 a, u64 b, int lane) {
  const u64 br = __shfl_sync(kFull, b, 31 - lane);
  u64 v = a > br ? a : br;
#pragma unroll
  for (int j = 16; j > 0; j >>= 1) {
    const u64 o = __shfl_xor_sync(kFull, v, j);
    const bool want_max = (lane & j) == 0;  // the lower lane of a pair keeps the larger key
    v = (want_max == (v > o)) ? v : o;
  }
  return v;
}

// ---- lists of up to 128 keys in registers: element i = r * 32 + lane of v[r], sorted descending, 0 = empty ----
// (32 < k <= 128 of the single-query kernel; the k <= 32 code above with four registers per lane instead of one)

// Offer one key per lane; `thr` mirrors element k - 1.  The list itself keeps up to 128 keys.
__device__ __forceinline__ void offer_big(u64 (&v)[4], int k, u64 key, bool have, u64& thr, int lane) {
  unsigned m = __ballot_sync(kFull, have && key > thr);
  if (m == 0) return;  // the common case after warm-up
  const int kr = (k - 1) >> 5, kl = (k - 1) & 31;
  while (m) {
    const int src = __ffs(m) - 1;
    m &= m - 1;
    const u64 cand = __shfl_sync(kFull, key, src);
    if (cand > thr) {
      int pos = 0;
#pragma unroll
      for (int r = 0; r < 4; ++r) pos += __popc(__ballot_sync(kFull, v[r] > cand));
#pragma unroll
      for (int r = 3; r >= 0; --r) {   // highest register first: v[r - 1] is still the old one when it is read
        const u64 up = __shfl_up_sync(kFull, v[r], 1);
        const u64 carry = r > 0 ? __shfl_sync(kFull, v[r > 0 ? r - 1 : 0], 31) : 0ull;
        const int idx = r * 32 + lane;
        if (idx > pos) v[r] = lane == 0 ? carry : up;
        else if (idx == pos) v[r] = cand;
      }
      const u64 sel = kr == 0 ? v[0] : (kr == 1 ? v[1] : (kr == 2 ? v[2] : v[3]));
      thr = __shfl_sync(kFull, sel, kl);
    }
  }
}

// a := the 128 largest keys of the union of two such lists, sorted.  C[i] = max(A[i], B[127 - i]) is bitonic and
// holds exactly those keys; two register-level and five shuffle-level compare-exchange stages sort it.
__device__ __forceinline__ void warp_merge_desc_big(u64 (&a)[4], const u64 (&b)[4], int lane) {
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const u64 br = __shfl_sync(kFull, b[3 - r], 31 - lane);
    a[r] = a[r] > br ? a[r] : br;
  }
#pragma unroll
  for (int r = 0; r < 2; ++r) {   // distance 64: registers r and r + 2
    const u64 hi = a[r] > a[r + 2] ? a[r] : a[r + 2], lo = a[r] > a[r + 2] ? a[r + 2] : a[r];
    a[r] = hi;
    a[r + 2] = lo;
  }
#pragma unroll
  for (int r = 0; r < 4; r += 2) {   // distance 32: registers r and r + 1
    const u64 hi = a[r] > a[r + 1] ? a[r] : a[r + 1], lo = a[r] > a[r + 1] ? a[r + 1] : a[r];
    a[r] = hi;
    a[r + 1] = lo;
  }
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    u64 x = a[r];
#pragma unroll
    for (int j = 16; j > 0; j >>= 1) {
      const u64 o = __shfl_xor_sync(kFull, x, j);
      const bool want_max = (lane & j) == 0;
      x = (want_max == (x > o)) ? x : o;
    }
    a[r] = x;
  }
}

// Block-wide exact select: out[r] = the candidate with exactly r larger candidates, for r < k.
// Keys are unique (they embed the row) except for empties (0), which are skipped.  `out[0..k)` must be
// zeroed and visible to the block before the call; all threads of the block call it.  O(M^2 / threads)
// comparisons: meant for the small candidate sets of the merge steps.
__device__ __forceinline__ void block_rank_select(const u64* cand, int M, int k, u64* out) {
  for (int i = threadIdx.x; i < M; i += blockDim.x) {
    const u64 key = cand[i];
    if (key == 0ull) continue;
    int rank = 0;
    for (int j = 0; j < M; ++j) rank += (cand[j] > key) ? 1 : 0;
    if (rank < k) out[rank] = key;
  }
}

// ------------------------------------------------------------------------------------------------
// PTX wrappers: mbarrier, bulk async copy (TMA 1-D), TMA 2-D tiles, tcgen05.
// ------------------------------------------------------------------------------------------------

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}

// generic-proxy accesses before this point are ordered before later async-proxy accesses
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// 1-D bulk copy global -> shared, completion signalled on `bar` (SASS: UBLKCP).
// dst, src and bytes must be multiples of 16.
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_u32(dst)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}

// same, with an L2 eviction-priority hint (createpolicy-encoded 64-bit policy)
__device__ __forceinline__ void bulk_g2s_hint(void* dst, const void* src, uint32_t bytes, uint64_t* bar,
                                              unsigned long long policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::
          "r"(smem_u32(dst)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
      : "memory");
}

// 2-D tiled TMA load (SASS: UTMALDG).  c0 = innermost (column) coordinate, c1 = row coordinate.
__device__ __forceinline__ void tma_load_2d(void* dst, const void* tmap, int c0, int c1,
                                            uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, "
      "{%3, %4}], [%2];" ::"r"(smem_u32(dst)),
      "l"(tmap), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const void* tmap) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(tmap) : "memory");
}

__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(pred));
  return pred != 0;
}

__device__ __forceinline__ void named_bar_sync(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

// named barrier that also returns the OR of `pred` over the participating threads
__device__ __forceinline__ bool named_bar_or(int id, int threads, bool pred) {
  uint32_t out;
  asm volatile(
      "{\n\t.reg .pred p, q;\n\t"
      "setp.ne.u32 q, %3, 0;\n\t"
      "barrier.cta.red.or.pred p, %1, %2, q;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(out)
      : "r"(id), "r"(threads), "r"((uint32_t)pred)
      : "memory");
  return out != 0;
}

// -- tcgen05 ---------------------------------------------------------------------------------------
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
template <int kCols>
__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                   smem_u32(dst_smem)),
               "n"(kCols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <int kCols>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(kCols)
               : "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc], tf32 inputs, fp32 accumulate.  One thread issues.
__device__ __forceinline__ void umma_tf32_ss(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b,
                                             uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// mbarrier arrive once all previously issued tcgen05.mma of this thread have completed
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                   smem_u32(bar))
               : "memory");
}
// 32 lanes x 8 consecutive columns: thread t of the warp gets lane (base_lane + t)
__device__ __forceinline__ void tmem_ld_32x32b_x8(uint32_t taddr, uint32_t (&v)[8]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]),
        "=r"(v[7])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() {
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// D[tmem] (+)= A[tmem] * B[smem]: the A operand (M = 128 rows on lanes 0..127, K consecutive 32-bit
// columns) is read from tensor memory instead of shared memory
__device__ __forceinline__ void umma_tf32_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b,
                                             uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem_d),
      "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// registers -> TMEM: lane (32 * (warp % 4) + laneid) receives this thread's 32 values in 32 consecutive columns
__device__ __forceinline__ void tmem_st_32x32b_x32(uint32_t taddr, const uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() {
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}

// -- CTA-pair (cluster of 2, tcgen05 cta_group::2) helpers ------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive on the barrier at the same shared-memory offset in CTA `cta` of the cluster.  Default
// semantics (.release.cta) as in CUTLASS' ClusterBarrier::arrive(cta_id): an explicit .release.cluster
// costs a cluster-scope MEMBAR (~1800 clk per arrive measured on B200), which made the converter warps
// the bottleneck of the whole K2 pipeline.
__device__ __forceinline__ void mbar_arrive_cluster(uint64_t* bar, uint32_t cta) {
  asm volatile(
      "{\n\t.reg .b32 ra;\n\t"
      "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
      "mbarrier.arrive.shared::cluster.b64 _, [ra];\n\t}" ::"r"(smem_u32(bar)),
      "r"(cta)
      : "memory");
}
// wait on a barrier whose arrivals may come from the peer CTA (default acquire semantics, as CUTLASS does;
// acquire.cluster makes ptxas emit CCTL.IVALL + MEMBAR in the spin loop)
__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
  } while (!ok);
}
// 2-D tiled TMA load with an L2 eviction-priority hint
static constexpr unsigned long long kEvictFirst = 0x12F0000000000000ull;
static constexpr unsigned long long kEvictLast = 0x14F0000000000000ull;
__device__ __forceinline__ void tma_load_2d_hint(void* dst, const void* tmap, int c0, int c1,
                                                 uint64_t* bar, unsigned long long policy) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.L2::cache_hint "
      "[%0], [%1, {%3, %4}], [%2], %5;" ::"r"(smem_u32(dst)),
      "l"(tmap), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "l"(policy)
      : "memory");
}
// 3-D tiled TMA load (one CTA), with an L2 cache-policy hint
__device__ __forceinline__ void tma_load_3d_hint(void* dst, const void* tmap, int c0, int c1, int c2,
                                                 uint64_t* bar, unsigned long long policy) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.L2::cache_hint "
      "[%0], [%1, {%3, %4, %5}], [%2], %6;" ::"r"(smem_u32(dst)),
      "l"(tmap), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "l"(policy)
      : "memory");
}
// 3-D tiled TMA load issued by either CTA of a pair; completion bytes go to the barrier at this offset in
// the LEADER CTA (rank 0) of the cluster
__device__ __forceinline__ void tma_load_3d_pair(void* dst, const void* tmap, int c0, int c1, int c2,
                                                 uint64_t* bar, unsigned long long policy) {
  asm volatile(
      "{\n\t.reg .b32 rb;\n\t"
      "mapa.shared::cluster.u32 rb, %2, 0;\n\t"
      "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.tile.mbarrier::complete_tx::bytes.L2::cache_hint "
      "[%0], [%1, {%3, %4, %5}], [rb], %6;\n\t}" ::"r"(smem_u32(dst)),
      "l"(tmap), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "l"(policy)
      : "memory");
}
template <int kCols>
__device__ __forceinline__ void tmem_alloc_pair(uint32_t* dst_smem) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                   smem_u32(dst_smem)),
               "n"(kCols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
template <int kCols>
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(kCols)
               : "memory");
}
// D[tmem of both CTAs] (+)= A * B over the CTA pair (M = 256: 128 rows per CTA; each CTA holds N/2 of B)
__device__ __forceinline__ void umma_tf32_ss_pair(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b,
                                                  uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on the barrier at this offset in BOTH CTAs once all prior tcgen05.mma of this thread are done
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {
  const uint16_t mask = 3;
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
          smem_u32(bar)),
      "h"(mask)
      : "memory");
}
// 32 lanes x 16 consecutive columns
__device__ __forceinline__ void tmem_ld_32x32b_x16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, "
      "%13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]),
        "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]),
        "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
}

}  // namespace cslam
