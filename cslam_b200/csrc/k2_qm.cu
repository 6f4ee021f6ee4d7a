// K2 (query-major pair variant) — batched cosine-similarity search, 97..256 queries per pass (smaller batches: k2_batched.cu).
//
// Replaces B calls of cslam/nns_matching.py:61-63 by ONE pass over the database.  Two CTAs of a cluster
// form one tcgen05 cta_group::2 pair and compute, per 256-row super tile of the database,
//     S[256 queries x 256 rows] = Q[256 x dim] * X[256 x dim]^T            (3xTF32, fp32 accumulate in TMEM)
// with the QUERIES on the MMA's M side: CTA r owns queries 128r .. 128r+127 (A operand, its TMEM lanes)
// and stages database rows 128r .. 128r+127 of the super tile (its half of the B operand).  Compared with
// a row-major tiling (database rows on M, k2_batched.cu) each CTA stages only HALF of the query block per
// K block — 128 database rows (raw -> x_hi, x_lo) and 128 queries (q_hi, q_lo) — which halves the L2 ->
// shared-memory query traffic per flop, and the accumulator comes out with TMEM lane == query, so every epilogue thread owns ONE query and keeps its top-k list privately:
// no warp collectives, no locks, a register threshold, and a warm-up that costs k*ln(256/k) scalar
// insertions per thread instead of k warp-wide extractions per (warp, query).
//
//   warp  roles (512 threads per CTA; both CTAs run every role except the MMA issue)
//   0-3, 12-15  epilogue   thread <-> query 128*rank + t and one half of the 256 columns (warps 0-3 columns
//                    0..127, warps 12-15 columns 128..255): tcgen05.ld 32 columns (= database rows) at a time,
//                    s = (acc + mean(q) * sum(x)) * inv|x| * inv|q|, compare with the thread's threshold,
//                    insert into the thread's sorted list (shared memory, [k][128] so lanes never conflict)
//   4     X producer  TMA of this CTA's 128 database rows (L2 evict-first)
//   5     MMA         leader CTA only: 3 products x (BK/8) tcgen05.mma.cta_group::2 per K block;
//                     tcgen05.commit multicast frees the ring slot / publishes the accumulator in both CTAs
//   6     Q producer  ONE 3-D TMA for this CTA's 128 rows of q_hi and q_lo (L2 evict-last); the bytes are
//                     counted on the LEADER's barrier (cp.async.bulk.tensor .cta_group::2)
//   8-11  converter   split the landed rows (x_hi in place, x_lo beside it), accumulate |x|^2 and sum(x);
//                     at the end of a tile write 1/max(|x|,eps) and sum(x) into BOTH CTAs' norm arrays
//                     (st.shared::cluster) because both epilogues need all 256 rows of the super tile
//
// Tensor-bound: algorithmic flops per launch = 2 * n * dim * 256-padded batch (issued: 3x).
// Algorithmic HBM bytes per launch = n * ld * 4 (the database is read exactly once per pass).
#include <cuda.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "kernels.h"

namespace cslam {
namespace {

constexpr int BM = 128;   // queries per CTA (MMA M = 256 over the pair) and database rows per CTA and tile
constexpr int BN2 = 256;  // database rows per super tile (MMA N)
constexpr int kThreads = 512;  // warps 0-3 and 12-15 epilogue, 4 X producer, 5 MMA, 6 Q producer, 8-11 converter
constexpr int kTmemCols = 512;
constexpr int kMaxStages = 12;

struct QDev {
  long long n;
  long long tiles2;  // 256-row super tiles
  int dim, B, k, stages, kblocks, dbg;
  uint32_t idesc;
  const float* inv_qn;
  const float* q_mean;
  unsigned* gthr;  // [B] grid-wide lower bound of each query's k-th best similarity (ordered uint), 0 = none
  u64* part_keys;  // [2 * pairs][B][k]
  u64* prof;
  // masked fallback pass: only queries whose bit is set take part, and the grid exits when *active_count == 0
  const unsigned* active_mask;
  const unsigned* active_count;
};

struct QSmem {
  uint32_t tile_bytes, x_lo_off, q_hi_off, q_lo_off, stage_bytes;
  uint32_t lists_off, norm_off, bars_off, tmem_off, total;
};

// SCREEN: the screening pass stages the raw database tile (the tensor core cuts it to tf32 itself) and ONE
// query plane: 2 tiles per stage instead of 4, so the ring is twice as deep and the L2 -> shared-memory
// traffic per K block drops from 3 units (x, q_hi, q_lo) to 2
template <int BK, bool SCREEN>
__host__ __device__ inline QSmem qcarve(int k, int stages) {
  QSmem s;
  s.tile_bytes = BM * BK * 4;
  s.x_lo_off = s.tile_bytes;
  s.q_hi_off = (SCREEN ? 1 : 2) * s.tile_bytes;
  s.q_lo_off = 3 * s.tile_bytes;
  s.stage_bytes = (SCREEN ? 2 : 4) * s.tile_bytes;  // 16 KB (BK = 32) or 8 KB (BK = 16) per tile: multiple of 1024
  uint32_t o = s.stage_bytes * (uint32_t)stages;
  s.lists_off = o;  o += 2u * BM * (k + 4) * 8;        // [2 column halves][k entries + 4 group minima][128] keys
  s.norm_off = o;   o += 2u * 2u * BN2 * 4;            // [2 accumulators][inv|x| (256), sum(x) (256)]
  s.bars_off = o;   o += (4u * kMaxStages + 10u) * 8;
  s.tmem_off = o;   o += 16;
  s.total = o;
  return s;
}

template <int BK>
__device__ __forceinline__ uint64_t qdesc(uint32_t saddr) {
  // K-major operand; rows of BK*4 bytes swizzled in 8-row atoms (SBO = 8 rows); LBO unused (1); version 1
  constexpr uint64_t kLayout = BK == 32 ? 2 : 4;  // SWIZZLE_128B : SWIZZLE_64B
  constexpr uint64_t kSbo = (8 * BK * 4) >> 4;
  return (uint64_t)((saddr & 0x3ffffu) >> 4) | ((uint64_t)1 << 16) | (kSbo << 32) | ((uint64_t)1 << 46) |
         (kLayout << 61);
}

__device__ __forceinline__ float q_tf32_hi(float x) {
  return __uint_as_float(__float_as_uint(x) & 0xffffe000u);
}
__device__ __forceinline__ float q_tf32_rna(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}

// Copy `bytes` (multiple of 16) of this CTA's shared memory to the SAME offset in CTA `peer` of the cluster with the
// bulk-copy engine; the bytes are counted (complete_tx) on the mbarrier at `bar`'s offset in the peer.  Waiters on
// that barrier see the data without any cluster-scope fence, like the bytes of a TMA load.
__device__ __forceinline__ void bulk_s2peer(const void* src, uint32_t bytes, uint64_t* bar, uint32_t peer) {
  asm volatile(
      "{\n\t.reg .b32 rd, rb;\n\t"
      "mapa.shared::cluster.u32 rd, %0, %3;\n\t"
      "mapa.shared::cluster.u32 rb, %2, %3;\n\t"
      "cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [rd], [%0], %1, [rb];\n\t}" ::"r"(
          smem_u32(src)),
      "r"(bytes), "r"(smem_u32(bar)), "r"(peer)
      : "memory");
}
#define QPROF_WAIT(acc, stmt)                   \
  do {                                          \
    if (p.prof) {                               \
      const long long t0__ = clock64();         \
      stmt;                                     \
      acc += clock64() - t0__;                  \
    } else {                                    \
      stmt;                                     \
    }                                           \
  } while (0)

// State of one thread-private list, passed and returned BY VALUE so that it stays in registers across the
// out-of-line call (inlining the insertion 32x per chunk blew the epilogue up to ~15k instructions and the
// instruction cache with it: C5 went from 5.9 to 15.5 ms).
struct TopState {
  int cnt, min_pos, min_grp;
  u64 min_key;
  float thr;      // similarity below which a candidate cannot matter: max(own smallest entry once the list
                  // is full, best bound published by any list of this query)
  float thr_u;    // the same bound for u = s / inv|q| (saves a multiply per column), 2^-21 slack
  unsigned g_seen;  // best bound known to this list: its own smallest entry once full, or a published one (ordered uint)
  unsigned dirty;   // g_seen comes from this list and has not been published yet (the caller does, once per tile)
};
__device__ __forceinline__ void top_set_threshold(TopState& st, float thr, float invq) {
  st.thr = thr;
  const float t_ = thr / invq;
  st.thr_u = t_ - fabsf(t_) * 4.8e-7f;
}
// max that returns NaN when either operand is NaN (fmaxf drops NaNs; a NaN similarity must reach the exact test)
__device__ __forceinline__ float max_nan(float a, float b) {
  float r;
  asm("max.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
  return r;
}
// v[j] for a lane-dependent j without local memory: a 4-level select tree (15 selects)
__device__ __forceinline__ uint32_t select16(const uint32_t (&v)[16], int j) {
  uint32_t a[8], b[4];
#pragma unroll
  for (int i = 0; i < 8; ++i) a[i] = (j & 1) ? v[2 * i + 1] : v[2 * i];
#pragma unroll
  for (int i = 0; i < 4; ++i) b[i] = (j & 2) ? a[2 * i + 1] : a[2 * i];
  const uint32_t c0 = (j & 4) ? b[1] : b[0], c1 = (j & 4) ? b[3] : b[2];
  return (j & 8) ? c1 : c0;
}
// tcgen05.wait::ld that names the 16 registers of the load it completes, so that no use of them can be scheduled
// above it (another load may be in flight into a second set of registers at the same time)
__device__ __forceinline__ void tmem_ld_wait16(uint32_t (&v)[16]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]),
                 "+r"(v[8]), "+r"(v[9]), "+r"(v[10]), "+r"(v[11]), "+r"(v[12]), "+r"(v[13]), "+r"(v[14]),
                 "+r"(v[15])
               :
               : "memory");
}
// Smallest entry of lst[lo .. hi) and its position (entries are [j * BM] apart).
__device__ __forceinline__ void group_min(const u64* lst, int lo, int hi, u64& mk, int& mp) {
  mk = ~0ull;
  mp = lo;
  for (int j = lo; j < hi; ++j) {
    const u64 e = lst[j * BM];
    if (e < mk) {
      mk = e;
      mp = j;
    }
  }
}
// Called by all 32 lanes together, each with its own list and candidate (key 0 = nothing to offer).
// The list is an UNSORTED set of k keys cut into 4 groups of gs = ceil(k / 4) entries whose smallest entries are
// kept behind it (lst[(k + g) * BM]).  Replacing the list's smallest entry re-scans ONE group and compares four
// cached minima: ~80 instructions, where a re-scan of all k entries cost ~400 (k = 24) — and this function is
// what a list's warm-up costs: k ln(rows / k) replacements, each a lane-divergent round of the epilogue.
__device__ __noinline__ TopState offer_key(u64* lst, int k, int gs, u64 key, float invq, unsigned* gthr, TopState st) {
  if (key == 0ull) return st;
  u64* gm = lst + (size_t)k * BM;
  u64 mk;
  if (st.cnt < k) {
    lst[st.cnt * BM] = key;
    if (++st.cnt < k) return st;
    // the list has just filled up: all four group minima from scratch
    mk = ~0ull;
    for (int g = 0; g < 4; ++g) {
      const int lo = g * gs, hi = (g + 1) * gs < k ? (g + 1) * gs : k;
      u64 m;
      int mp;
      group_min(lst, lo < k ? lo : k, hi, m, mp);   // an empty group (small k) keeps the sentinel and never wins
      gm[g * BM] = m;
      if (m < mk) {
        mk = m;
        st.min_pos = mp;
        st.min_grp = g;
      }
    }
  } else {
    if (key <= st.min_key) return st;
    lst[st.min_pos * BM] = key;
    const int g = st.min_grp;
    const int lo = g * gs, hi = (g + 1) * gs < k ? (g + 1) * gs : k;
    u64 mg;
    int mpg;
    group_min(lst, lo, hi, mg, mpg);
    gm[g * BM] = mg;
    const u64 m0 = g == 0 ? mg : gm[0], m1 = g == 1 ? mg : gm[BM], m2 = g == 2 ? mg : gm[2 * BM],
              m3 = g == 3 ? mg : gm[3 * BM];
    const u64 a = m1 < m0 ? m1 : m0, b = m3 < m2 ? m3 : m2;
    const int ga = m1 < m0 ? 1 : 0, gb = m3 < m2 ? 3 : 2;
    mk = b < a ? b : a;
    const int bg = b < a ? gb : ga;
    if (bg == g) {
      st.min_pos = mpg;
    } else {  // the smallest entry now sits in another group: find where
      const int lo2 = bg * gs, hi2 = (bg + 1) * gs < k ? (bg + 1) * gs : k;
      u64 m_;
      group_min(lst, lo2, hi2, m_, st.min_pos);
      st.min_grp = bg;
    }
  }
  st.min_key = mk;
  // the list is full: its smallest entry is a proven lower bound of the query's final k-th best
  // (published by the caller at the end of the tile: 2 x 74 lists raising their bounds entry by entry during
  // the warm-up are thousands of atomics on ONE word per query, which the L2 serialises)
  const unsigned ord = (unsigned)(mk >> 32);
  if (ord > st.g_seen && ord != 0xffffffffu) {  // NaN entries are never a bound
    st.g_seen = ord;
    st.dirty = gthr != nullptr ? 1u : 0u;
    top_set_threshold(st, key_sim(mk), invq);
  }
  return st;
}

template <int BK, bool SCREEN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
k2q_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_q,
           const QDev p) {
  // masked fallback pass, queued unconditionally behind the screening pass: nothing to do when every query's
  // candidate set was proven complete (uniform over the grid, before any barrier or allocation)
  if (p.active_count != nullptr && __ldcg(p.active_count) == 0u) return;
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  // identical layout in both CTAs (the pair MMA addresses the peer's operands by the leader's offsets)
  unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const QSmem L = qcarve<BK, SCREEN>(p.k, p.stages);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int S = p.stages;
  const uint32_t rank = cluster_ctarank();
  const long long pair = blockIdx.x >> 1, npairs = gridDim.x >> 1;

  u64* lists = reinterpret_cast<u64*>(smem + L.lists_off);
  float* norms = reinterpret_cast<float*>(smem + L.norm_off);  // [a][0..255] inv|x|, [a][256..511] sum(x)
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L.bars_off);
  uint64_t* full_bar = bars;                     // [S] local: own database rows landed
  uint64_t* conv_bar = bars + kMaxStages;        // [S] LEADER: both CTAs' rows are split (8 warp arrivals)
  uint64_t* empty_bar = bars + 2 * kMaxStages;   // [S] local: the pair MMAs that read the slot are done
  uint64_t* qfull_bar = bars + 3 * kMaxStages;   // [S] LEADER: both CTAs' query tiles landed (tx bytes)
  uint64_t* acc_full = bars + 4 * kMaxStages;    // [2] local: accumulator complete (multicast commit)
  uint64_t* acc_empty = acc_full + 2;            // [2] both CTAs: all 8 epilogue warps are done with
                                                 //     accumulator a and norms[a] (each warp arrives on both)
  uint64_t* norm_bar = acc_full + 4;             // [2] local: all 256 rows' norms are in (8 converter warps)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + L.tmem_off);

  for (int i = tid; i < 2 * BM * (p.k + 4); i += kThreads) lists[i] = 0ull;
  if (tid == 0) {
    for (int s = 0; s < S; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&conv_bar[s], 8);
      mbar_init(&empty_bar[s], 1);
      mbar_init(&qfull_bar[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&acc_full[a], 1);
      mbar_init(&acc_empty[a], 16);
      mbar_init(&norm_bar[a], 4);   // the 4 local converter warps; the peer's half arrives as 4 x 256 tx bytes
    }
    mbar_fence_init();
  }
  if (warp == 5) tmem_alloc_pair<kTmemCols>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const long long my_tiles = pair < p.tiles2 ? (p.tiles2 - pair + npairs - 1) / npairs : 0;

  if (warp == 4) {
    // ===== X producer ===============================================================================
    if (lane == 0) {
      tma_prefetch_desc(&map_x);
      long long w0 = 0;
      const long long tstart = clock64();
      int s = 0;
      uint32_t ph = 0;  // ring position and wrap parity, advanced incrementally (no div/mod per stage)
      for (long long t = 0; t < my_tiles; ++t) {
        const int row0 = (int)(((pair + t * npairs) * 2 + rank) * BM);
        for (int kb = 0; kb < p.kblocks; ++kb, s = (s + 1 == S ? 0 : s + 1), ph ^= (s == 0)) {
          QPROF_WAIT(w0, mbar_wait(&empty_bar[s], ph ^ 1));
          mbar_expect_tx(&full_bar[s], L.tile_bytes);
          tma_load_2d_hint(smem + (size_t)s * L.stage_bytes, &map_x, kb * BK, row0, &full_bar[s], kEvictFirst);
        }
      }
      if (p.prof) {
        p.prof[blockIdx.x * 16 + 0] = (u64)w0;
        p.prof[blockIdx.x * 16 + 7] = (u64)(clock64() - tstart);
      }
    }
  } else if (warp == 6) {
    // ===== Q producer ===============================================================================
    if (lane == 0) {
      tma_prefetch_desc(&map_q);
      long long w0 = 0;
      int s = 0;
      uint32_t ph = 0;
      for (long long t = 0; t < my_tiles; ++t) {
        for (int kb = 0; kb < p.kblocks; ++kb, s = (s + 1 == S ? 0 : s + 1), ph ^= (s == 0)) {
          QPROF_WAIT(w0, mbar_wait(&empty_bar[s], ph ^ 1));
          if (rank == 0) mbar_expect_tx(&qfull_bar[s], (SCREEN ? 2u : 4u) * L.tile_bytes);  // both CTAs' boxes
          tma_load_3d_pair(smem + (size_t)s * L.stage_bytes + L.q_hi_off, &map_q, kb * BK, (int)rank * BM, 0,
                           &qfull_bar[s], kEvictLast);
        }
      }
      if (p.prof) p.prof[blockIdx.x * 16 + 1] = (u64)w0;
    }
  } else if (warp == 5) {
    // ===== MMA issuer (leader CTA only) =============================================================
    // The whole warp runs the loop and one elected lane issues: in an `if (lane == 0)` region ptxas wraps
    // every tcgen05.mma in an ELECT / BRA.U.ANY loop, which costs about as much as the MMA itself.
    if (rank == 0) {
      long long w0 = 0, w1 = 0, w2 = 0;
      int s = 0;
      uint32_t ph = 0;
      const uint32_t ring = smem_u32(smem);
      for (long long t = 0; t < my_tiles; ++t) {
        const int a = (int)(t & 1);
        QPROF_WAIT(w2, mbar_wait_cluster(&acc_empty[a], (uint32_t)(((t >> 1) & 1) ^ 1)));
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(a * BN2);
        for (int kb = 0; kb < p.kblocks; ++kb, s = (s + 1 == S ? 0 : s + 1), ph ^= (s == 0)) {
          QPROF_WAIT(w0, mbar_wait_cluster(&qfull_bar[s], ph));
          QPROF_WAIT(w1, mbar_wait_cluster(&conv_bar[s], ph));
          tc_fence_after();
          const uint32_t st = ring + (uint32_t)s * L.stage_bytes;
          const uint64_t x_hi = qdesc<BK>(st), x_lo = qdesc<BK>(st + L.x_lo_off);
          const uint64_t q_hi = qdesc<BK>(st + L.q_hi_off), q_lo = qdesc<BK>(st + L.q_lo_off);
          const uint32_t first = kb != 0 ? 1u : 0u;
          if (elect_one()) {
#pragma unroll
            for (int ks = 0; ks < BK / 8; ++ks) {
              const uint64_t adv = (uint64_t)(2 * ks);  // 32 bytes along K = +2 in the 16-byte address field
              if (SCREEN) {  // one product: tf32(q - c) x tf32(x), both cut by the tensor core itself
                umma_tf32_ss_pair(d_tmem, q_hi + adv, x_hi + adv, p.idesc, ks ? 1u : first);
              } else if (!(p.dbg & 2)) {
                umma_tf32_ss_pair(d_tmem, q_hi + adv, x_lo + adv, p.idesc, ks ? 1u : first);
                umma_tf32_ss_pair(d_tmem, q_lo + adv, x_hi + adv, p.idesc, 1u);
                umma_tf32_ss_pair(d_tmem, q_hi + adv, x_hi + adv, p.idesc, 1u);
              } else {
                umma_tf32_ss_pair(d_tmem, q_hi + adv, x_hi + adv, p.idesc, ks ? 1u : first);
              }
            }
            umma_commit_pair(&empty_bar[s]);
          }
          __syncwarp();
        }
        if (elect_one()) umma_commit_pair(&acc_full[a]);
        __syncwarp();
      }
      if (p.prof && lane == 0) {
        p.prof[blockIdx.x * 16 + 2] = (u64)w0;
        p.prof[blockIdx.x * 16 + 3] = (u64)w1;
        p.prof[blockIdx.x * 16 + 4] = (u64)w2;
      }
    }
  } else if (warp >= 8 && warp < 12) {
    // ===== converter ===============================================================================
    const int r = tid - 8 * 32;  // row of this CTA's 128-row tile
    constexpr int kChunks = BK / 4;  // 16-byte chunks per row
    int s = 0;
    uint32_t ph = 0;
    long long w0 = 0, wce = 0, wcf = 0;
    for (long long t = 0; t < my_tiles; ++t) {
      const int a = (int)(t & 1);
      float nn = 0.f, sx = 0.f;
      for (int kb = 0; kb < p.kblocks; ++kb, s = (s + 1 == S ? 0 : s + 1), ph ^= (s == 0)) {
        QPROF_WAIT(w0, mbar_wait(&full_bar[s], ph));
        unsigned char* st = smem + (size_t)s * L.stage_bytes;
        if (!(p.dbg & 1)) {
          float4* raw = reinterpret_cast<float4*>(st + (size_t)r * (BK * 4));
          float4* lo = reinterpret_cast<float4*>(st + L.x_lo_off + (size_t)r * (BK * 4));
          float4 x[kChunks];
#pragma unroll
          for (int i = 0; i < kChunks; ++i) {
            // LOGICAL 16-byte chunk i of the row sits at physical chunk i ^ (swizzle bits of the row): a
            // quarter-warp touches 8 distinct bank groups, and |x|^2 / sum(x) are accumulated in column
            // order whatever the row's position in the tile, so a row's similarity does not depend on how
            // the database is partitioned (bit-identical results for any shard count)
            const int c = BK == 32 ? (i ^ (r & 7)) : (i ^ ((r >> 1) & 3));
            x[i] = raw[c];
          }
#pragma unroll
          for (int i = 0; i < kChunks; ++i) {
            const int c = BK == 32 ? (i ^ (r & 7)) : (i ^ ((r >> 1) & 3));
            float4 h, l;
            h.x = q_tf32_hi(x[i].x); h.y = q_tf32_hi(x[i].y); h.z = q_tf32_hi(x[i].z); h.w = q_tf32_hi(x[i].w);
            l.x = q_tf32_rna(x[i].x - h.x); l.y = q_tf32_rna(x[i].y - h.y);
            l.z = q_tf32_rna(x[i].z - h.z); l.w = q_tf32_rna(x[i].w - h.w);
            nn = fmaf(x[i].x, x[i].x, nn); nn = fmaf(x[i].y, x[i].y, nn);
            nn = fmaf(x[i].z, x[i].z, nn); nn = fmaf(x[i].w, x[i].w, nn);
            sx += (x[i].x + x[i].y) + (x[i].z + x[i].w);
            if (!SCREEN && !(p.dbg & 16)) {  // the screening pass only reads: |x|^2 and sum(x), no split tiles
              if (!(p.dbg & 8)) raw[c] = h;
              lo[c] = l;
            }
          }
          if (!SCREEN && !(p.dbg & 16)) fence_proxy_async();  // generic-proxy stores -> visible to the tensor core's async proxy
        }
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(&conv_bar[s], 0);
      }
      // both epilogues must be done with norms[a] (tile t-2) before either converter overwrites it
      QPROF_WAIT(wce, mbar_wait(&acc_empty[a], (uint32_t)(((t >> 1) & 1) ^ 1)));
      float* ninv = norms + a * (2 * BN2) + rank * BM + r;
      const float inv = rsqrtf(fmaxf(nn, kEps * kEps));  // 1 / max(|x|, eps)
      // Both epilogues need all 256 rows: this warp's 32 rows go into the local arrays with plain stores and into
      // the peer's with two 128-byte bulk copies that complete on the PEER's norm barrier, while the local barrier
      // is told to expect the peer's 256 bytes for the same rows of its half.  (st.shared::cluster + fence.acq_rel
      // .cluster on both sides cost the converter ~2.2k clk and every epilogue warp ~0.9k clk per tile, a quarter
      // of a 256-float tile's time.)  The source is not rewritten before both epilogues have released tile t
      // (acc_empty above), by which time the peer has long consumed the copy.
      ninv[0] = inv;
      ninv[BN2] = sx;
      QPROF_WAIT(wcf, fence_proxy_async());   // generic-proxy stores -> visible to the bulk-copy engine
      __syncwarp();
      if (lane == 0) {
        mbar_expect_tx(&norm_bar[a], 256u);   // = arrive + expect: our 32 rows are in, the peer's 2 x 128 B are due
        bulk_s2peer(ninv, 128u, &norm_bar[a], rank ^ 1u);
        bulk_s2peer(ninv + BN2, 128u, &norm_bar[a], rank ^ 1u);
      }
    }
    if (p.prof && tid == 8 * 32) {
      p.prof[blockIdx.x * 16 + 5] = (u64)w0;
      p.prof[blockIdx.x * 16 + 11] = (u64)wce;
      p.prof[blockIdx.x * 16 + 12] = (u64)wcf;
    }
  } else if (warp < 4 || warp >= 12) {
    // ===== epilogue: thread <-> (query, half of the super tile's 256 columns) ========================
    // Warp w may only touch TMEM lanes 32 (w % 4) .. +31, so warps 0-3 take columns 0..127 and warps 12-15
    // columns 128..255 of the same 128 queries: two private lists per query, merged by K4 like any other part.
    const int eh = warp >= 12 ? 1 : 0;
    const int tq = (warp & 3) * 32 + lane;
    const int q = (int)rank * BM + tq;
    const bool valid = q < p.B && (p.active_mask == nullptr || ((__ldcg(p.active_mask + (q >> 5)) >> (q & 31)) & 1u));
    const float qmean_r = valid ? p.q_mean[q] : 0.f;
    const float invq_r = valid ? p.inv_qn[q] : 0.f;
    const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
    u64* lst = lists + (size_t)eh * BM * (p.k + 4) + tq;
    const int gs = (p.k + 3) >> 2;   // entries per group of the list (offer_key)
    // Thread-private UNSORTED top-k: the first k candidates fill the list, after that a candidate replaces
    // the smallest entry and the minimum is re-found with k independent loads (a sorted list costs a
    // dependent load-compare-store chain per shifted entry; K4 does not need sorted parts).
    TopState ts;
    ts.cnt = 0;
    ts.min_pos = 0;
    ts.min_grp = 0;
    ts.min_key = 0ull;
    ts.thr = -INFINITY;
    ts.thr_u = -INFINITY;
    ts.g_seen = 0u;
    ts.dirty = 0u;
    // Grid-wide threshold sharing.  The 2 x 74 lists of a query each warm up from nothing; whenever one of
    // them is full its smallest entry is a proven lower bound of the query's final k-th best (k real rows
    // at or above it exist), so it is published with one atomicMax (inside offer_key) and every list of the
    // query — full or not — drops candidates below the best bound published so far (read once per tile).
    // This removes most of the k ln(rows/k) insertions per list, which diverge lane by lane.
    unsigned* gq = (valid && p.gthr != nullptr && !(p.dbg & 64)) ? p.gthr + q : nullptr;
    // Seeding sample pass (k == 1, no shared bound): the thread's best row is a running maximum in registers, no
    // list and no insertion rounds at all
    const bool max_only = p.k == 1 && p.gthr == nullptr;   // uniform over the grid (gq is per lane: a warp-wide vote follows)
    float best_u = -INFINITY;
    uint32_t best_row = 0xffffffffu;
    long long wepi = 0, wnorm = 0, wfence = 0, wscan = 0;
    for (long long t = 0; t < my_tiles; ++t) {
      const int a = (int)(t & 1);
      const uint32_t ph = (uint32_t)((t >> 1) & 1);
      const long long row_base = (pair + t * npairs) * (long long)BN2;
      const int valid_cols = (int)((p.n - row_base) < BN2 ? (p.n - row_base) : BN2);
      if (gq) {
        const unsigned g = __ldcg(gq);
        if (g > ts.g_seen) {
          ts.g_seen = g;
          top_set_threshold(ts, ord_to_sim(g), invq_r);
        }
      }
      QPROF_WAIT(wnorm, mbar_wait(&norm_bar[a], ph));
      // (no cluster fence: the peer's half of norms[a] came as bulk-copy bytes counted on norm_bar)
      QPROF_WAIT(wepi, mbar_wait(&acc_full[a], ph));
      tc_fence_after();
      const long long tscan0 = p.prof ? clock64() : 0;
      const float* ninv = norms + a * (2 * BN2);
      const float* nsum = ninv + BN2;
      // 8 chunks of 16 columns, two register sets: the tcgen05.ld of chunk i + 1 is in flight while chunk i is
      // scanned (one 32-column load followed by its scan left the ~250 clk of the load exposed four times per tile,
      // and with two epilogue warps per scheduler there is little else to hide it).
      auto scan16 = [&](uint32_t (&v)[16], int c0) {
        if (c0 >= valid_cols || (p.dbg & 4)) return;
        if (p.dbg & 32) {   // development: TMEM loads only (keeps the registers live), no scan
          if (v[0] == 0x7fc12345u && v[15] == 0x7fc54321u) ts.g_seen ^= 1u;
          return;
        }
        // fast path: u = (c sum(x) + acc) / |x| for the 16 columns with packed fp32x2 instructions (same roundings as
        // fmaf / a multiply) and their NaN-propagating maximum: 2.5 issue slots per column.  Only when some lane's
        // maximum reaches its bound (the bound is only read here; a stale value just sends the warp to the exact
        // test below for nothing) is the per-column survivor mask built.  The per-column FSETP / SEL / IADD3 this
        // replaced made the epilogue issue-bound at 256-float rows: 2.9k clk per tile against 1k for the bare loads.
        float mx = -INFINITY;
        const float2 qm2 = make_float2(qmean_r, qmean_r);
#pragma unroll
        for (int j4 = 0; j4 < 16; j4 += 4) {
          const float4 iv = *reinterpret_cast<const float4*>(ninv + c0 + j4);
          const float4 sm = *reinterpret_cast<const float4*>(nsum + c0 + j4);
          const float2 u01 = __fmul2_rn(__ffma2_rn(qm2, make_float2(sm.x, sm.y),
                                                   make_float2(__uint_as_float(v[j4]), __uint_as_float(v[j4 + 1]))),
                                        make_float2(iv.x, iv.y));
          const float2 u23 = __fmul2_rn(__ffma2_rn(qm2, make_float2(sm.z, sm.w),
                                                   make_float2(__uint_as_float(v[j4 + 2]), __uint_as_float(v[j4 + 3]))),
                                        make_float2(iv.z, iv.w));
          v[j4] = __float_as_uint(u01.x);
          v[j4 + 1] = __float_as_uint(u01.y);
          v[j4 + 2] = __float_as_uint(u23.x);
          v[j4 + 3] = __float_as_uint(u23.y);
          mx = max_nan(max_nan(mx, u01.x), max_nan(u01.y, max_nan(u23.x, u23.y)));
        }
        if (max_only) {
          if (valid && mx > best_u) {   // NaN and the zero rows past the end of the last tile never win
            int jb = 0;
#pragma unroll
            for (int jj = 15; jj >= 0; --jj) jb = (__uint_as_float(v[jj]) == mx) ? jj : jb;   // first column of the maximum
            if (c0 + jb < valid_cols) {
              best_u = mx;
              best_row = (uint32_t)(row_base + c0 + jb);
            }
          }
          return;
        }
        if (!__any_sync(kFull, valid && !(mx < ts.thr_u))) return;
        // ties go to the exact key comparison (a lower row index wins); NaN passes and sorts first
        uint32_t m = 0u;
#pragma unroll
        for (int jj = 0; jj < 16; ++jj) m |= !(__uint_as_float(v[jj]) < ts.thr_u) ? (1u << jj) : 0u;
        if (c0 + 16 > valid_cols) m &= (1u << (valid_cols - c0)) - 1u;   // the database's last, partial tile
        if (!valid || (p.dbg & 16)) m = 0u;
        // slow path, CONVERGENT: every lane pops its own next survivor (its value is picked out of the 16
        // registers by a select tree) and all lanes run the insertion together, each on its own list; a lane
        // without a survivor rides along with key 0.  Rounds per chunk = the largest survivor count of any
        // lane (1-2 after warm-up), where a branch per column served one lane at a time.
        while (__any_sync(kFull, m != 0u)) {
          const int j = m != 0u ? __ffs(m) - 1 : 0;
          const bool has = m != 0u;
          m &= m - 1u;
          const float s_ = __uint_as_float(select16(v, j)) * invq_r;
          const u64 key = (has && !(s_ < ts.thr))
                              ? (((u64)sim_to_ord(s_) << 32) | (u64)(~(uint32_t)(row_base + c0 + j)))
                              : 0ull;
          const float thr_was = ts.thr_u;
          ts = offer_key(lst, p.k, gs, key, invq_r, gq, ts);
          // A list whose bound has just moved re-tests its remaining survivors at once: the mask above was built
          // with the bound the chunk started with, and while a list fills up or warms up that bound is far too
          // low — every remaining column then cost a round of its own (the whole first tile of a list: 128 rounds;
          // the seeding sample pass, one tile per list, spent 20 us there for 2.6 us of products).
          if (__any_sync(kFull, ts.thr_u != thr_was)) {
            uint32_t keep = 0u;
#pragma unroll
            for (int jj = 0; jj < 16; ++jj) keep |= !(__uint_as_float(v[jj]) < ts.thr_u) ? (1u << jj) : 0u;
            m &= keep;
          }
        }
      };
      {
        uint32_t va[16], vb[16];
        const uint32_t t0 = tmem_base + lane_base + (uint32_t)(a * BN2 + eh * (BN2 / 2));
        const int cb = eh * (BN2 / 2);
        tmem_ld_32x32b_x16(t0, va);
#pragma unroll 1
        for (int i = 0; i < (BN2 / 2) / 16; i += 2) {
          tmem_ld_wait16(va);
          tmem_ld_32x32b_x16(t0 + (uint32_t)((i + 1) * 16), vb);
          scan16(va, cb + i * 16);
          tmem_ld_wait16(vb);
          if (i + 2 < (BN2 / 2) / 16) tmem_ld_32x32b_x16(t0 + (uint32_t)((i + 2) * 16), va);
          scan16(vb, cb + (i + 1) * 16);
        }
      }
      if (ts.dirty) {  // this list's bound grew during the tile: one atomic per list and tile at most
        atomicMax(gq, ts.g_seen);
        ts.dirty = 0u;
      }
      if (p.prof) wscan += clock64() - tscan0;
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive_cluster(&acc_empty[a], 0);
        mbar_arrive_cluster(&acc_empty[a], 1);
      }
    }
    if (p.prof && tid == 0) {
      p.prof[blockIdx.x * 16 + 6] = (u64)wepi;
      p.prof[blockIdx.x * 16 + 8] = (u64)wnorm;
      p.prof[blockIdx.x * 16 + 9] = (u64)wfence;
      p.prof[blockIdx.x * 16 + 10] = (u64)wscan;
    }
    // this thread's list -> part_keys[2 * pair + column half][q][k]
    if (max_only && valid) lst[0] = best_row != 0xffffffffu ? make_key(best_u * invq_r, best_row) : 0ull;
    if (valid) {
      u64* out = p.part_keys + ((size_t)(2 * pair + eh) * p.B + q) * p.k;
      for (int j = 0; j < p.k; ++j) out[j] = lst[j * BM];
    }
  }

  // ---- teardown -----------------------------------------------------------------------------------
  __syncwarp();
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 5) {
    tc_fence_after();
    tmem_dealloc_pair<kTmemCols>(tmem_base);
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

bool qencode_x(void* fn, CUtensorMap* m, const float* base, long long rows, int dim, int ld, int bk) {
  const cuuint64_t gdim[2] = {(cuuint64_t)dim, (cuuint64_t)rows};
  const cuuint64_t gstr[1] = {(cuuint64_t)ld * 4};
  const cuuint32_t box[2] = {(cuuint32_t)bk, (cuuint32_t)BM};
  const cuuint32_t estr[2] = {1, 1};
  return reinterpret_cast<EncodeTiledFn>(fn)(
             m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), gdim, gstr, box, estr,
             CU_TENSOR_MAP_INTERLEAVE_NONE, bk == 32 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B,
             CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}
// the two planes written by k2_prepare (q_hi [B][q_ld] then q_lo [B][q_ld]) as ONE 3-D tensor
// (planes == 1: the screening pass loads the q_hi plane only)
bool qencode_q(void* fn, CUtensorMap* m, const float* planes, int B, int dim, int q_ld, int bk, int nplanes) {
  const cuuint64_t gdim[3] = {(cuuint64_t)dim, (cuuint64_t)B, (cuuint64_t)nplanes};
  const cuuint64_t gstr[2] = {(cuuint64_t)q_ld * 4, (cuuint64_t)B * q_ld * 4};
  const cuuint32_t box[3] = {(cuuint32_t)bk, (cuuint32_t)BM, (cuuint32_t)nplanes};
  const cuuint32_t estr[3] = {1, 1, 1};
  return reinterpret_cast<EncodeTiledFn>(fn)(
             m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(planes), gdim, gstr, box, estr,
             CU_TENSOR_MAP_INTERLEAVE_NONE, bk == 32 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B,
             CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

constexpr size_t kSmemCap = 227 * 1024;

template <int BK, bool SCREEN>
int qpick_stages(int k) {
  for (int s = kMaxStages; s >= 2; --s)
    if ((size_t)qcarve<BK, SCREEN>(k, s).total + 1024 <= kSmemCap) return s;
  return 0;
}

int q_env_bk() {
  static int bk = 0;
  if (!bk) {
    const char* v = getenv("CSLAM_NNS_K2_BK");
    bk = (v && atoi(v) == 16) ? 16 : 32;
  }
  return bk;
}

int q_pairs(long long n, int num_sms) {
  const long long tiles2 = (n + BN2 - 1) / BN2;
  const long long pairs = num_sms / 2;
  return (int)(tiles2 < pairs ? (tiles2 < 1 ? 1 : tiles2) : pairs);
}

template <int BK, bool SCREEN>
cudaError_t qlaunch(const K2Args& a, int num_sms, cudaStream_t stream, const char** err) {
  QDev p;
  p.n = a.n;
  p.tiles2 = (a.n + BN2 - 1) / BN2;
  p.dim = a.dim;
  p.B = a.B;
  p.k = a.k;
  p.stages = qpick_stages<BK, SCREEN>(a.k);
  {
    const char* v = getenv("CSLAM_NNS_K2_STAGES");
    if (v && atoi(v) >= 2 && atoi(v) < p.stages) p.stages = atoi(v);
    v = getenv("CSLAM_NNS_K2_DBG");
    p.dbg = v ? atoi(v) : 0;
  }
  p.kblocks = (a.dim + BK - 1) / BK;
  p.inv_qn = a.inv_qn + kQParInv * a.B;
  p.q_mean = a.inv_qn + kQParMean * a.B;
  // zeroed by k2_prepare; the masked fallback pass has its own plane (the screening pass's bounds are bounds
  // of the SCREENED similarities and are still needed by k6_rescore's proof when the fallback runs)
  p.gthr = a.gthr_plane < 0 ? nullptr   // no bound sharing between the lists (the seeding sample pass)
                            : reinterpret_cast<unsigned*>(const_cast<float*>(a.inv_qn)) + a.gthr_plane * a.B;
  p.active_mask = a.active_mask;
  p.active_count = a.active_count;
  p.part_keys = a.part_keys;
  // kind::tf32, fp32 accumulate, K-major operands; N>>3 at bit 17, M>>4 at bit 24 (M = 256 over the pair)
  p.idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(BN2 >> 3) << 17) | ((uint32_t)((2 * BM) >> 4) << 24);
  if (p.stages < 2) {
    if (err) *err = "k2: k does not fit shared memory";
    return cudaErrorInvalidConfiguration;
  }
  CUtensorMap mx, mq;
  if (!qencode_x(a.encode_fn, &mx, a.db, a.n, a.dim, a.ld, BK) ||
      !qencode_q(a.encode_fn, &mq, a.q_lo, a.B, a.dim, a.q_ld, BK, SCREEN ? 1 : 2)) {
    if (err) *err = "k2: cuTensorMapEncodeTiled failed";
    return cudaErrorInvalidValue;
  }
  const size_t smem = qcarve<BK, SCREEN>(p.k, p.stages).total + 1024;
  static bool attr_set[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64 || !attr_set[dev]) {
    cudaError_t e = cudaFuncSetAttribute(k2q_kernel<BK, SCREEN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemCap);
    if (e != cudaSuccess) return e;
    if (dev >= 0 && dev < 64) attr_set[dev] = true;
  }
  const int grid = 2 * q_pairs(a.n, num_sms);
  static u64* prof_buf = nullptr;
  static const bool want_prof = getenv("CSLAM_NNS_K2_PROF") != nullptr;
  p.prof = nullptr;
  if (want_prof) {
    if (!prof_buf) cudaMalloc(&prof_buf, 1024 * 16 * sizeof(u64));
    cudaMemsetAsync(prof_buf, 0, 1024 * 16 * sizeof(u64), stream);
    p.prof = prof_buf;
  }
  k2q_kernel<BK, SCREEN><<<grid, kThreads, smem, stream>>>(mx, mq, p);
  count_launch();
  if (want_prof) {
    static u64 hbuf[1024 * 16];
    cudaStreamSynchronize(stream);
    cudaMemcpy(hbuf, prof_buf, sizeof(hbuf), cudaMemcpyDeviceToHost);
    const char* names[13] = {"Xprod wait empty", "Qprod wait empty", "MMA wait qfull", "MMA wait conv",
                             "MMA wait acc_empty", "conv wait full", "epi wait acc_full", "total (Xprod)",
                             "epi wait norm", "epi fence", "epi scan", "conv wait acc_empty", "conv fence"};
    for (int r = 0; r < 2; ++r) {
      fprintf(stderr, "[k2q prof] rank %d (avg cycles over %d CTAs, %d stages):", r, grid / 2, p.stages);
      for (int i = 0; i < 13; ++i) {
        double acc = 0;
        for (int c = r; c < grid; c += 2) acc += (double)hbuf[c * 16 + i];
        fprintf(stderr, "  %s=%.0f", names[i], acc / (grid / 2));
      }
      fprintf(stderr, "\n");
    }
  }
  return cudaGetLastError();
}

// which kernel serves a batch of B: 0 = query-major pair kernel (this file: the MMA work is that of 256
// queries whatever B is), 2 = row-major one-CTA kernel (k2_batched.cu: MMA N = B padded to 16, cheaper for
// small batches).  Measured crossover on B200 (2M x 256 and 400k x 4096): between 64 and 128 queries.
// CSLAM_NNS_K2_FORCE=qm|rows overrides (development).
// Which kernel serves a batch.  Measured on B200 (2M x 256 and 400k x 4096, k = 10; profiles/r1_k2rows_*):
// the row-major kernel runs at max(HBM time, ~0.13 + 0.0053 * B ns per row) whatever the dimension (its
// epilogue sees every row x query pair), the query-major pair kernel at the tensor-pipe time of 256 queries
// (~0.002 ns per row per float of dimension) whatever the batch, so query-major wins once B * 256 > 72 * dim
// (B > 72 at dim 256, never below 97 from dim 384 up).
int q_route(int B, int dim) {
  static int forced = -2;
  if (forced == -2) {
    const char* v = getenv("CSLAM_NNS_K2_FORCE");
    forced = !v ? -1 : !strcmp(v, "qm") ? 0 : (!strcmp(v, "rows") || !strcmp(v, "legacy")) ? 2 : -1;
  }
  if (forced == 0 || (forced == 2 && B <= kRowsMaxBatch)) return forced;
  return (B > 96 || (long long)B * 256 > 72LL * dim) ? 0 : 2;
}

}  // namespace

bool k2_supported(int B, int k, int dim) {
  if (B < 1 || B > kK2MaxBatch || k < 1 || k > 32) return false;
  switch (q_route(B, dim)) {
    case 0: return (q_env_bk() == 16 ? qpick_stages<16, false>(k) : qpick_stages<32, false>(k)) >= 2;
    default: return k2s_supported(B, k);
  }
}

int k2_parts(long long n, int B, int dim, int num_sms) {
  switch (q_route(B, dim)) {
    case 0: return 2 * q_pairs(n, num_sms);
    default: return k2s_grid(n, num_sms);
  }
}

const unsigned* k2_floor_plane(const float* inv_qn, int B, int dim) {
  (void)dim;  // both kernels publish their bounds in the same plane
  return reinterpret_cast<const unsigned*>(inv_qn + kQParGthr * B);
}

// The screening route (one tf32 product + exact re-scoring, k6_rescore.cu) serves the batches the query-major
// kernel serves.  CSLAM_NNS_SCREEN=0 switches it off (every batch then takes the 3xTF32 pass).
bool k2_screen_supported(int B, int k, int dim) {
  static const bool off = getenv("CSLAM_NNS_SCREEN") && atoi(getenv("CSLAM_NNS_SCREEN")) == 0;
  if (off || B < 1 || B > kK2MaxBatch || k < 1 || k > 32 || q_env_bk() != 32) return false;
  return q_route(B, dim) == 0 && qpick_stages<32, true>(k2_screen_list_len(k)) >= 2;
}

// Lists of the screening pass hold more than k entries: 2k + 4, at least 12, at most 32 (k = 1 -> 12, 10 -> 24,
// >= 14 -> 32).  See k6_rescore.cu: the spare entries keep the published bound below the k-th best when a
// query's neighbours all fall into one list.  Larger lists cost shared memory (ring depth 6 -> 4 stages at 32),
// not time: insertions are warm-up work.  CSLAM_NNS_SCREEN_LIST overrides (development).
int k2_screen_list_len(int k) {
  static const int forced = getenv("CSLAM_NNS_SCREEN_LIST") ? atoi(getenv("CSLAM_NNS_SCREEN_LIST")) : 0;
  int kl = forced > 0 ? forced : 2 * k + 4;
  if (forced <= 0 && kl < 12) kl = 12;
  if (kl < k) kl = k;
  if (kl > 32) kl = 32;
  return kl;
}

cudaError_t k2_launch(const K2Args& a, int num_sms, cudaStream_t stream, const char** err) {
  switch (q_route(a.B, a.dim)) {
    case 0:
      if (a.screen) return qlaunch<32, true>(a, num_sms, stream, err);
      return q_env_bk() == 16 ? qlaunch<16, false>(a, num_sms, stream, err) : qlaunch<32, false>(a, num_sms, stream, err);
    default: return k2s_launch(a, num_sms, stream, err);
  }
}

}  // namespace cslam
