// K2 (pair variant) — batched cosine-similarity search on CTA PAIRS of the 5th-generation tensor cores.
//
// Same contract as k2_batched.cu (replaces B calls of cslam/nns_matching.py:61-63 by one pass over the
// database, 3xTF32 split precision, centred query, fused per-query top-k), re-tiled for the thing that
// bounded the one-CTA kernel: every 128-row database tile needs the WHOLE query block (B x 32 floats x
// {hi, lo} = 64 KB per K block at B = 256) re-streamed from L2 into shared memory, which (a) exceeds the
// ~6.3 KB/clk L2 cap when all 148 SMs do it and (b) leaves room for only two 96 KB ring stages, too few
// to cover the TMA -> convert -> MMA chain.  Here two CTAs of a cluster form one tcgen05 cta_group::2
// pair: M = 256 database rows (128 per CTA), N = B queries of which each CTA stages only HALF, so the
// query traffic per CTA and per flop halves, and the K block shrinks to 16 floats (64-byte swizzle rows)
// so that a stage is 32 KB and the ring is 5-6 deep.
//
//   warp  roles (320 threads per CTA, both CTAs run all roles except the MMA issue)
//   0-3   epilogue   tcgen05.ld own accumulator (lane == own database row), scale, pre-filter against
//                    the per-query threshold, insert survivors into shared per-query lists under a
//                    per-query spin lock (no CTA-wide barriers)
//   4-7   converter  split the landed database tile (x_hi in place, x_lo beside it), |x|^2 and sum(x);
//                    one elected lane per warp arrives on the LEADER's conv barrier (remote for CTA 1)
//   8     producer   TMA: own 128 database rows (evict-first) + own half of q_hi / q_lo (evict-last)
//   9     MMA        leader CTA only: 3 products x (BK/8) tcgen05.mma.cta_group::2 per K block;
//                    tcgen05.commit multicast frees the ring slot / publishes the accumulator in BOTH CTAs
#include <cuda.h>
#include <stdio.h>
#include <stdlib.h>

#include "common.cuh"
#include "kernels.h"

namespace cslam {
namespace {

constexpr int BM = 128;  // database rows per CTA and tile (UMMA M = 256 over the pair)
constexpr int kEpiThreads = 128, kConvThreads = 128, kConvWarps = 4, kEpiWarps = 4;
constexpr int kMaxGroups = 3;  // converter groups (4 warps each), group g splits the K blocks it % groups == g
constexpr int kMaxThreads = 256 + 128 * kMaxGroups;
constexpr int kTmemCols = 512;
constexpr int kMaxStages = 12;

struct PDev {
  long long n;
  long long tiles2;  // 256-row super tiles
  int dim, B, BN, k, stages, kblocks, groups, dbg;
  uint32_t idesc;
  const float* inv_qn;
  const float* q_mean;
  u64* part_keys;  // [grid][B][k]
  u64* prof;       // development: [grid][8] cycles spent waiting, per role (CSLAM_NNS_K2_PROF=1)
};

struct PSmem {
  uint32_t tile_bytes, q_bytes, a_lo_off, b_hi_off, b_lo_off, stage_bytes;
  uint32_t lists_off, thr_off, invq_off, qmean_off, lock_off, norm_off, bars_off, tmem_off, total;
};

template <int BK>
__host__ __device__ inline PSmem pcarve(int BN, int B, int k, int stages) {
  PSmem s;
  s.tile_bytes = BM * BK * 4;
  s.q_bytes = (uint32_t)(BN / 2) * BK * 4;  // the 3-D query load packs the hi and lo planes back to back
  s.a_lo_off = s.tile_bytes;
  s.b_hi_off = 2 * s.tile_bytes;
  s.b_lo_off = s.b_hi_off + s.q_bytes;
  s.stage_bytes = (s.b_lo_off + s.q_bytes + 1023u) & ~1023u;
  uint32_t o = s.stage_bytes * (uint32_t)stages;
  s.lists_off = o;  o += (uint32_t)B * k * 8;
  o = (o + 15u) & ~15u;
  s.thr_off = o;    o += (uint32_t)BN * 4;
  s.invq_off = o;   o += (uint32_t)BN * 4;
  s.qmean_off = o;  o += (uint32_t)BN * 4;
  s.lock_off = o;   o += (uint32_t)BN * 4;
  s.norm_off = o;   o += 4u * BM * 4 * kMaxGroups;  // [2][groups][BM] |x|^2 then the same for sum(x)
  o = (o + 7u) & ~7u;
  s.bars_off = o;   o += (4u * kMaxStages + 10u) * 8;
  s.tmem_off = o;   o += 16;
  s.total = o;
  return s;
}

template <int BK>
__device__ __forceinline__ uint64_t pdesc(uint32_t saddr) {
  // K-major operand; rows of BK*4 bytes swizzled in 8-row atoms (SBO = 8 rows); LBO unused (1); version 1
  constexpr uint64_t kLayout = BK == 32 ? 2 : 4;  // SWIZZLE_128B : SWIZZLE_64B
  constexpr uint64_t kSbo = (8 * BK * 4) >> 4;
  return (uint64_t)((saddr & 0x3ffffu) >> 4) | ((uint64_t)1 << 16) | (kSbo << 32) | ((uint64_t)1 << 46) |
         (kLayout << 61);
}

__device__ __forceinline__ float p_tf32_hi(float x) {
  return __uint_as_float(__float_as_uint(x) & 0xffffe000u);
}
__device__ __forceinline__ float p_tf32_rna(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}

__device__ __forceinline__ float4 lds_volatile_f4(const float* p) {
  float4 v;
  asm volatile("ld.volatile.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "r"(smem_u32(p)));
  return v;
}

// Offer this warp's candidates for query q (warp-uniform); `key` is 0 on lanes without one.
__device__ __forceinline__ void list_offer(u64* lst, volatile float* thrf, unsigned* lock, int q, int k,
                                           u64 key, int lane) {
  if (lane == 0) {
    while (atomicCAS(lock, 0u, 1u) != 0u) {
    }
  }
  __syncwarp();
  __threadfence_block();
  u64 mine = lane < k ? *reinterpret_cast<volatile u64*>(lst + lane) : 0ull;
  bool changed = false;
  while (true) {
    const u64 kth = __shfl_sync(kFull, mine, k - 1);
    const u64 cand = key > kth ? key : 0ull;
    if (__ballot_sync(kFull, cand != 0ull) == 0u) break;
    const uint32_t hi = __reduce_max_sync(kFull, (uint32_t)(cand >> 32));
    const uint32_t lo = __reduce_max_sync(kFull, (uint32_t)(cand >> 32) == hi ? (uint32_t)cand : 0u);
    const u64 mx = ((u64)hi << 32) | lo;
    const int pos = __popc(__ballot_sync(kFull, lane < k && mine > mx));
    const u64 up = __shfl_up_sync(kFull, mine, 1);
    if (lane < k) {
      if (lane > pos) mine = up;
      else if (lane == pos) mine = mx;
    }
    if (key == mx) key = 0ull;
    changed = true;
  }
  if (changed) {
    if (lane < k) lst[lane] = mine;
    if (lane == k - 1 && mine != 0ull) thrf[q] = key_sim(mine);
  }
  __threadfence_block();
  __syncwarp();
  if (lane == 0) atomicExch(lock, 0u);
}

#define PROF_WAIT(acc, stmt)                    \
  do {                                          \
    if (p.prof) {                               \
      const long long t0__ = clock64();         \
      stmt;                                     \
      acc += clock64() - t0__;                  \
    } else {                                    \
      stmt;                                     \
    }                                           \
  } while (0)

template <int BK>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kMaxThreads, 1)
k2p_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_q,
           const PDev p) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  // identical layout in both CTAs (the pair MMA addresses the peer's operands by the leader's offsets)
  unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const PSmem L = pcarve<BK>(p.BN, p.B, p.k, p.stages);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int S = p.stages;
  const uint32_t rank = cluster_ctarank();
  const long long pair = blockIdx.x >> 1, npairs = gridDim.x >> 1;

  u64* lists = reinterpret_cast<u64*>(smem + L.lists_off);
  float* thrf = reinterpret_cast<float*>(smem + L.thr_off);
  float* invq = reinterpret_cast<float*>(smem + L.invq_off);
  float* qmean = reinterpret_cast<float*>(smem + L.qmean_off);
  unsigned* locks = reinterpret_cast<unsigned*>(smem + L.lock_off);
  float* norms = reinterpret_cast<float*>(smem + L.norm_off);
  float* sums = norms + 2 * kMaxGroups * BM;
  const int G = p.groups;
  const int nthreads = blockDim.x;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L.bars_off);
  uint64_t* full_bar = bars;                      // [S] local: own TMA loads landed
  uint64_t* conv_bar = bars + kMaxStages;         // [S] LEADER: both CTAs' tiles are split (8 warp arrivals)
  uint64_t* empty_bar = bars + 2 * kMaxStages;    // [S] local: the pair MMAs that read the slot are done
  uint64_t* qfull_bar = bars + 3 * kMaxStages;    // [S] LEADER: both CTAs' query halves landed (tx bytes only)
  uint64_t* acc_full = bars + 4 * kMaxStages;     // [2] local: accumulator complete (multicast commit)
  uint64_t* acc_empty = acc_full + 2;             // [2] LEADER: both CTAs' epilogues drained it (8 warps)
  uint64_t* acc_free = acc_full + 4;              // [2] local: own epilogue is done with norms[a]
  uint64_t* norm_bar = acc_full + 6;              // [2] local: row norms of the tile are in smem
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + L.tmem_off);

  for (int i = tid; i < p.B * p.k; i += nthreads) lists[i] = 0ull;
  for (int i = tid; i < p.BN; i += nthreads) {
    thrf[i] = -INFINITY;
    locks[i] = 0u;
    invq[i] = i < p.B ? p.inv_qn[i] : 0.f;
    qmean[i] = i < p.B ? p.q_mean[i] : 0.f;
  }
  if (tid == 0) {
    for (int s = 0; s < S; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&conv_bar[s], 2 * kConvWarps);
      mbar_init(&empty_bar[s], 1);
      mbar_init(&qfull_bar[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&acc_full[a], 1);
      mbar_init(&acc_empty[a], 2 * kEpiWarps);
      mbar_init(&acc_free[a], kEpiThreads);
      mbar_init(&norm_bar[a], kConvThreads * G);
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
    // ===== X producer: this CTA's 128 database rows, one K block per stage ==========================
    if (lane == 0) {
      tma_prefetch_desc(&map_a);
      long long w0 = 0;
      const long long tstart = clock64();
      int s = 0;
      uint32_t ph = 0;  // ring position and wrap parity, advanced incrementally (no 64-bit div/mod per stage)
      for (long long t = 0; t < my_tiles; ++t) {
        const int row0 = (int)(((pair + t * npairs) * 2 + rank) * BM);
        for (int kb = 0; kb < p.kblocks; ++kb, s = (s + 1 == S ? 0 : s + 1), ph ^= (s == 0)) {
          PROF_WAIT(w0, mbar_wait(&empty_bar[s], ph ^ 1));
          mbar_expect_tx(&full_bar[s], (uint32_t)(BM * BK * 4));
          tma_load_2d_hint(smem + (size_t)s * L.stage_bytes, &map_a, kb * BK, row0, &full_bar[s], kEvictFirst);
        }
      }
      if (p.prof) {
        p.prof[blockIdx.x * 8 + 0] = (u64)w0;
        p.prof[blockIdx.x * 8 + 7] = (u64)(clock64() - tstart);
      }
    }
  } else if (warp == 6) {
    // ===== Q producer: this CTA's half of q_hi and q_lo in ONE 3-D load; bytes count on the leader =====
    if (lane == 0) {
      tma_prefetch_desc(&map_q);
      const int qrow0 = (int)rank * (p.BN / 2);
      long long w0 = 0;
      int s = 0;
      uint32_t ph = 0;
      for (long long t = 0; t < my_tiles; ++t) {
        for (int kb = 0; kb < p.kblocks; ++kb, s = (s + 1 == S ? 0 : s + 1), ph ^= (s == 0)) {
          PROF_WAIT(w0, mbar_wait(&empty_bar[s], ph ^ 1));
          if (rank == 0) mbar_expect_tx(&qfull_bar[s], (uint32_t)(2 * p.BN * BK * 4));
          tma_load_3d_pair(smem + (size_t)s * L.stage_bytes + L.b_hi_off, &map_q, kb * BK, qrow0, 0,
                           &qfull_bar[s], kEvictLast);
        }
      }
      if (p.prof) p.prof[blockIdx.x * 8 + 1] = (u64)w0;
    }
  } else if (warp == 5) {
    // ===== MMA issuer (leader CTA only) =============================================================
    if (rank == 0 && lane == 0) {
      long long w0 = 0, w1 = 0, w2 = 0;
      int s = 0;
      uint32_t ph = 0;
      for (long long t = 0; t < my_tiles; ++t) {
        const int a = (int)(t & 1);
        PROF_WAIT(w2, mbar_wait_cluster(&acc_empty[a], (uint32_t)(((t >> 1) & 1) ^ 1)));
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(a * p.BN);
        for (int kb = 0; kb < p.kblocks; ++kb, s = (s + 1 == S ? 0 : s + 1), ph ^= (s == 0)) {
          PROF_WAIT(w0, mbar_wait_cluster(&qfull_bar[s], ph));
          PROF_WAIT(w1, mbar_wait_cluster(&conv_bar[s], ph));
          tc_fence_after();
          const uint32_t st = smem_u32(smem + (size_t)s * L.stage_bytes);
#pragma unroll
          for (int ks = 0; ks < BK / 8; ++ks) {
            const uint64_t a_hi = pdesc<BK>(st + ks * 32);
            const uint64_t a_lo = pdesc<BK>(st + L.a_lo_off + ks * 32);
            const uint64_t b_hi = pdesc<BK>(st + L.b_hi_off + ks * 32);
            const uint64_t b_lo = pdesc<BK>(st + L.b_lo_off + ks * 32);
            if (p.dbg & 64) {
            } else if (!(p.dbg & 2)) {
              umma_tf32_ss_pair(d_tmem, a_lo, b_hi, p.idesc, (kb | ks) != 0 ? 1u : 0u);
              umma_tf32_ss_pair(d_tmem, a_hi, b_lo, p.idesc, 1u);
              umma_tf32_ss_pair(d_tmem, a_hi, b_hi, p.idesc, 1u);
            } else {
              umma_tf32_ss_pair(d_tmem, a_hi, b_hi, p.idesc, (kb | ks) != 0 ? 1u : 0u);
            }
          }
          umma_commit_pair(&empty_bar[s]);
        }
        umma_commit_pair(&acc_full[a]);
      }
      if (p.prof) {
        p.prof[blockIdx.x * 8 + 2] = (u64)w0;
        p.prof[blockIdx.x * 8 + 3] = (u64)w1;
        p.prof[blockIdx.x * 8 + 4] = (u64)w2;
      }
    }
  } else if (warp >= 8) {
    // ===== converter ===============================================================================
    const int g = (warp - 8) >> 2;
    const int r = (tid - 8 * 32) & 127;
    constexpr int kChunks = BK / 4;  // 16-byte chunks per row
    int s = 0, turn = 0;  // turn: which converter group owns this K block
    uint32_t ph = 0;
    long long w0 = 0;
    for (long long t = 0; t < my_tiles; ++t) {
      const int a = (int)(t & 1);
      float nn = 0.f, sx = 0.f;
      for (int kb = 0; kb < p.kblocks;
           ++kb, s = (s + 1 == S ? 0 : s + 1), ph ^= (s == 0), turn = (turn + 1 == G ? 0 : turn + 1)) {
        if (turn != g) continue;
        PROF_WAIT(w0, mbar_wait(&full_bar[s], ph));
        unsigned char* st = smem + (size_t)s * L.stage_bytes;
        if (p.dbg & 1) {
          if (!(p.dbg & 8)) fence_proxy_async();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(&conv_bar[s], 0);
          continue;
        }
        float4* raw = reinterpret_cast<float4*>(st + (size_t)r * (BK * 4));
        float4* lo = reinterpret_cast<float4*>(st + L.a_lo_off + (size_t)r * (BK * 4));
        float4 x[kChunks];
#pragma unroll
        for (int i = 0; i < kChunks; ++i) {
          // rotate the chunk order so that a quarter-warp touches 8 distinct 16-byte bank groups
          const int c = BK == 32 ? ((i + r) & 7) : ((i + (r >> 1)) & 3);
          x[i] = raw[c];
        }
#pragma unroll
        for (int i = 0; i < kChunks; ++i) {
          const int c = BK == 32 ? ((i + r) & 7) : ((i + (r >> 1)) & 3);
          float4 h, l;
          h.x = p_tf32_hi(x[i].x); h.y = p_tf32_hi(x[i].y); h.z = p_tf32_hi(x[i].z); h.w = p_tf32_hi(x[i].w);
          l.x = p_tf32_rna(x[i].x - h.x); l.y = p_tf32_rna(x[i].y - h.y);
          l.z = p_tf32_rna(x[i].z - h.z); l.w = p_tf32_rna(x[i].w - h.w);
          nn = fmaf(x[i].x, x[i].x, nn); nn = fmaf(x[i].y, x[i].y, nn);
          nn = fmaf(x[i].z, x[i].z, nn); nn = fmaf(x[i].w, x[i].w, nn);
          sx += (x[i].x + x[i].y) + (x[i].z + x[i].w);
          raw[c] = h;
          lo[c] = l;
        }
        if (!(p.dbg & 8)) fence_proxy_async();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(&conv_bar[s], 0);
      }
      mbar_wait(&acc_free[a], (uint32_t)(((t >> 1) & 1) ^ 1));
      norms[(a * kMaxGroups + g) * BM + r] = nn;
      sums[(a * kMaxGroups + g) * BM + r] = sx;
      mbar_arrive(&norm_bar[a]);
    }
    if (p.prof && tid == 8 * 32) p.prof[blockIdx.x * 8 + 5] = (u64)w0;
  } else if (warp < 4) {
    // ===== epilogue ================================================================================
    const int r = tid;
    const uint32_t lane_base = (uint32_t)(warp * 32) << 16;
    long long wepi = 0;
    for (long long t = 0; t < my_tiles; ++t) {
      const int a = (int)(t & 1);
      const uint32_t ph = (uint32_t)((t >> 1) & 1);
      const long long row0 = ((pair + t * npairs) * 2 + rank) * BM;
      mbar_wait(&norm_bar[a], ph);
      PROF_WAIT(wepi, mbar_wait(&acc_full[a], ph));
      tc_fence_after();
      const bool row_ok = row0 + r < p.n;
      const uint32_t nrow = ~(uint32_t)(row0 + r);
      float nn_ = 0.f, sum_x = 0.f;
      for (int g = 0; g < G; ++g) {
        nn_ += norms[(a * kMaxGroups + g) * BM + r];
        sum_x += sums[(a * kMaxGroups + g) * BM + r];
      }
      const float inv_nx = rsqrtf(fmaxf(nn_, kEps * kEps));
      for (int c0 = 0; c0 < p.BN; c0 += 16) {
        if (c0 >= p.B || (p.dbg & 4)) break;
        uint32_t v[16];
        tmem_ld_32x32b_x16(tmem_base + lane_base + (uint32_t)(a * p.BN + c0), v);
        tmem_ld_wait();
#pragma unroll
        for (int j4 = 0; j4 < 16; j4 += 4) {
          const float4 qm = *reinterpret_cast<const float4*>(qmean + c0 + j4);
          const float4 iq = *reinterpret_cast<const float4*>(invq + c0 + j4);
          const float4 th = lds_volatile_f4(thrf + c0 + j4);
          const float qm_[4] = {qm.x, qm.y, qm.z, qm.w};
          const float iq_[4] = {iq.x, iq.y, iq.z, iq.w};
          const float th_[4] = {th.x, th.y, th.z, th.w};
#pragma unroll
          for (int jj = 0; jj < 4; ++jj) {
            const int q = c0 + j4 + jj;
            const float s_ = fmaf(qm_[jj], sum_x, __uint_as_float(v[j4 + jj])) * inv_nx * iq_[jj];
            // ties go to the exact key comparison (a lower row index wins)
            const bool surv = row_ok && q < p.B && !(s_ < th_[jj]);
            if (__any_sync(kFull, surv)) {
              const u64 key = surv ? (((u64)sim_to_ord(s_) << 32) | (u64)nrow) : 0ull;
              list_offer(lists + (size_t)q * p.k, thrf, locks + q, q, p.k, key, lane);
            }
          }
        }
      }
      tc_fence_before();
      mbar_arrive(&acc_free[a]);
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(&acc_empty[a], 0);
    }
    if (p.prof && tid == 0) p.prof[blockIdx.x * 8 + 6] = (u64)wepi;
    named_bar_sync(1, kEpiThreads);
    u64* out = p.part_keys + (size_t)blockIdx.x * p.B * p.k;
    for (int i = tid; i < p.B * p.k; i += kEpiThreads) out[i] = lists[i];
  }

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

bool pencode(void* fn, CUtensorMap* m, const float* base, long long rows, int dim, int ld, int bk,
             int box_rows, CUtensorMapL2promotion promo) {
  const cuuint64_t gdim[2] = {(cuuint64_t)dim, (cuuint64_t)rows};
  const cuuint64_t gstr[1] = {(cuuint64_t)ld * 4};
  const cuuint32_t box[2] = {(cuuint32_t)bk, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1, 1};
  return reinterpret_cast<EncodeTiledFn>(fn)(
             m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), gdim, gstr, box, estr,
             CU_TENSOR_MAP_INTERLEAVE_NONE, bk == 32 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B,
             promo, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

bool pencode_q(void* fn, CUtensorMap* m, const float* planes, int B, int dim, int q_ld, int bk, int box_rows) {
  const cuuint64_t gdim[3] = {(cuuint64_t)dim, (cuuint64_t)B, 2};
  const cuuint64_t gstr[2] = {(cuuint64_t)q_ld * 4, (cuuint64_t)B * q_ld * 4};
  const cuuint32_t box[3] = {(cuuint32_t)bk, (cuuint32_t)box_rows, 2};
  const cuuint32_t estr[3] = {1, 1, 1};
  return reinterpret_cast<EncodeTiledFn>(fn)(
             m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(planes), gdim, gstr, box, estr,
             CU_TENSOR_MAP_INTERLEAVE_NONE, bk == 32 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B,
             CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

constexpr size_t kSmemCap = 227 * 1024;

int ppad(int B) { return (B + 15) / 16 * 16; }

template <int BK>
int ppick_stages(int B, int k) {
  const int BN = ppad(B);
  for (int s = kMaxStages; s >= 2; --s)
    if ((size_t)pcarve<BK>(BN, B, k, s).total + 1024 <= kSmemCap) return s;
  return 0;
}

int env_bk() {
  static int bk = 0;
  if (!bk) {
    const char* v = getenv("CSLAM_NNS_K2_BK");
    bk = (v && atoi(v) == 32) ? 32 : 16;
  }
  return bk;
}
bool env_legacy() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("CSLAM_NNS_K2_LEGACY");
    v = (e && atoi(e) != 0) ? 1 : 0;
  }
  return v == 1;
}
int env_stage_cap() {
  const char* v = getenv("CSLAM_NNS_K2_STAGES");
  return (v && atoi(v) >= 2) ? atoi(v) : kMaxStages;
}

template <int BK>
cudaError_t plaunch(const K2Args& a, int num_sms, cudaStream_t stream, const char** err) {
  PDev p;
  p.n = a.n;
  p.tiles2 = (a.n + 2 * BM - 1) / (2 * BM);
  p.dim = a.dim;
  p.B = a.B;
  p.BN = ppad(a.B);
  p.k = a.k;
  p.stages = ppick_stages<BK>(a.B, a.k);
  if (p.stages > env_stage_cap()) p.stages = env_stage_cap();
  p.kblocks = (a.dim + BK - 1) / BK;
  {
    const char* v = getenv("CSLAM_NNS_K2_GROUPS");
    p.groups = v ? atoi(v) : 1;
    if (p.groups < 1) p.groups = 1;
    if (p.groups > kMaxGroups) p.groups = kMaxGroups;
    v = getenv("CSLAM_NNS_K2_DBG");
    p.dbg = v ? atoi(v) : 0;
  }
  p.inv_qn = a.inv_qn;
  p.q_mean = a.inv_qn + a.B;
  p.part_keys = a.part_keys;
  // kind::tf32, fp32 accumulate, K-major operands; N>>3 at bit 17, M>>4 at bit 24 (M = 256 over the pair)
  p.idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(p.BN >> 3) << 17) | ((uint32_t)((2 * BM) >> 4) << 24);
  if (p.stages < 2) {
    if (err) *err = "k2: batch * k does not fit shared memory";
    return cudaErrorInvalidConfiguration;
  }
  CUtensorMap ma, mq;
  // a.q_lo holds the two planes written by k2_prepare: q_hi [B][q_ld] then q_lo [B][q_ld]
  if (!pencode(a.encode_fn, &ma, a.db, a.n, a.dim, a.ld, BK, BM, CU_TENSOR_MAP_L2_PROMOTION_L2_128B) ||
      !pencode_q(a.encode_fn, &mq, a.q_lo, a.B, a.dim, a.q_ld, BK, p.BN / 2)) {
    if (err) *err = "k2: cuTensorMapEncodeTiled failed";
    return cudaErrorInvalidValue;
  }
  const size_t smem = pcarve<BK>(p.BN, p.B, p.k, p.stages).total + 1024;
  static bool attr_set[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64 || !attr_set[dev]) {
    cudaError_t e = cudaFuncSetAttribute(k2p_kernel<BK>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemCap);
    if (e != cudaSuccess) return e;
    if (dev >= 0 && dev < 64) attr_set[dev] = true;
  }
  const int grid = k2p_grid(a.n, num_sms);
  static u64* prof_buf = nullptr;
  static const bool want_prof = getenv("CSLAM_NNS_K2_PROF") != nullptr;
  p.prof = nullptr;
  if (want_prof) {
    if (!prof_buf) cudaMalloc(&prof_buf, 1024 * 8 * sizeof(u64));
    cudaMemsetAsync(prof_buf, 0, 1024 * 8 * sizeof(u64), stream);
    p.prof = prof_buf;
  }
  k2p_kernel<BK><<<grid, 256 + 128 * p.groups, smem, stream>>>(ma, mq, p);
  count_launch();
  if (want_prof) {
    static u64 hbuf[1024 * 8];
    cudaStreamSynchronize(stream);
    cudaMemcpy(hbuf, prof_buf, sizeof(hbuf), cudaMemcpyDeviceToHost);
    const char* names[8] = {"Xprod wait empty", "Qprod wait empty", "MMA wait qfull", "MMA wait conv",
                            "MMA wait acc_empty", "conv wait full", "epi wait acc_full", "total (Xprod)"};
    for (int r = 0; r < 2; ++r) {
      fprintf(stderr, "[k2 prof] rank %d (avg cycles over %d CTAs, %d stages):", r, grid / 2, p.stages);
      for (int i = 0; i < 8; ++i) {
        double acc = 0;
        for (int c = r; c < grid; c += 2) acc += (double)hbuf[c * 8 + i];
        fprintf(stderr, "  %s=%.0f", names[i], acc / (grid / 2));
      }
      fprintf(stderr, "\n");
    }
  }
  return cudaGetLastError();
}

}  // namespace

bool k2p_supported(int B, int k) {
  return B >= 1 && B <= kK2MaxBatch && k >= 1 && k <= 32 &&
         (env_bk() == 32 ? ppick_stages<32>(B, k) : ppick_stages<16>(B, k)) >= 2;
}

// CTAs launched (= partial lists written): one CTA pair per 256-row super tile, at most one pair per 2 SMs
int k2p_grid(long long n, int num_sms) {
  const long long tiles2 = (n + 2 * BM - 1) / (2 * BM);
  const long long pairs = num_sms / 2;
  return (int)(2 * (tiles2 < pairs ? (tiles2 < 1 ? 1 : tiles2) : pairs));
}

cudaError_t k2p_launch(const K2Args& a, int num_sms, cudaStream_t stream, const char** err) {
  return env_bk() == 32 ? plaunch<32>(a, num_sms, stream, err) : plaunch<16>(a, num_sms, stream, err);
}

}  // namespace cslam
