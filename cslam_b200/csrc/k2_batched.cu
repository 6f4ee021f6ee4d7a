// K2 — batched cosine-similarity search on the 5th-generation tensor cores.
//
// Replaces B calls of cslam/nns_matching.py:61-63 (one per query; the reference API is
// single-query) by ONE pass over the database: S[rows x B] = X[rows x dim] * Q[B x dim]^T on
// tcgen05.mma kind::tf32 with fp32 accumulators in TMEM, 3xTF32 split precision
//     x*q ~= x_lo*q_hi + x_hi*q_lo + x_hi*q_hi,   x_hi = tf32(x), x_lo = tf32(x - x_hi)
// so similarities agree with the reference's fp32 arithmetic to ~1e-6 relative, and a fused
// per-query top-k epilogue, so the rows x B similarity matrix never exists in memory.
//
// The tensor core accumulates in fp32 with truncation, a bias that grows with the number of
// accumulations (dim/8 * 3) and with the magnitude of the running sum.  To keep the running sum small
// the QUERY IS CENTRED:  x.q = x.(q - c) + c * sum(x),  c = mean(q).  The converter warps accumulate
// sum(x) next to |x|^2 (they touch every element anyway) and the epilogue adds c * sum(x) back in
// fp32, so the identity is exact and the accumulated part is ~10x smaller for non-negative data.
//
// Tensor-bound: algorithmic flops per launch = 2 * n * dim * B (issued: 3x that).
// Algorithmic HBM bytes per launch = n * ld * 4 (the DB is read exactly once per batch of <= 256).
//
// This file: the ROW-MAJOR one-CTA kernel for small batches (database rows on the MMA's M side, N = the batch
// padded to 16, so tensor work scales with the batch and a small batch runs at HBM speed), k2_prepare (query
// planes and per-query parameters, shared with the query-major pair kernel in k2_qm.cu) and the routing entry
// k2_launch.  Large batches take k2_qm.cu (screening pass + k6_rescore.cu).
//
// One persistent CTA per SM, tiles of 128 database rows, K blocks of 32 floats (one 128-byte swizzle row).
// Warp roles (512 threads):
//   warps 0-3          epilogue   tcgen05.ld the accumulator (lane == database row), a branch-free pass over 16
//                                 queries, barrier.red.or "did anything survive", owner warps insert the queued
//                                 survivors into the per-query sorted lists; full lists publish their k-th best
//   warps 4-7, 12-15   converter  two groups taking alternate K blocks: thread r owns tile row r, splits its 128
//                                 bytes into x_hi / x_lo, accumulates |x|^2 and sum(x), and writes the split row
//                                 with tcgen05.st into a ring of operand slots in TENSOR MEMORY (the split tiles
//                                 never exist in shared memory)
//   warp  8            producer   TMA (cp.async.bulk.tensor, 128B swizzle, evict-first) of the raw database tile
//   warp  9            MMA        the whole warp runs the loop, elect.sync picks the lane that issues 3 x 4
//                                 tcgen05.mma per K block with the A operand read from TMEM; tcgen05.commit frees
//                                 the operand slot / publishes the accumulator; owns the TMEM allocation
//   warp 10            Q producer both query planes of a K block as one 3-D TMA box (evict-last)
// The accumulator is double-buffered in TMEM so the epilogue of tile i overlaps the MMAs of tile i+1.
// Design history and measurements: DESIGN.md §3 "Row-major one-CTA kernel".
#include <cuda.h>

#include <cstdlib>
#include "common.cuh"
#include "kernels.h"

namespace cslam {
namespace {

constexpr int BM = 128;          // database rows per tile (UMMA M)
constexpr int BK = 32;           // floats per K block = 128 bytes = one swizzle row
constexpr int kEpiThreads = 128, kConvThreads = 128;
constexpr int kConvGroups = 2;   // converter warp groups, alternating K blocks (warps 4-7 and 12-15)
constexpr int kThreads = 512;    // warps 0-3 epilogue, 4-7 converter group 0, 8 X producer, 9 MMA issuer,
                                 // 10 Q producer, 11 idle, 12-15 converter group 1
constexpr int kQueueCap = 512;   // per epilogue warp and 16-query chunk: 4 queries x 128 rows
constexpr int kTmemCols = 512;
constexpr int kMaxXSlots = 8;    // raw database tiles in flight (shared memory)
constexpr int kMaxASlots = 4;    // split tiles waiting for the tensor core (tensor memory) + their query tiles
constexpr int kASlotCols = 2 * BK;  // x_hi (32 columns) then x_lo (32 columns)

struct K2Dev {
  long long n;
  int dim;
  int B;        // real queries
  int BN;       // padded to a multiple of 16 (UMMA N)
  int k;
  int xslots;   // raw-tile ring depth
  int aslots;   // operand ring depth
  int kblocks;  // ceil(dim / 32)
  int dbg;      // CSLAM_NNS_K2_DBG experiments (wrong results): 1 no conversion, 2 one product, 4 no scan,
                // 64 no grid-wide threshold sharing (results stay right)
  uint32_t idesc;
  const float* inv_qn;
  const float* q_mean;
  unsigned* gthr;  // [B] grid-wide lower bound of each query's k-th best similarity (ordered uint), 0 = none
  u64* part_keys;  // [grid][B][k]
};

// shared-memory carve-up (offsets from a 1024-byte aligned base)
struct Smem {
  uint32_t x_bytes, q_bytes;  // one raw-tile slot; one query slot (q_hi plane then q_lo plane)
  uint32_t q_off;             // start of the query ring (the raw-tile ring starts at 0)
  uint32_t lists_off, qpar_off, queue_off, qcnt_off, norm_off, bars_off, tmem_off, total;
};

__host__ __device__ inline Smem carve(int BN, int B, int k, int xslots, int aslots) {
  Smem s;
  s.x_bytes = BM * BK * 4;
  s.q_bytes = 2u * (uint32_t)BN * BK * 4;  // BN % 16 == 0 -> each plane is a multiple of 2048 bytes
  uint32_t o = s.x_bytes * (uint32_t)xslots;
  s.q_off = o;      o += s.q_bytes * (uint32_t)aslots;
  s.lists_off = o;  o += (uint32_t)B * k * 8;
  o = (o + 15u) & ~15u;
  s.qpar_off = o;   o += (uint32_t)BN * 16;  // per query {mean(q), 1/|q|, threshold (ordered uint), -}
  s.queue_off = o;  o += 4u * kQueueCap * 8;
  s.qcnt_off = o;   o += 16;
  s.norm_off = o;   o += 4u * kConvGroups * BM * 4;  // [2][groups][BM] partial |x|^2, then the same for sum(x)
  o = (o + 7u) & ~7u;
  s.bars_off = o;   o += 40u * 8;      // see the barrier table in the kernel
  s.tmem_off = o;   o += 16;
  s.total = o;
  return s;
}

__device__ __forceinline__ uint64_t make_desc(uint32_t saddr) {
  // K-major operand, 128-byte swizzle: 8-row atoms of 1024 bytes (SBO), LBO unused (1), version 1
  return (uint64_t)((saddr & 0x3ffffu) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) |
         ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
}

__device__ __forceinline__ float tf32_hi(float x) {
  return __uint_as_float(__float_as_uint(x) & 0xffffe000u);
}
__device__ __forceinline__ float tf32_rna(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}

__global__ void __launch_bounds__(kThreads, 1)
k2_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_q,
          const K2Dev p) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  // dynamic shared memory is only guaranteed 16-byte aligned: round up to the 1024 the swizzle needs.  Offset
  // the array (do NOT round-trip through uintptr_t: that makes every later access a generic LD.E / ST.E /
  // ATOM.E instead of LDS / STS / ATOMS, which cost the epilogue ~150 clk per similarity)
  unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const Smem L = carve(p.BN, p.B, p.k, p.xslots, p.aslots);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int SX = p.xslots, SA = p.aslots;

  u64* lists = reinterpret_cast<u64*>(smem + L.lists_off);
  uint4* qpar = reinterpret_cast<uint4*>(smem + L.qpar_off);
  u64* queue = reinterpret_cast<u64*>(smem + L.queue_off);
  int* qcnt = reinterpret_cast<int*>(smem + L.qcnt_off);
  float* norms = reinterpret_cast<float*>(smem + L.norm_off);
  float* sums = norms + 2 * kConvGroups * BM;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L.bars_off);
  uint64_t* x_full = bars;               // [SX] raw tile landed (TMA)                   X producer -> converter
  uint64_t* x_empty = bars + 8;          // [SX] raw tile is in registers                converter  -> X producer
  uint64_t* q_full = bars + 16;          // [SA] query planes of the K block landed      Q producer -> MMA
  uint64_t* a_full = bars + 20;          // [SA] x_hi / x_lo are in tensor memory        converter  -> MMA
  uint64_t* a_empty = bars + 24;         // [SA] the MMAs reading the slot are done      MMA -> converter, Q producer
  uint64_t* acc_full = bars + 28;        // [2]  accumulator complete
  uint64_t* acc_empty = bars + 30;       // [2]  accumulator drained by the epilogue
  uint64_t* norm_bar = bars + 32;        // [2]  row norms of the tile are in smem
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + L.tmem_off);

  // ---- one-time setup -----------------------------------------------------------------------------
  for (int i = tid; i < p.B * p.k; i += kThreads) lists[i] = 0ull;
  for (int i = tid; i < p.BN; i += kThreads)
    qpar[i] = make_uint4(__float_as_uint(i < p.B ? p.q_mean[i] : 0.f), __float_as_uint(i < p.B ? p.inv_qn[i] : 0.f),
                         0u, 0u);
  if (tid < 4) qcnt[tid] = 0;
  if (tid == 0) {
    for (int s = 0; s < SX; ++s) {
      mbar_init(&x_full[s], 1);
      mbar_init(&x_empty[s], kConvThreads / 32);
    }
    for (int s = 0; s < SA; ++s) {
      mbar_init(&q_full[s], 1);
      mbar_init(&a_full[s], kConvThreads);
      mbar_init(&a_empty[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&acc_full[a], 1);
      mbar_init(&acc_empty[a], kEpiThreads);
      mbar_init(&norm_bar[a], kConvThreads * kConvGroups);
    }
    mbar_fence_init();
  }
  if (warp == 9) tmem_alloc<kTmemCols>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // tensor-memory map: columns [0, 2*BN) two accumulators, then SA operand slots of 64 columns
  const uint32_t a_col0 = (uint32_t)(2 * p.BN);

  const long long tiles = (p.n + BM - 1) / BM;
  const long long my_tiles =
      (long long)blockIdx.x < tiles ? (tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;

  if (warp == 8) {
    // ===== X producer: raw database tiles, as deep a ring as shared memory allows ====================
    if (lane == 0) {
      tma_prefetch_desc(&map_a);
      int s = 0;
      uint32_t ph = 0;  // ring position / wrap parity advanced incrementally: a 64-bit div/mod per slot in
                        // a single-thread role costs ~500 clk of dependent latency
      for (long long t = 0; t < my_tiles; ++t) {
        const int row0 = (int)((blockIdx.x + t * gridDim.x) * BM);
        for (int kb = 0; kb < p.kblocks; ++kb, s = (s + 1 == SX ? 0 : s + 1), ph ^= (s == 0)) {
          mbar_wait(&x_empty[s], ph ^ 1);
          mbar_expect_tx(&x_full[s], L.x_bytes);
          tma_load_2d_hint(smem + (size_t)s * L.x_bytes, &map_a, kb * BK, row0, &x_full[s], kEvictFirst);
        }
      }
    }
  } else if (warp == 10) {
    // ===== Q producer: both query planes of the K block as ONE 3-D box (they stay L2 resident) =======
    if (lane == 0) {
      tma_prefetch_desc(&map_q);
      int s = 0;
      uint32_t ph = 0;
      for (long long t = 0; t < my_tiles; ++t) {
        for (int kb = 0; kb < p.kblocks; ++kb, s = (s + 1 == SA ? 0 : s + 1), ph ^= (s == 0)) {
          mbar_wait(&a_empty[s], ph ^ 1);
          mbar_expect_tx(&q_full[s], L.q_bytes);
          tma_load_3d_hint(smem + L.q_off + (size_t)s * L.q_bytes, &map_q, kb * BK, 0, 0, &q_full[s],
                           kEvictLast);
        }
      }
    }
  } else if (warp == 9) {
    // ===== MMA issuer: A = split database tile in TENSOR MEMORY, B = query planes in shared memory ====
    // The WHOLE warp runs the loop and one elected lane issues: inside an `if (lane == 0)` region ptxas
    // treats every operand as divergent and wraps each tcgen05.mma in an ELECT / R2UR.BROADCAST loop
    // (~90 clk per instruction, measured: 12 small MMAs per K block cost more than the HBM time of the block).
    {
      int s = 0;
      uint32_t ph = 0;
      const uint32_t lo_plane = (uint32_t)p.BN * BK * 4;
      const uint32_t q_ring = smem_u32(smem + L.q_off);
      for (long long t = 0; t < my_tiles; ++t) {
        const int a = (int)(t & 1);
        mbar_wait(&acc_empty[a], (uint32_t)(((t >> 1) & 1) ^ 1));
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(a * p.BN);
        for (int kb = 0; kb < p.kblocks; ++kb, s = (s + 1 == SA ? 0 : s + 1), ph ^= (s == 0)) {
          mbar_wait(&q_full[s], ph);
          mbar_wait(&a_full[s], ph);
          tc_fence_after();
          const uint32_t qs = q_ring + (uint32_t)s * L.q_bytes;
          const uint32_t at = tmem_base + a_col0 + (uint32_t)(s * kASlotCols);
          const uint64_t b_hi = make_desc(qs), b_lo = make_desc(qs + lo_plane);
          const uint32_t first = kb != 0 ? 1u : 0u;
          if (elect_one()) {
#pragma unroll
            for (int ks = 0; ks < BK / 8; ++ks) {
              // advancing 32 bytes along K inside the swizzle atom = +2 in the descriptor's address field
              umma_tf32_ts(d_tmem, at + BK + ks * 8, b_hi + 2 * ks, p.idesc, ks ? 1u : first);  // x_lo * q_hi
              if (p.dbg & 2) continue;
              umma_tf32_ts(d_tmem, at + ks * 8, b_lo + 2 * ks, p.idesc, 1u);                      // x_hi * q_lo
              umma_tf32_ts(d_tmem, at + ks * 8, b_hi + 2 * ks, p.idesc, 1u);                      // x_hi * q_hi
            }
            umma_commit(&a_empty[s]);  // implies tcgen05.fence::before_thread_sync
          }
          __syncwarp();
        }
        if (elect_one()) umma_commit(&acc_full[a]);
        __syncwarp();
      }
    }
  } else if ((warp >= 4 && warp < 8) || warp >= 12) {
    // ===== converter: raw tile (smem) -> x_hi, x_lo (tensor memory), row norms ========================
    // Two groups of 4 warps take alternate K blocks (group = kb & 1, so a row's partial sums do not depend on
    // its tile or shard): one group alone costs ~12 % over the HBM time of a block.
    // Thread r owns tile row r == TMEM lane r (warp w may only address lanes 32*(w%4)..+31, and 4 % 4 == 0).
    // The split tiles never touch shared memory: that removes the two 16 KB stores and the three 16 KB
    // operand fetches per K block that kept the shared-memory pipe ~90% busy (profiles/r1_k2rows_*).
    const int g = warp >= 12 ? 1 : 0;
    const int r = (warp & 3) * 32 + lane;
    const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
    int sx = 0, sa = 0;
    uint32_t phx = 0, pha = 0;
    for (long long t = 0; t < my_tiles; ++t) {
      const int a = (int)(t & 1);
      float nn = 0.f, sum = 0.f;
      for (int kb = 0; kb < p.kblocks; ++kb) {
        if ((kb & 1) != g) {  // the other group's block: only keep the ring positions in step
          sx = sx + 1 == SX ? 0 : sx + 1; phx ^= (sx == 0);
          sa = sa + 1 == SA ? 0 : sa + 1; pha ^= (sa == 0);
          continue;
        }
        uint32_t hi[BK], lo[BK];
        mbar_wait(&x_full[sx], phx);
        if (!(p.dbg & 1)) {
          const float4* raw = reinterpret_cast<const float4*>(smem + (size_t)sx * L.x_bytes + (size_t)r * 128);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            // 128-byte swizzle: logical 16-byte chunk j of row r sits at chunk j ^ (r & 7); a quarter-warp
            // therefore reads 8 distinct bank groups
            const float4 x = raw[j ^ (r & 7)];
            const float h0 = tf32_hi(x.x), h1 = tf32_hi(x.y), h2 = tf32_hi(x.z), h3 = tf32_hi(x.w);
            hi[4 * j + 0] = __float_as_uint(h0); hi[4 * j + 1] = __float_as_uint(h1);
            hi[4 * j + 2] = __float_as_uint(h2); hi[4 * j + 3] = __float_as_uint(h3);
            lo[4 * j + 0] = __float_as_uint(tf32_rna(x.x - h0)); lo[4 * j + 1] = __float_as_uint(tf32_rna(x.y - h1));
            lo[4 * j + 2] = __float_as_uint(tf32_rna(x.z - h2)); lo[4 * j + 3] = __float_as_uint(tf32_rna(x.w - h3));
            nn = fmaf(x.x, x.x, nn); nn = fmaf(x.y, x.y, nn);
            nn = fmaf(x.z, x.z, nn); nn = fmaf(x.w, x.w, nn);
            sum += (x.x + x.y) + (x.z + x.w);
          }
        }
        // the raw slot is in registers: hand it back to the producer before waiting for an operand slot
        __syncwarp();
        if (lane == 0) mbar_arrive(&x_empty[sx]);
        mbar_wait(&a_empty[sa], pha ^ 1);
        tc_fence_after();
        if (!(p.dbg & 1)) {
          const uint32_t at = tmem_base + lane_base + a_col0 + (uint32_t)(sa * kASlotCols);
          tmem_st_32x32b_x32(at, hi);
          tmem_st_32x32b_x32(at + BK, lo);
          tmem_st_wait();
        }
        tc_fence_before();
        mbar_arrive(&a_full[sa]);
        sx = sx + 1 == SX ? 0 : sx + 1; phx ^= (sx == 0);
        sa = sa + 1 == SA ? 0 : sa + 1; pha ^= (sa == 0);
      }
      // the epilogue of tile t-2 must be done with norms[a] before it is overwritten
      mbar_wait(&acc_empty[a], (uint32_t)(((t >> 1) & 1) ^ 1));
      norms[(a * kConvGroups + g) * BM + r] = nn;
      sums[(a * kConvGroups + g) * BM + r] = sum;
      mbar_arrive(&norm_bar[a]);
    }
  } else if (warp < 4) {
    // ===== epilogue: TMEM -> similarities -> per-query top-k ========================================
    const int r = tid;  // accumulator lane == tile row; warp w may only touch lanes 32w .. 32w+31
    const uint32_t lane_base = (uint32_t)(warp * 32) << 16;
    u64* myq = queue + (size_t)warp * kQueueCap;  // the queue this warp DRAINS (queries q & 3 == warp)
    for (long long t = 0; t < my_tiles; ++t) {
      const int a = (int)(t & 1);
      const uint32_t ph = (uint32_t)((t >> 1) & 1);
      const long long row0 = (blockIdx.x + t * gridDim.x) * (long long)BM;
      // grid-wide threshold sharing: every full list's k-th best is a proven lower bound of the final k-th
      // best, so other CTAs may drop anything BELOW it (ties are kept: a lower row index wins a tie).
      // Thread q refreshes query q once per tile; the load is issued before the waits that hide it.
      const unsigned g_new = (tid < p.B && !(p.dbg & 64)) ? __ldcg(p.gthr + tid) : 0u;
      mbar_wait(&norm_bar[a], ph);
      mbar_wait(&acc_full[a], ph);
      tc_fence_after();
      // (racing with the owner warp's update below is benign: both values are valid lower bounds)
      if (g_new > 1u && g_new - 1u > qpar[tid].z) qpar[tid].z = g_new - 1u;
      const bool row_ok = row0 + r < p.n;
      // 1 / max(|x|, eps)  ==  rsqrt(max(|x|^2, eps^2))
      const float inv_nx = rsqrtf(fmaxf(norms[(a * kConvGroups) * BM + r] + norms[(a * kConvGroups + 1) * BM + r],
                                        kEps * kEps));
      const float sum_x = sums[(a * kConvGroups) * BM + r] + sums[(a * kConvGroups + 1) * BM + r];
      for (int c0 = 0; c0 < ((p.dbg & 4) ? 0 : p.BN); c0 += 16) {
        uint32_t v[16];
        tmem_ld_32x32b_x16(tmem_base + lane_base + (uint32_t)(a * p.BN + c0), v);
        tmem_ld_wait();
        if (c0 >= p.B) continue;  // padding columns (uniform over the CTA)
        // branch-free pass over the chunk: the 16 parameter loads are independent and pipeline, survivors
        // only set a bit.  (Testing and pushing query by query serialises ~3 shared-memory latencies per
        // similarity on the one epilogue warp of each scheduler.)
        uint32_t o[16];
        uint32_t m = 0;
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const uint4 qp = qpar[c0 + j];
          const float s_ =
              fmaf(__uint_as_float(qp.x), sum_x, __uint_as_float(v[j])) * inv_nx * __uint_as_float(qp.y);
          o[j] = sim_to_ord(s_);
          m |= (o[j] > qp.z ? 1u : 0u) << j;
        }
        const int live = p.B - c0;  // real queries in this chunk
        m &= row_ok ? (live >= 16 ? 0xffffu : (1u << live) - 1u) : 0u;
        const bool pushed = m != 0;
        if (pushed) {
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            if (m & (1u << j)) {
              const int slot = atomicAdd(&qcnt[j & 3], 1);  // (c0 + j) & 3 == j & 3
              queue[(size_t)(j & 3) * kQueueCap + slot] = ((u64)o[j] << 32) | ((u64)(c0 + j) << 8) | (u64)r;
            }
          }
        }
        // one barrier tells every epilogue thread whether ANY row of the tile survived a threshold in this
        // chunk; once the lists have warmed up almost none does and the drain (and its second barrier) is skipped
        if (!named_bar_or(1, kEpiThreads, pushed)) continue;
        // drain: this warp owns the queries with q & 3 == warp, so no two warps touch one list
        const int cnt = qcnt[warp];
        for (int e = 0; e < cnt; ++e) {
          const u64 ent = myq[e];
          const int q = (int)((ent >> 8) & 0xff);
          const uint32_t row = (uint32_t)(row0 + (long long)(ent & 0x7f));
          const u64 key = (ent & 0xffffffff00000000ull) | (u64)(~row);
          u64* lst = lists + (size_t)q * p.k;
          u64 mine = lane < p.k ? lst[lane] : 0ull;
          const u64 last = __shfl_sync(kFull, mine, p.k - 1);
          if (key > last) {
            const int pos = __popc(__ballot_sync(kFull, lane < p.k && mine > key));
            const u64 up = __shfl_up_sync(kFull, mine, 1);
            if (lane < p.k) {
              if (lane > pos) mine = up;
              else if (lane == pos) mine = key;
              lst[lane] = mine;
            }
            const u64 nl = __shfl_sync(kFull, mine, p.k - 1);
            if (lane == 0 && nl != 0ull) {
              const uint32_t ord = (uint32_t)(nl >> 32);
              if (ord > qpar[q].z) qpar[q].z = ord;
              if (ord != 0xffffffffu && !(p.dbg & 64)) atomicMax(p.gthr + q, ord);  // NaN is never a bound
            }
          }
          __syncwarp();
        }
        __syncwarp();
        if (lane == 0) qcnt[warp] = 0;
        named_bar_sync(1, kEpiThreads);
      }
      tc_fence_before();
      mbar_arrive(&acc_empty[a]);
    }
    // this CTA's lists -> global scratch (all CTAs write, also the ones without tiles)
    named_bar_sync(1, kEpiThreads);
    u64* out = p.part_keys + (size_t)blockIdx.x * p.B * p.k;
    for (int i = tid; i < p.B * p.k; i += kEpiThreads) out[i] = lists[i];
  }

  // ---- teardown -----------------------------------------------------------------------------------
  __syncwarp();
  tc_fence_before();
  __syncthreads();
  if (warp == 9) {
    tc_fence_after();
    tmem_dealloc<kTmemCols>(tmem_base);
  }
}

// q -> c = mean(q), (q - c) split into q_hi (tf32-exact) and q_lo (tf32 residual), 1 / max(|q|, eps);
// one block per query
__global__ void k2_prepare_kernel(const float* __restrict__ q, int dim, int q_ld, float* q_hi,
                                  float* q_lo, float* inv_qn, float* q_mean) {
  const int b = blockIdx.x;
  const float* src = q + (size_t)b * q_ld;
  __shared__ float red[2][32];
  __shared__ float mean_s;
  float part = 0.f, sum = 0.f;
  for (int i = threadIdx.x; i < dim; i += blockDim.x) {
    const float x = src[i];
    part = fmaf(x, x, part);
    sum += x;
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    part += __shfl_xor_sync(kFull, part, off);
    sum += __shfl_xor_sync(kFull, sum, off);
  }
  if ((threadIdx.x & 31) == 0) {
    red[0][threadIdx.x >> 5] = part;
    red[1][threadIdx.x >> 5] = sum;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f, u = 0.f;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) {
      t += red[0][w];
      u += red[1][w];
    }
    const float invq = 1.0f / fmaxf(sqrtf(t), kEps);
    inv_qn[b] = invq;
    const float c = u / (float)dim;
    mean_s = (c == c && fabsf(c) < 3.0e38f) ? c : 0.f;  // NaN/Inf queries: no centring
    q_mean[b] = mean_s;
    const int B = (int)gridDim.x;
    // planes 2 and 3: the grid-wide per-query lower bounds of the k-th best similarity (ordered uint, 0 = none
    // yet) of the first pass (the only one, or the screening pass) and of the masked fallback pass
    reinterpret_cast<unsigned*>(inv_qn + kQParGthr * B)[b] = 0u;
    reinterpret_cast<unsigned*>(inv_qn + kQParGthrFull * B)[b] = 0u;
    // plane 4: E_q, a bound on |screened similarity - exact fp32 similarity| for ANY row (k6_rescore.cu has the
    // derivation).  Operands: the tensor core cuts x to tf32 (relative error < 2^-10), q - c was rounded to
    // tf32 above (<= 2^-11): together (2^-10 + 2^-11 + 2^-21) * sum |x_i| |q_i - c| <= 1.5 * 2^-10 |x| |q - c|.
    // Accumulation in tensor memory: dim / 8 instructions, each at most one fp32 ulp of the running sum for the
    // add and one for aligning its 8 products: dim / 4 ulps of at most |x| |q - c|.  The converter's fp32
    // |x|^2 and sum(x) (dim / 4 sequential adds each, the latter scaled by c), the epilogue's and the
    // re-scoring's roundings: the last two terms.  x 1.125 on top.
    const float dev2 = fmaxf(t - (float)dim * mean_s * mean_s, 0.f);      // |q - c|^2
    const float rho = sqrtf(dev2) * invq;
    const float fd = (float)dim;
    const float e = (1.4653e-3f + (0.25f * fd + 16.f) * 1.1920929e-7f) * rho +
                    (0.25f * fd + 8.f) * 5.9604645e-8f * (1.f + fabsf(mean_s) * sqrtf(fd) * invq) + 2e-6f;
    inv_qn[kQParEbound * B + b] = 1.125f * e;      // NaN for a NaN query: every comparison fails -> fallback
    if (b == 0) {
      unsigned* w = reinterpret_cast<unsigned*>(inv_qn + kQParPlanes * B);
      for (int i = 0; i < kQParTailWords; ++i) w[i] = 0u;   // fail mask [0..7], fail count [8]
    }
  }
  __syncthreads();
  const float c = mean_s;
  for (int i = threadIdx.x; i < q_ld; i += blockDim.x) {
    const float x = i < dim ? src[i] - c : 0.f;
    // round to nearest: |q - c - h| <= 2^-11 |q - c| (a cut would be 2^-10), which the screening pass's error
    // bound relies on; h is tf32-exact either way and the residual below completes it for the 3xTF32 pass
    const float h = tf32_rna(x);
    q_hi[(size_t)b * q_ld + i] = h;
    q_lo[(size_t)b * q_ld + i] = tf32_rna(x - h);
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

bool encode_map(void* fn, CUtensorMap* m, const float* base, long long rows, int dim, int ld,
                int box_rows, CUtensorMapL2promotion promo) {
  const cuuint64_t gdim[2] = {(cuuint64_t)dim, (cuuint64_t)rows};
  const cuuint64_t gstr[1] = {(cuuint64_t)ld * 4};
  const cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1, 1};
  return reinterpret_cast<EncodeTiledFn>(fn)(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2,
                                             const_cast<float*>(base), gdim, gstr, box, estr,
                                             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                             promo, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// both query planes as one 3-D map {dim, B, plane}; rows B..BN-1 of the box are out of bounds -> zero filled
bool encode_q_map(void* fn, CUtensorMap* m, const float* planes, int B, int dim, int q_ld, int box_rows) {
  const cuuint64_t gdim[3] = {(cuuint64_t)dim, (cuuint64_t)B, 2};
  const cuuint64_t gstr[2] = {(cuuint64_t)q_ld * 4, (cuuint64_t)B * q_ld * 4};
  const cuuint32_t box[3] = {(cuuint32_t)BK, (cuuint32_t)box_rows, 2};
  const cuuint32_t estr[3] = {1, 1, 1};
  return reinterpret_cast<EncodeTiledFn>(fn)(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3,
                                             const_cast<float*>(planes), gdim, gstr, box, estr,
                                             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                             CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

constexpr size_t kSmemCap = 227 * 1024;

int pad_batch(int B) { return (B + 15) / 16 * 16; }

// ring depths: as many operand slots as tensor memory holds next to the two accumulators (<= 4), then as
// many raw-tile slots as the remaining shared memory holds (<= 8, >= 3)
bool pick_slots(int B, int k, int* xslots, int* aslots) {
  const int BN = pad_batch(B);
  int amax = (kTmemCols - 2 * BN) / kASlotCols;
  if (amax > kMaxASlots) amax = kMaxASlots;
  for (int sa = amax; sa >= 2; --sa) {
    const size_t fixed = (size_t)carve(BN, B, k, 0, sa).total + 1024;
    if (fixed >= kSmemCap) continue;
    int sx = (int)((kSmemCap - fixed) / (BM * BK * 4));
    if (sx > kMaxXSlots) sx = kMaxXSlots;
    if (sx >= 3) {
      *xslots = sx;
      *aslots = sa;
      return true;
    }
  }
  return false;
}

}  // namespace

bool k2s_supported(int B, int k) {
  int sx, sa;
  return B >= 1 && B <= kRowsMaxBatch && k >= 1 && k <= 32 && pick_slots(B, k, &sx, &sa);
}

int k2s_grid(long long n, int num_sms) {
  const long long tiles = (n + BM - 1) / BM;
  return (int)(tiles < num_sms ? (tiles < 1 ? 1 : tiles) : num_sms);
}

// q_lo buffer layout: [q_hi plane (B x q_ld)][q_lo plane (B x q_ld)] — callers size it 2x
cudaError_t k2_prepare(const float* q, int B, int dim, int q_ld, float* q_planes, float* inv_qn,
                       cudaStream_t stream) {
  // inv_qn buffer layout: kQParPlanes planes of B words + kQParTailWords words (kernels.h) — callers size it so
  const int threads = dim >= 2048 ? 512 : (dim >= 512 ? 256 : 128);  // three latency-bound passes per block
  k2_prepare_kernel<<<B, threads, 0, stream>>>(q, dim, q_ld, q_planes, q_planes + (size_t)B * q_ld, inv_qn,
                                          inv_qn + B);
  count_launch();
  return cudaGetLastError();
}

cudaError_t k2s_launch(const K2Args& a, int num_sms, cudaStream_t stream, const char** err) {
  K2Dev p;
  p.n = a.n;
  p.dim = a.dim;
  p.B = a.B;
  p.BN = pad_batch(a.B);
  p.k = a.k;
  if (!pick_slots(a.B, a.k, &p.xslots, &p.aslots)) {
    if (err) *err = "k2: batch * k does not fit shared memory";
    return cudaErrorInvalidConfiguration;
  }
  static const int env_stages = getenv("CSLAM_NNS_K2_STAGES") ? atoi(getenv("CSLAM_NNS_K2_STAGES")) : 0;
  static const int env_dbg = getenv("CSLAM_NNS_K2_DBG") ? atoi(getenv("CSLAM_NNS_K2_DBG")) : 0;
  if (env_stages >= 2 && env_stages < p.xslots) p.xslots = env_stages;
  p.dbg = env_dbg;
  p.kblocks = (a.dim + BK - 1) / BK;
  p.inv_qn = a.inv_qn + kQParInv * a.B;
  p.q_mean = a.inv_qn + kQParMean * a.B;
  p.part_keys = a.part_keys;
  p.gthr = reinterpret_cast<unsigned*>(const_cast<float*>(a.inv_qn)) + kQParGthr * a.B;  // zeroed by k2_prepare
  // kind::tf32, fp32 accumulate, both operands K-major: c_format=F32(1)<<4, a/b_format=TF32(2)<<7/<<10,
  // N>>3 at bit 17, M>>4 at bit 24
  p.idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(p.BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
  CUtensorMap ma, mq;
  // a.q_lo holds the planes written by k2_prepare: [q_hi (B x q_ld)][q_lo (B x q_ld)]
  static const int env_promo = getenv("CSLAM_NNS_K2_PROMO") ? atoi(getenv("CSLAM_NNS_K2_PROMO")) : 128;
  const CUtensorMapL2promotion promo = env_promo == 256  ? CU_TENSOR_MAP_L2_PROMOTION_L2_256B
                                       : env_promo == 64 ? CU_TENSOR_MAP_L2_PROMOTION_L2_64B
                                       : env_promo == 0  ? CU_TENSOR_MAP_L2_PROMOTION_NONE
                                                         : CU_TENSOR_MAP_L2_PROMOTION_L2_128B;
  if (!encode_map(a.encode_fn, &ma, a.db, a.n, a.dim, a.ld, BM, promo) ||
      !encode_q_map(a.encode_fn, &mq, a.q_lo, a.B, a.dim, a.q_ld, p.BN)) {
    if (err) *err = "k2: cuTensorMapEncodeTiled failed";
    return cudaErrorInvalidValue;
  }
  const size_t smem = carve(p.BN, p.B, p.k, p.xslots, p.aslots).total + 1024;
  static bool attr_set[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64 || !attr_set[dev]) {
    cudaError_t e = cudaFuncSetAttribute(k2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemCap);
    if (e != cudaSuccess) return e;
    if (dev >= 0 && dev < 64) attr_set[dev] = true;
  }
  const int grid = k2s_grid(a.n, num_sms);
  k2_kernel<<<grid, kThreads, smem, stream>>>(ma, mq, p);
  count_launch();
  return cudaGetLastError();
}

}  // namespace cslam
