// K1 — single-query cosine similarity (a GEMV over the descriptor DB) fused with a warp-level
// top-k select.  Replaces cslam/nns_matching.py:61-63 (cosine_similarity + argsort + gather), which
// on the reference costs 6-8 ATen launches, ~5x the algorithmic bytes and a full radix sort.
//
// HBM-bound: algorithmic bytes per launch = n * ld * 4 (every DB row is read exactly once).
//
// Structure (one persistent CTA per SM, 8 warps):
//   * every warp owns a private ring of `stages` shared-memory buffers and private mbarriers;
//     its lane 0 streams row tiles into the ring with 1-D TMA bulk copies
//     (cp.async.bulk.shared.global -> SASS UBLKCP).  A row tile is one contiguous byte range because
//     the DB is row-major with a padded pitch, so no tensor map is needed.  Warp-private rings mean
//     no CTA-wide synchronisation anywhere in the streaming loop.
//   * lanes read the staged rows with conflict-free 128-bit ld.shared; LPR lanes cooperate on a row
//     (LPR = 32 for dim >= 128) and U = 4 rows are processed per pass so the query's 128-bit loads
//     are amortised; dot(x,q) and |x|^2 come from the same pass (no norm cache to go stale when a
//     caller writes matcher.data directly).
//   * per-warp sorted top-k list in shared memory (64-bit keys, see common.cuh); candidates below
//     the current k-th key are rejected with one ballot.
//   * CTA merge -> scratch[grid][k] -> atomic ticket; the last CTA merges the grid's lists, decodes
//     and writes the result (optionally straight into mapped pinned host memory).
#include <cstdlib>
#include <string.h>

#include "common.cuh"
#include "kernels.h"

namespace cslam {

namespace {

constexpr int W = kK1Warps;
constexpr int U = 4;  // rows per lane group per pass

#ifdef K1_PROF
// Development build only (make K1_PROF=1): per-CTA %globaltimer stamps -> tools/k1_timeline.py
__device__ unsigned long long g_k1_prof[8 * 512];
__device__ __forceinline__ void k1_stamp(int slot) {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  g_k1_prof[blockIdx.x * 8 + slot] = t;
}
#define K1_STAMP(cond, slot) do { if (cond) k1_stamp(slot); } while (0)
#else
#define K1_STAMP(cond, slot) do { } while (0)
#endif

struct K1Dev {
  const float4* db;
  long long n;
  int ld4;
  int dim;
  const float* q;
  int k;
  int rows_per_tile;
  int n_chunks;
  int chunk_f4;
  int lpr_log2;
  int stages;
  int stage_bytes;
  u64* scratch;
  unsigned* ticket;
  u64* out_keys;
  long long* out_idx;
  float* out_sim;
  long long idx_stride, idx_offset;
  float* all_sims;
  int dep_wait;            // execute griddepcontrol.wait after the stream, before the tail (chained searches, see k1_kernel)
  unsigned* done_flag;     // mapped host word (or null): receives done_value after the results are written
  unsigned done_value;
  long long keep_rows;  // rows below this are loaded with L2 evict-last, the rest evict-first (0: no hints)
  long long tiles_div, tiles_rem;  // total tiles / and % (grid * warps): a warp's tile count without a device division
};

__device__ __forceinline__ void fma4(const float4& x, const float4& q, float& dot, float& nn) {
  dot = fmaf(x.x, q.x, dot);
  dot = fmaf(x.y, q.y, dot);
  dot = fmaf(x.z, q.z, dot);
  dot = fmaf(x.w, q.w, dot);
  nn = fmaf(x.x, x.x, nn);
  nn = fmaf(x.y, x.y, nn);
  nn = fmaf(x.z, x.z, nn);
  nn = fmaf(x.w, x.w, nn);
}

__device__ __forceinline__ float group_sum(float v, int lpr) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1)
    if (off < lpr) v += __shfl_xor_sync(kFull, v, off);
  return v;
}


// sim = x.q / (max(|x|,eps) * max(|q|,eps)) with one MUFU.RSQ: max(sqrt(nn),eps) == sqrt(max(nn,eps^2))
__device__ __forceinline__ float cos_fast(float dot, float nn, float inv_qn) {
  return dot * rsqrtf(fmaxf(nn, kEps * kEps)) * inv_qn;
}

// NQ > 0: the query lives in registers (NQ float4 per lane), rows are whole inside one tile.
// NQ == 0: rows longer than a stage; the query is staged in shared memory and a row is accumulated
//          across column chunks.
template <int NQ>
__global__ void __launch_bounds__(W * 32, 1) k1_kernel(const K1Dev p) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int S = p.stages;
  // A search of ANOTHER database queued right behind this one with programmatic stream serialisation
  // (cslam_nns_search_multi: same query, disjoint scratch and outputs, nothing of ours is read by it) may
  // take over each SM as soon as this grid's CTA there exits, instead of waiting for the whole grid to
  // drain and a fresh launch.  A no-op when the next launch is an ordinary one.
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  K1_STAMP(tid == 0, 0);  // CTA start

  unsigned char* stage_base = smem;
  float4* q_s = reinterpret_cast<float4*>(smem + (size_t)W * S * p.stage_bytes);
  u64* lists = reinterpret_cast<u64*>(reinterpret_cast<unsigned char*>(q_s) +
                                      (NQ == 0 ? (size_t)p.ld4 * 16 : 0));
  uint64_t* bars = reinterpret_cast<uint64_t*>(lists + (size_t)W * p.k);
  float* red = reinterpret_cast<float*>(bars + W * S);
  int* flag = reinterpret_cast<int*>(red + W + 1);

  // tile t of the database belongs to warp t % GW of the grid; the launcher did the 64-bit divisions
  // (two of them here cost ~0.4 us before the first copy could be issued)
  const long long GW = (long long)gridDim.x * W;
  const long long gw = (long long)blockIdx.x * W + warp;
  const int R = p.rows_per_tile;
  const long long my_tiles = p.tiles_div + (gw < p.tiles_rem ? 1 : 0);
  const long long my_items = my_tiles * p.n_chunks;

  uint64_t* my_bars = bars + warp * S;
  unsigned char* my_stage = stage_base + (size_t)warp * S * p.stage_bytes;

  // lane 0 only.  The ring slot, tile index and column chunk are tracked incrementally by the callers:
  // a 64-bit div/mod per tile costs as much as the tile's arithmetic.
  auto issue = [&](int s, long long tj, int c) {
    const long long row0 = (gw + tj * GW) * R;
    const float4* src;
    uint32_t bytes;
    if (NQ > 0) {
      long long rows = p.n - row0;
      if (rows > R) rows = R;
      src = p.db + row0 * p.ld4;
      bytes = (uint32_t)(rows * p.ld4 * 16);
    } else {
      int f4 = p.ld4 - c * p.chunk_f4;
      if (f4 > p.chunk_f4) f4 = p.chunk_f4;
      src = p.db + row0 * p.ld4 + (long long)c * p.chunk_f4;
      bytes = (uint32_t)f4 * 16u;
    }
    mbar_expect_tx(&my_bars[s], bytes);
    if (p.keep_rows > 0)
      bulk_g2s_hint(my_stage + (size_t)s * p.stage_bytes, src, bytes, &my_bars[s],
                    row0 < p.keep_rows ? kEvictLast : kEvictFirst);
    else
      bulk_g2s(my_stage + (size_t)s * p.stage_bytes, src, bytes, &my_bars[s]);
  };

  // -- get the first copies in flight before anything else --------------------------------------
  K1_STAMP(tid == 0, 7);
  if (lane == 0) {
    for (int s = 0; s < S; ++s) mbar_init(&my_bars[s], 1);
    mbar_fence_init();
    fence_proxy_async();
    long long tj = 0;
    int c = 0;
    for (int i = 0; i < S && i < my_items; ++i) {
      issue(i, tj, c);
      if (++c == p.n_chunks) { c = 0; ++tj; }
    }
  }
  K1_STAMP(tid == 0, 4);  // ring primed (mbarrier init + first S bulk copies issued)
#ifdef K1_PROF
  if (tid == 1 && my_items > 2) {  // a plain load from a part of the database nobody has touched yet
    const float4* far = p.db + ((gw + 2 * GW) * p.rows_per_tile) * p.ld4 + 8;
    float4 v;
    asm volatile("ld.global.cv.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(far));
    if (v.x == 123456.f) g_k1_prof[0] = 0;  // keep the dependency
    k1_stamp(6);
  }
#endif
  // position of the NEXT item to issue (item index it + S while consuming item it)
  long long nx_tile = p.n_chunks == 1 ? (long long)S : (long long)(S / p.n_chunks);
  int nx_chunk = p.n_chunks == 1 ? 0 : S % p.n_chunks;
  u64* my_list = lists + (size_t)warp * p.k;
  u64 thr = 0ull, mine = 0ull;
  u64 big[4] = {0ull, 0ull, 0ull, 0ull};   // 32 < k <= 128: the warp's list, four keys per lane (offer_big)

  const int lpr = 1 << p.lpr_log2;
  const int G = 32 >> p.lpr_log2;
  const int sub = lane & (lpr - 1);
  const int grp = lane >> p.lpr_log2;

  if (NQ > 0) {
    // -- query into registers: lane holds q4[sub + i*lpr]; every lane group covers q once ---------
    float4 qreg[NQ > 0 ? NQ : 1];
    float part = 0.f;
#pragma unroll
    for (int i = 0; i < NQ; ++i) {
      const int j = sub + i * lpr;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (j < p.ld4) {
        if (4 * j + 3 < p.dim) {
          v = __ldg(reinterpret_cast<const float4*>(p.q) + j);
        } else {  // last, partial float4 of a dim that is not a multiple of 4
          if (4 * j + 0 < p.dim) v.x = __ldg(p.q + 4 * j + 0);
          if (4 * j + 1 < p.dim) v.y = __ldg(p.q + 4 * j + 1);
          if (4 * j + 2 < p.dim) v.z = __ldg(p.q + 4 * j + 2);
        }
      }
      qreg[i] = v;
      part = fmaf(v.x, v.x, part);
      part = fmaf(v.y, v.y, part);
      part = fmaf(v.z, v.z, part);
      part = fmaf(v.w, v.w, part);
    }
    part = group_sum(part, lpr);
    const float inv_qn = 1.0f / fmaxf(sqrtf(part), kEps);
    K1_STAMP(tid == 0 && inv_qn != 123456.f, 5);  // query in registers

    int s = 0;
    uint32_t parity = 0;
    for (long long it = 0; it < my_items; ++it) {
      const long long row0 = (gw + it * GW) * R;
      const long long rows_ll = p.n - row0;
      const int rows = rows_ll > R ? R : (int)rows_ll;
      mbar_wait(&my_bars[s], parity);
      K1_STAMP(tid == 0 && it == 0, 1);  // first bytes have landed
      const float4* buf = reinterpret_cast<const float4*>(my_stage + (size_t)s * p.stage_bytes);
      for (int r0 = 0; r0 < rows; r0 += G) {
        const int rr = r0 + grp;
        const int rc = rr < rows ? rr : rows - 1;
        const float4* rowp = buf + (size_t)rc * p.ld4 + sub;
        float4 x[NQ > 0 ? NQ : 1];
#pragma unroll
        for (int i = 0; i < NQ; ++i)
          x[i] = (sub + i * lpr < p.ld4) ? rowp[i * lpr] : make_float4(0.f, 0.f, 0.f, 0.f);
        float d0 = 0.f, n0 = 0.f, d1 = 0.f, n1 = 0.f;
#pragma unroll
        for (int i = 0; i < NQ; ++i) {
          if (i & 1) fma4(x[i], qreg[i], d1, n1);
          else fma4(x[i], qreg[i], d0, n0);
        }
        const float dot = group_sum(d0 + d1, lpr);
        const float nn = group_sum(n0 + n1, lpr);
        const float s_ = cos_fast(dot, nn, inv_qn);
        const bool valid = (rr < rows) && (sub == 0);
        const long long row = row0 + rr;
        if (p.all_sims) {
          if (valid) p.all_sims[row] = s_;
        } else {
          if (p.k <= 32) offer(my_list, p.k, make_key(s_, (uint32_t)row), valid, thr, mine, lane);
          else offer_big(big, p.k, make_key(s_, (uint32_t)row), valid, thr, lane);
        }
      }
      __syncwarp();
      if (lane == 0 && it + S < my_items) {
        fence_proxy_async();
        issue(s, it + S, 0);
      }
      if (++s == S) { s = 0; parity ^= 1u; }
    }
  } else {
    // -- stage the query in shared memory, compute 1/max(|q|, eps) --------------------------------
    float* qf = reinterpret_cast<float*>(q_s);
    float part = 0.f;
    for (int i = tid; i < p.ld4 * 4; i += W * 32) {
      const float v = i < p.dim ? p.q[i] : 0.f;
      qf[i] = v;
      part = fmaf(v, v, part);
    }
    part = group_sum(part, 32);
    if (lane == 0) red[warp] = part;
    __syncthreads();
    if (tid == 0) {
      float t = 0.f;
      for (int w = 0; w < W; ++w) t += red[w];
      red[W] = 1.0f / fmaxf(sqrtf(t), kEps);
    }
    __syncthreads();
    const float inv_qn = red[W];

    // one row per tile, accumulated across column chunks by the whole warp
    float dot = 0.f, nn = 0.f;
    int s = 0, c = 0;
    uint32_t parity = 0;
    long long tj = 0;
    for (long long it = 0; it < my_items; ++it) {
      const long long row = gw + tj * GW;
      int f4 = p.ld4 - c * p.chunk_f4;
      if (f4 > p.chunk_f4) f4 = p.chunk_f4;
      mbar_wait(&my_bars[s], parity);
      const float4* buf = reinterpret_cast<const float4*>(my_stage + (size_t)s * p.stage_bytes);
      const float4* qc = q_s + (size_t)c * p.chunk_f4;
      float d0 = 0.f, n0 = 0.f, d1 = 0.f, n1 = 0.f;
      int j = lane;
      for (; j + 32 < f4; j += 64) {
        const float4 x0 = buf[j], x1 = buf[j + 32];
        const float4 q0 = qc[j], q1 = qc[j + 32];
        fma4(x0, q0, d0, n0);
        fma4(x1, q1, d1, n1);
      }
      if (j < f4) fma4(buf[j], qc[j], d0, n0);
      dot += d0 + d1;
      nn += n0 + n1;
      if (c == p.n_chunks - 1) {
        dot = group_sum(dot, 32);
        nn = group_sum(nn, 32);
        const float s_ = cos_fast(dot, nn, inv_qn);
        if (p.all_sims) {
          if (lane == 0) p.all_sims[row] = s_;
        } else {
          if (p.k <= 32) offer(my_list, p.k, make_key(s_, (uint32_t)row), lane == 0, thr, mine, lane);
          else offer_big(big, p.k, make_key(s_, (uint32_t)row), lane == 0, thr, lane);
        }
        dot = 0.f;
        nn = 0.f;
      }
      __syncwarp();
      if (lane == 0 && it + S < my_items) {
        fence_proxy_async();
        issue(s, nx_tile, nx_chunk);
      }
      if (++nx_chunk == p.n_chunks) { nx_chunk = 0; ++nx_tile; }
      if (++c == p.n_chunks) { c = 0; ++tj; }
      if (++s == S) { s = 0; parity ^= 1u; }
    }
  }
  if (p.all_sims) return;
  K1_STAMP(tid == 0, 2);  // warp 0 of this CTA is through its share of the stream
  // Chained searches of the SAME database (programmatic dependent launch): this grid's CTAs were scheduled as the
  // previous search's CTAs left their SMs and have streamed their whole share of the database meanwhile.  Up to
  // here the kernel has only READ — rows, which change only through the library's own stream-ordered appends at
  // the END of the database (an append in between breaks the chain), and the query, which was complete before the
  // previous search started (anything queued between the two searches also breaks the chain) — and written only
  // shared memory and registers.  The ticket / scratch slot, the output and the done flag wait here until the
  // previous grid has completed and flushed: what overlaps with its tail is this search's launch latency,
  // prologue AND stream.  At most two searches are in flight (this grid's CTAs hold their SMs until the wait
  // returns; the launcher chains only grids of one CTA per SM).  A no-op for an ordinary launch.
  if (p.dep_wait) asm volatile("griddepcontrol.wait;" ::: "memory");

  // -- k == 1: the best key of the grid is one atomicMax per CTA; the CTA that draws the last ticket
  //    decodes and publishes it.  No scratch lists, no second merge level. ------------------------------
  if (p.k == 1) {
    if (lane == 0) my_list[0] = thr;
    __syncthreads();
    if (tid == 0) {
      u64 best = 0ull;
#pragma unroll
      for (int w = 0; w < W; ++w) best = lists[w] > best ? lists[w] : best;
      u64* gbest = reinterpret_cast<u64*>(p.ticket + 2);  // zero between launches, like the ticket
      if (best != 0ull) atomicMax(gbest, best);
      __threadfence();
      const unsigned t = atomicAdd(p.ticket, 1u);
      K1_STAMP(true, 3);  // ticket drawn
      if (t == gridDim.x - 1) {
        __threadfence();
        const u64 key = atomicExch(gbest, 0ull);
        const bool have = key != 0ull;
        const long long grow = (long long)key_row(key) * p.idx_stride + p.idx_offset;
        if (p.out_keys) p.out_keys[0] = have ? ((key & 0xffffffff00000000ull) | (u64)(~(uint32_t)grow)) : 0ull;
        if (p.out_idx) p.out_idx[0] = have ? grow : -1ll;
        if (p.out_sim) p.out_sim[0] = have ? key_sim(key) : -INFINITY;
        *p.ticket = 0u;
        if (p.done_flag) {  // host destination: the caller spins on this word instead of a stream synchronisation
          __threadfence_system();
          *reinterpret_cast<volatile unsigned*>(p.done_flag) = p.done_value;
        }
      }
    }
    return;
  }

  // -- 2 <= k <= 32: register lists all the way, merged with bitonic networks (warp_merge_desc): warp 0
  //    merges the CTA's 8 warp lists and publishes the result; in the CTA that draws the last ticket every
  //    warp merges 1/8 of the grid's CTA lists (loads batched 4 lists deep so the L2 round trips overlap),
  //    warp 0 merges those 8 and writes the result.  A merge costs the same ~12 shuffles whatever the lists
  //    hold (the one-key-at-a-time insertion this replaced cost ~200 clk per accepted key, ~30 of them per
  //    level: k = 10 at 100k x 512 took 14 us longer than k = 1). ------------------------------------------
  if (p.k <= 32) {
    const int k = p.k;
    if (lane < k) my_list[lane] = mine;
    __syncthreads();
    if (warp == 0) {
      u64 o[W - 1];
#pragma unroll
      for (int w = 1; w < W; ++w) o[w - 1] = lane < k ? lists[(size_t)w * k + lane] : 0ull;
#pragma unroll
      for (int w = 1; w < W; ++w) mine = warp_merge_desc(mine, o[w - 1], lane);
      if (lane < k) p.scratch[(size_t)blockIdx.x * k + lane] = mine;
      __threadfence();
      __syncwarp();
      unsigned t = 0;
      if (lane == 0) t = atomicAdd(p.ticket, 1u);
      t = __shfl_sync(kFull, t, 0);
      if (lane == 0) *flag = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (!*flag) return;
    __threadfence();
    // CTA lists warp, warp + W, ... of the grid
    mine = 0ull;
    const int G_ = (int)gridDim.x;
    for (int c0 = warp; c0 < G_; c0 += 4 * W) {
      u64 o[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int c = c0 + u * W;
        o[u] = (c < G_ && lane < k) ? __ldcg(p.scratch + (size_t)c * k + lane) : 0ull;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) mine = warp_merge_desc(mine, o[u], lane);
    }
    __syncthreads();  // everybody is past reading `lists` as the CTA merge's input
    if (lane < k) my_list[lane] = mine;
    __syncthreads();
    if (warp == 0) {
      u64 o[W - 1];
#pragma unroll
      for (int w = 1; w < W; ++w) o[w - 1] = lane < k ? lists[(size_t)w * k + lane] : 0ull;
#pragma unroll
      for (int w = 1; w < W; ++w) mine = warp_merge_desc(mine, o[w - 1], lane);
      if (lane < k) {
        const u64 key = mine;
        const bool have = key != 0ull;
        const long long grow = (long long)key_row(key) * p.idx_stride + p.idx_offset;
        if (p.out_keys) p.out_keys[lane] = have ? ((key & 0xffffffff00000000ull) | (u64)(~(uint32_t)grow)) : 0ull;
        if (p.out_idx) p.out_idx[lane] = have ? grow : -1ll;
        if (p.out_sim) p.out_sim[lane] = have ? key_sim(key) : -INFINITY;
      }
      if (lane == 0) *p.ticket = 0u;
      if (p.done_flag) {
        __threadfence_system();
        __syncwarp();
        if (lane == 0) *reinterpret_cast<volatile unsigned*>(p.done_flag) = p.done_value;
      }
    }
    return;
  }

  // -- 32 < k <= 128: the k <= 32 scheme with four keys per lane (offer_big / warp_merge_desc_big).  The
  //    shared-memory lists, rank selects and the tau pruning this replaced took 70 us at k = 33 against 44.5 us at
  //    k = 32 (100k x 512). ------------------------------------------------------------------------------------
  {
    const int k = p.k;
#pragma unroll
    for (int r = 0; r < 4; ++r)
      if (r * 32 + lane < k) my_list[r * 32 + lane] = big[r];
    __syncthreads();
    if (warp == 0) {
      for (int w = 1; w < W; ++w) {
        u64 o[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) o[r] = r * 32 + lane < k ? lists[(size_t)w * k + r * 32 + lane] : 0ull;
        warp_merge_desc_big(big, o, lane);
      }
#pragma unroll
      for (int r = 0; r < 4; ++r)
        if (r * 32 + lane < k) p.scratch[(size_t)blockIdx.x * k + r * 32 + lane] = big[r];
      __threadfence();
      __syncwarp();
      unsigned t = 0;
      if (lane == 0) t = atomicAdd(p.ticket, 1u);
      t = __shfl_sync(kFull, t, 0);
      if (lane == 0) *flag = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (!*flag) return;
    __threadfence();
    // the last CTA: warp w merges the CTA lists w, w + W, ... of the grid (two lists' loads in flight)
#pragma unroll
    for (int r = 0; r < 4; ++r) big[r] = 0ull;
    const int G_ = (int)gridDim.x;
    for (int c0 = warp; c0 < G_; c0 += 2 * W) {
      u64 o0[4], o1[4];
      const int c1 = c0 + W;
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const bool in = r * 32 + lane < k;
        o0[r] = in ? __ldcg(p.scratch + (size_t)c0 * k + r * 32 + lane) : 0ull;
        o1[r] = (in && c1 < G_) ? __ldcg(p.scratch + (size_t)c1 * k + r * 32 + lane) : 0ull;
      }
      warp_merge_desc_big(big, o0, lane);
      warp_merge_desc_big(big, o1, lane);
    }
    __syncthreads();  // everybody is past reading `lists` as the CTA merge's input
#pragma unroll
    for (int r = 0; r < 4; ++r)
      if (r * 32 + lane < k) my_list[r * 32 + lane] = big[r];
    __syncthreads();
    if (warp == 0) {
      for (int w = 1; w < W; ++w) {
        u64 o[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) o[r] = r * 32 + lane < k ? lists[(size_t)w * k + r * 32 + lane] : 0ull;
        warp_merge_desc_big(big, o, lane);
      }
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const int i = r * 32 + lane;
        if (i < k) {
          const u64 key = big[r];
          const bool have = key != 0ull;
          const long long grow = (long long)key_row(key) * p.idx_stride + p.idx_offset;
          if (p.out_keys) p.out_keys[i] = have ? ((key & 0xffffffff00000000ull) | (u64)(~(uint32_t)grow)) : 0ull;
          if (p.out_idx) p.out_idx[i] = have ? grow : -1ll;
          if (p.out_sim) p.out_sim[i] = have ? key_sim(key) : -INFINITY;
        }
      }
      if (lane == 0) *p.ticket = 0u;
    }
  }
}

inline int ceil_log2(int v) {
  int l = 0;
  while ((1 << l) < v) ++l;
  return l;
}

}  // namespace

#ifdef K1_PROF
extern "C" __attribute__((visibility("default"))) int cslam_k1_prof_read(unsigned long long* out, int ctas) {
  return (int)cudaMemcpyFromSymbol(out, g_k1_prof, sizeof(unsigned long long) * 8 * ctas);
}
#endif

int k1_max_grid(int num_sms) { return num_sms > 0 ? num_sms * 2 : 296; }

cudaError_t k1_launch(const K1Args& a, const K1Tuning& t, int num_sms, cudaStream_t stream) {
  K1Dev p;
  p.db = reinterpret_cast<const float4*>(a.db);
  p.n = a.n;
  p.ld4 = a.ld / 4;
  p.dim = a.dim;
  p.q = a.q;
  p.k = a.k;
  p.scratch = a.scratch;
  p.ticket = a.ticket;
  p.out_keys = a.out_keys;
  p.out_idx = a.out_idx;
  p.out_sim = a.out_sim;
  p.idx_stride = a.idx_stride;
  p.idx_offset = a.idx_offset;
  p.all_sims = a.all_sims;
  p.dep_wait = (a.overlap_prev && a.dep_wait) ? 1 : 0;
  p.done_flag = a.k <= 32 ? a.done_flag : nullptr;
  p.done_value = a.done_value;
  // Opt-in (off by default, not used for any reported number): keep the first CSLAM_NNS_L2_KEEP_MB of the
  // database resident in L2 across back-to-back searches of the same database.  B200, 100k x 512 (205 MB):
  // 5.58 -> 6.36 TB/s effective at 80 MB, 6.63 at 96 MB, thrashing back to 5.66 at 112 MB.
  static const long long keep_mb = getenv("CSLAM_NNS_L2_KEEP_MB") ? atoll(getenv("CSLAM_NNS_L2_KEEP_MB")) : 0;
  p.keep_rows = keep_mb > 0 ? keep_mb * 1000000ll / ((long long)a.ld * 4) : 0;

  const long long row_bytes = (long long)a.ld * 4;
  const size_t smem_cap = 227 * 1024;
  int stages = t.stages < 2 ? 2 : (t.stages > 16 ? 16 : t.stages);
  long long target = t.stage_bytes < 512 ? 512 : t.stage_bytes;
  target &= ~15ll;

  // lanes per row: 8 by default (a quarter-warp reads 128 contiguous bytes), fewer for tiny rows,
  // more when a lane would otherwise need more than 16 query float4 in registers
  int lpr = 8;
  while (lpr > 1 && lpr / 2 >= p.ld4) lpr /= 2;
  while ((p.ld4 + lpr - 1) / lpr > 16 && lpr < 32) lpr *= 2;
  int nq = (p.ld4 + lpr - 1) / lpr;
  const bool reg_path = row_bytes <= target && nq <= 16;

  // mbarriers (<= 16 stages) + reduction scratch + per-warp lists (+ the query on the chunked path)
  const size_t fixed = (size_t)W * a.k * 8 + (size_t)W * 16 * 8 + 128 + (size_t)a.k * 8 +
                       (reg_path ? 0 : (size_t)p.ld4 * 16);
  while ((size_t)W * stages * target + fixed > smem_cap) {
    if (stages > 2) --stages;
    else if (target > 1024) target = (target / 2) & ~15ll;
    else return cudaErrorInvalidConfiguration;
  }
  int NQ = 0;
  if (reg_path && row_bytes <= target) {
    p.rows_per_tile = (int)(target / row_bytes);
    const int G = 32 / lpr;
    if (p.rows_per_tile > G) p.rows_per_tile = p.rows_per_tile / G * G;  // whole passes
    p.n_chunks = 1;
    p.chunk_f4 = p.ld4;
    p.stage_bytes = (int)(((long long)p.rows_per_tile * row_bytes + 127) & ~127ll);
    NQ = nq <= 4 ? 4 : (nq <= 8 ? 8 : 16);
  } else {
    p.rows_per_tile = 1;
    p.chunk_f4 = (int)(target / 16);
    if (p.chunk_f4 > p.ld4) p.chunk_f4 = p.ld4;
    p.n_chunks = (p.ld4 + p.chunk_f4 - 1) / p.chunk_f4;
    p.stage_bytes = (int)(((long long)p.chunk_f4 * 16 + 127) & ~127ll);
    lpr = 32;
  }
  p.stages = stages;
  p.lpr_log2 = ceil_log2(lpr);

  const long long tiles = (a.n + p.rows_per_tile - 1) / p.rows_per_tile;
  long long grid = (tiles + W - 1) / W;
  int cap = t.ctas > 0 ? t.ctas : num_sms;
  if (cap > k1_max_grid(num_sms)) cap = k1_max_grid(num_sms);
  if (grid > cap) grid = cap;
  if (grid < 1) grid = 1;

  const size_t smem = (size_t)W * p.stages * p.stage_bytes + fixed;
  p.tiles_div = tiles / (grid * W);
  p.tiles_rem = tiles % (grid * W);
  static bool attr_set[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64 || !attr_set[dev]) {
    cudaError_t e = cudaSuccess;
    e = cudaFuncSetAttribute(k1_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_cap);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k1_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_cap);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k1_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_cap);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k1_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_cap);
    if (e != cudaSuccess) return e;
    if (dev >= 0 && dev < 64) attr_set[dev] = true;
  }
  const dim3 g((unsigned)grid), b(W * 32);
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = g;
  cfg.blockDim = b;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  // A chained search of the same database needs every SM to hold exactly one CTA of the previous search (then at
  // most two searches are ever in flight and the ticket / scratch ring of 4 cannot wrap) and a scan long enough
  // that the search is worth chaining at all (everything read before griddepcontrol.wait is rows [0, n) of this
  // search's own row count: an append between two searches breaks the chain in abi.cu).
  const bool chain_ok = !a.dep_wait || (grid == num_sms && p.tiles_div >= 4 * (long long)p.stages);
  if (!chain_ok) p.dep_wait = 0;
  if (a.overlap_prev && chain_ok) {  // see the top of k1_kernel
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
  }
  cudaError_t le;
  switch (NQ) {
    case 4: le = cudaLaunchKernelEx(&cfg, k1_kernel<4>, p); break;
    case 8: le = cudaLaunchKernelEx(&cfg, k1_kernel<8>, p); break;
    case 16: le = cudaLaunchKernelEx(&cfg, k1_kernel<16>, p); break;
    default: le = cudaLaunchKernelEx(&cfg, k1_kernel<0>, p); break;
  }
  count_launch();
  return le != cudaSuccess ? le : cudaGetLastError();
}

}  // namespace cslam
