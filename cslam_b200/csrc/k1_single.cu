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
#include "common.cuh"
#include "kernels.h"

namespace cslam {

namespace {

constexpr int W = kK1Warps;
constexpr int U = 4;  // rows per lane group per pass

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

__device__ __forceinline__ float cos_from(float dot, float nn, float qn) {
  // x.q / (max(|x|,eps) * max(|q|,eps)), nns_matching.py:61
  return dot / (fmaxf(sqrtf(nn), kEps) * qn);
}

__global__ void __launch_bounds__(W * 32, 1) k1_kernel(const K1Dev p) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int S = p.stages;

  unsigned char* stage_base = smem;
  float4* q_s = reinterpret_cast<float4*>(smem + (size_t)W * S * p.stage_bytes);
  u64* lists = reinterpret_cast<u64*>(reinterpret_cast<unsigned char*>(q_s) + (size_t)p.ld4 * 16);
  uint64_t* bars = reinterpret_cast<uint64_t*>(lists + (size_t)W * p.k);
  float* red = reinterpret_cast<float*>(bars + W * S);  // [W] partial |q|^2, then [W] = qn
  int* flag = reinterpret_cast<int*>(red + W + 1);

  const long long GW = (long long)gridDim.x * W;
  const long long gw = (long long)blockIdx.x * W + warp;
  const int R = p.rows_per_tile;
  const long long total_tiles = (p.n + R - 1) / R;
  const long long my_tiles = gw < total_tiles ? (total_tiles - gw + GW - 1) / GW : 0;
  const long long my_items = my_tiles * p.n_chunks;

  uint64_t* my_bars = bars + warp * S;
  unsigned char* my_stage = stage_base + (size_t)warp * S * p.stage_bytes;

  auto issue = [&](long long item) {  // lane 0 only
    const int s = (int)(item % S);
    const long long tj = item / p.n_chunks;
    const int c = (int)(item - tj * p.n_chunks);
    const long long row0 = (gw + tj * GW) * R;
    const float4* src;
    uint32_t bytes;
    if (p.n_chunks == 1) {
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
    bulk_g2s(my_stage + (size_t)s * p.stage_bytes, src, bytes, &my_bars[s]);
  };

  // -- get the first copies in flight before anything else --------------------------------------
  if (lane == 0) {
    for (int s = 0; s < S; ++s) mbar_init(&my_bars[s], 1);
    mbar_fence_init();
    fence_proxy_async();
    for (int i = 0; i < S && i < my_items; ++i) issue(i);
  }

  // -- stage the query, compute max(|q|, eps) ----------------------------------------------------
  {
    float* qf = reinterpret_cast<float*>(q_s);
    float part = 0.f;
    for (int i = tid; i < p.ld4 * 4; i += W * 32) {
      float v = i < p.dim ? p.q[i] : 0.f;
      qf[i] = v;
      part = fmaf(v, v, part);
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) part += __shfl_xor_sync(kFull, part, off);
    if (lane == 0) red[warp] = part;
    u64* my_list0 = lists + (size_t)warp * p.k;
    warp_list_clear(my_list0, p.k, lane);
    __syncthreads();
    if (tid == 0) {
      float t = 0.f;
      for (int w = 0; w < W; ++w) t += red[w];
      red[W] = fmaxf(sqrtf(t), kEps);
    }
    __syncthreads();
  }
  const float qn = red[W];
  u64* my_list = lists + (size_t)warp * p.k;
  u64 thr = 0ull;

  const int lpr = 1 << p.lpr_log2;
  const int G = 32 >> p.lpr_log2;
  const int sub = lane & (lpr - 1);
  const int grp = lane >> p.lpr_log2;

  // -- stream the DB ----------------------------------------------------------------------------
  if (p.n_chunks == 1) {
    for (long long it = 0; it < my_items; ++it) {
      const int s = (int)(it % S);
      const uint32_t parity = (uint32_t)((it / S) & 1);
      const long long row0 = (gw + it * GW) * R;
      long long rows_ll = p.n - row0;
      const int rows = rows_ll > R ? R : (int)rows_ll;
      mbar_wait(&my_bars[s], parity);
      const float4* buf = reinterpret_cast<const float4*>(my_stage + (size_t)s * p.stage_bytes);

      for (int r0 = 0; r0 < rows; r0 += G * U) {
        float dot[U], nn[U];
        int rr[U];
        const float4* rowp[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          dot[u] = 0.f;
          nn[u] = 0.f;
          rr[u] = r0 + u * G + grp;
          int rc = rr[u] < rows ? rr[u] : rows - 1;
          rowp[u] = buf + (size_t)rc * p.ld4;
        }
#pragma unroll 2
        for (int j = sub; j < p.ld4; j += lpr) {
          const float4 qv = q_s[j];
          float4 x[U];
#pragma unroll
          for (int u = 0; u < U; ++u) x[u] = rowp[u][j];
#pragma unroll
          for (int u = 0; u < U; ++u) fma4(x[u], qv, dot[u], nn[u]);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          dot[u] = group_sum(dot[u], lpr);
          nn[u] = group_sum(nn[u], lpr);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const bool valid = (rr[u] < rows) && (sub == 0);
          const float s_ = cos_from(dot[u], nn[u], qn);
          const long long row = row0 + rr[u];
          if (p.all_sims) {
            if (valid) p.all_sims[row] = s_;
          } else {
            warp_list_offer(my_list, p.k, make_key(s_, (uint32_t)row), valid, thr, lane);
          }
        }
      }
      __syncwarp();
      if (lane == 0 && it + S < my_items) {
        fence_proxy_async();
        issue(it + S);
      }
    }
  } else {
    // rows longer than one stage: accumulate across column chunks (one row per tile, LPR = 32)
    float dot = 0.f, nn = 0.f;
    for (long long it = 0; it < my_items; ++it) {
      const int s = (int)(it % S);
      const uint32_t parity = (uint32_t)((it / S) & 1);
      const long long tj = it / p.n_chunks;
      const int c = (int)(it - tj * p.n_chunks);
      const long long row = gw + tj * GW;
      int f4 = p.ld4 - c * p.chunk_f4;
      if (f4 > p.chunk_f4) f4 = p.chunk_f4;
      mbar_wait(&my_bars[s], parity);
      const float4* buf = reinterpret_cast<const float4*>(my_stage + (size_t)s * p.stage_bytes);
      const float4* qc = q_s + (size_t)c * p.chunk_f4;
      float d0 = 0.f, n0 = 0.f, d1 = 0.f, n1 = 0.f;
      int j = lane;
      for (; j + 32 < f4; j += 64) {
        float4 x0 = buf[j], x1 = buf[j + 32];
        float4 q0 = qc[j], q1 = qc[j + 32];
        fma4(x0, q0, d0, n0);
        fma4(x1, q1, d1, n1);
      }
      if (j < f4) fma4(buf[j], qc[j], d0, n0);
      dot += d0 + d1;
      nn += n0 + n1;
      if (c == p.n_chunks - 1) {
        dot = group_sum(dot, 32);
        nn = group_sum(nn, 32);
        const float s_ = cos_from(dot, nn, qn);
        if (p.all_sims) {
          if (lane == 0) p.all_sims[row] = s_;
        } else {
          warp_list_offer(my_list, p.k, make_key(s_, (uint32_t)row), lane == 0, thr, lane);
        }
        dot = 0.f;
        nn = 0.f;
      }
      __syncwarp();
      if (lane == 0 && it + S < my_items) {
        fence_proxy_async();
        issue(it + S);
      }
    }
  }
  if (p.all_sims) return;

  // -- CTA merge -> scratch ---------------------------------------------------------------------
  __syncthreads();
  if (warp == 0) {
    for (int w = 1; w < W; ++w) warp_list_stream(my_list, p.k, thr, lists + (size_t)w * p.k, p.k, lane);
    for (int i = lane; i < p.k; i += 32) p.scratch[(size_t)blockIdx.x * p.k + i] = my_list[i];
    __threadfence();
    __syncwarp();
    if (lane == 0) {
      unsigned t = atomicAdd(p.ticket, 1u);
      *flag = (t == gridDim.x - 1);
    }
  }
  __syncthreads();
  if (!*flag) return;

  // -- last CTA: merge the grid's lists, decode, publish ------------------------------------------
  __threadfence();
  const long long total = (long long)gridDim.x * p.k;
  const long long per = (total + W - 1) / W;
  warp_list_clear(my_list, p.k, lane);
  thr = 0ull;
  {
    long long lo = per * warp, hi = lo + per;
    if (hi > total) hi = total;
    for (long long base = lo; base < hi; base += 32) {
      long long i = base + lane;
      u64 key = 0ull;
      if (i < hi) key = __ldcg(p.scratch + i);
      warp_list_offer(my_list, p.k, key, key != 0ull, thr, lane);
    }
  }
  __syncthreads();
  if (warp == 0) {
    for (int w = 1; w < W; ++w) warp_list_stream(my_list, p.k, thr, lists + (size_t)w * p.k, p.k, lane);
    for (int i = lane; i < p.k; i += 32) {
      const u64 key = my_list[i];
      const bool have = key != 0ull;
      const long long grow = (long long)key_row(key) * p.idx_stride + p.idx_offset;
      if (p.out_keys) p.out_keys[i] = have ? ((key & 0xffffffff00000000ull) | (u64)(~(uint32_t)grow)) : 0ull;
      if (p.out_idx) p.out_idx[i] = have ? grow : -1ll;
      if (p.out_sim) p.out_sim[i] = have ? key_sim(key) : -INFINITY;
    }
    if (lane == 0) *p.ticket = 0u;
  }
}

inline int ceil_log2(int v) {
  int l = 0;
  while ((1 << l) < v) ++l;
  return l;
}

}  // namespace

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

  const long long row_bytes = (long long)a.ld * 4;
  // query + per-warp lists + mbarriers (<= 16 stages) + reduction scratch
  const size_t fixed = (size_t)p.ld4 * 16 + (size_t)W * a.k * 8 + (size_t)W * 16 * 8 + 128;
  const size_t smem_cap = 227 * 1024;
  int stages = t.stages < 2 ? 2 : (t.stages > 16 ? 16 : t.stages);
  long long target = t.stage_bytes < 512 ? 512 : t.stage_bytes;
  target &= ~15ll;
  // shrink the ring until it fits beside the query and the lists
  while ((size_t)W * stages * target + fixed > smem_cap) {
    if (stages > 2) --stages;
    else if (target > 1024) target = (target / 2) & ~15ll;
    else return cudaErrorInvalidConfiguration;
  }
  if (row_bytes <= target) {
    p.rows_per_tile = (int)(target / row_bytes);
    p.n_chunks = 1;
    p.chunk_f4 = p.ld4;
    p.stage_bytes = (int)(((long long)p.rows_per_tile * row_bytes + 127) & ~127ll);
  } else {
    p.rows_per_tile = 1;
    p.chunk_f4 = (int)(target / 16);
    p.n_chunks = (p.ld4 + p.chunk_f4 - 1) / p.chunk_f4;
    p.stage_bytes = (int)((target + 127) & ~127ll);
  }
  p.stages = stages;
  p.lpr_log2 = ceil_log2(p.ld4);
  if (p.lpr_log2 > 5) p.lpr_log2 = 5;
  if (p.n_chunks > 1) p.lpr_log2 = 5;

  const long long tiles = (a.n + p.rows_per_tile - 1) / p.rows_per_tile;
  long long grid = (tiles + W - 1) / W;
  int cap = t.ctas > 0 ? t.ctas : num_sms;
  if (cap > k1_max_grid(num_sms)) cap = k1_max_grid(num_sms);
  if (grid > cap) grid = cap;
  if (grid < 1) grid = 1;

  const size_t smem = (size_t)W * p.stages * p.stage_bytes + fixed;
  static bool attr_set[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64 || !attr_set[dev]) {
    cudaError_t e = cudaFuncSetAttribute(k1_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)smem_cap);
    if (e != cudaSuccess) return e;
    if (dev >= 0 && dev < 64) attr_set[dev] = true;
  }
  k1_kernel<<<(unsigned)grid, W * 32, smem, stream>>>(p);
  count_launch();
  return cudaGetLastError();
}

}  // namespace cslam
