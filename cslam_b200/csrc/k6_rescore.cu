// K6 — exact fp32 re-scoring of the screening pass's candidates, with a per-query completeness PROOF.
//
// The batched search needs fp32-faithful similarities (cslam/nns_matching.py:61), which on the tensor core costs
// three tf32 products per K slice (3xTF32).  The screening pass (k2_qm.cu, SCREEN) runs ONE product,
//     s~(x) = ( tf32(x) . tf32(q - c) + c * sum(x) ) / (max(|x|, eps) * max(|q|, eps)),
// and keeps, per query, 2 * (SMs / 2) thread-private lists of the kl >= k best rows BY s~ plus the shared bound
// tau_q every list publishes when it is full.  (kl > k matters: the list that holds the query's true neighbours
// — consecutive keyframes of a revisited place sit in consecutive rows, hence in ONE list — would otherwise
// evict rows right below its k-th entry and publish a bound right at the k-th best; with kl - k spare entries
// the bound sits kl - k candidates lower.)  Two facts make an exact answer out of that:
//
//  (1) |s~(x) - s(x)| <= E_q for EVERY row x, s being the fp32 similarity this kernel computes.  The operands are
//      cut to tf32 (relative error < 2^-10 each), so the dot product is off by at most
//      1.5 * 2^-10 * sum_i |x_i| |q_i - c| <= 1.5 * 2^-10 |x| |q - c| (x is cut to tf32 by the tensor core, q - c
//      was rounded to tf32 by k2_prepare; Cauchy-Schwarz); dividing by |x| |q| leaves 1.5 * 2^-10 * rho_q,
//      rho_q = |q - c| / |q| — independent of the row.  The remaining terms (fp32 accumulation in
//      tensor memory, the converter's fp32 |x|^2 and sum(x), epilogue and re-scoring roundings) are bounded in
//      k2_prepare, which writes E_q (x 1.125) next to the other per-query parameters.
//  (2) Every row that is in NONE of the query's lists has s~ <= tau_q: a row is only ever dropped, or evicted
//      from a list, against a full list's smallest entry or against the shared bound, and the shared bound at the
//      end of the pass is >= every full list's smallest entry (offer_key publishes it whenever it grows).
//
// Per query (one CTA): T = the k-th largest screened similarity in the union of the lists (radix select over the
// ordered-uint keys in shared memory).  The (>= k) candidates at or above T are re-scored exactly; the k-th largest
// of their exact similarities, L0, is a lower bound of the final k-th best, and by (1) no row with s~ < lim =
// L0 - E_q can reach it.  Every other candidate of the union with s~ >= lim (the "band") is re-scored as well.  If
// tau_q < lim, by (2) no row outside the union can reach L0 either, so the k best exact keys among the re-scored
// rows ARE the exact top-k of the whole database: they are written out (sorted, lower row first on equal similarity, like every
// other path).  Otherwise — tau_q too close (databases full of near-duplicates), more than kBandCap rows in the
// band, NaN anywhere — the query's bit is set in the fallback mask and the masked 3xTF32 pass queued right behind
// this kernel recomputes it; that pass exits at once when no bit is set.
//
// Latency-bound: B * (k + band) rows of ld * 4 bytes (C3: 256 x ~40 x 16 KB = 170 MB, ~1 % of the pass).
#include "common.cuh"
#include "kernels.h"

namespace cslam {
namespace {

constexpr int W6 = 16;           // warps per CTA (2 CTAs per SM: every query of a 256-batch is resident at once)
constexpr int kBandCap = 384;    // candidates one query may re-score before giving up
constexpr int kKeysPerThread = 10;   // the union of a query's lists: <= 148 parts x 32 entries = 4736 keys <= 512 x 10

struct K6Dev {
  const float* db;
  long long n;
  int ld, dim;
  const float* q;
  int q_ld;
  const float* inv_qn;
  const float* ebound;
  const unsigned* gthr;
  const u64* part_keys;
  int parts, B, k;
  u64* out_keys;
  long long* out_idx;
  float* out_sim;
  long long idx_stride, idx_offset;
  unsigned* fail_mask;
  unsigned* fail_count;
  unsigned long long* stats;
  int kl;  // capacity of the screening pass's lists (>= k)
  int seed;            // seed mode: see K6Args
  unsigned* gthr_w;    // seed mode: bound plane of the first pass (written) ...
  unsigned* gthr_clr;  // ... and of the sample pass (cleared)
  u64* debug;  // development: 64 words per query (top-k keys by s~, counters), or null
};

__device__ __forceinline__ void fma4x(const float4& x, const float4& q, float& dot, float& nn) {
  dot = fmaf(x.x, q.x, dot);
  dot = fmaf(x.y, q.y, dot);
  dot = fmaf(x.z, q.z, dot);
  dot = fmaf(x.w, q.w, dot);
  nn = fmaf(x.x, x.x, nn);
  nn = fmaf(x.y, x.y, nn);
  nn = fmaf(x.z, x.z, nn);
  nn = fmaf(x.w, x.w, nn);
}

constexpr int kRowsPerPass = 64;  // rows re-scored between two block barriers
constexpr int kMaxSegs = 16;      // warps that share one row (a row is spread over 32 .. 512 threads)

// Re-score `nr` candidates (keys cand[0 .. nr), nr <= kRowsPerPass) exactly; partial sums go to pd / pn, one entry per
// (row, warp of the row's thread group), and are added up in warp order by the caller.
//
// Work layout: the block's 512 threads are cut into groups of G = 2^g_log2 threads (32 <= G <= 512, G >= the row's
// float4 count when that is <= 512); thread `sub` of a group owns the float4 columns sub, sub + G, ... (NQ of
// them per chunk of G * NQ columns) of EVERY row its group re-scores, so its slice of the query stays in registers,
// and a group takes the rows grp, grp + groups, ... UR = 8 / NQ at a time: 8 independent 16-byte loads per thread
// are in flight whatever the row length (64 KB per block).  The version this replaced gave (row, 4 KB segment)
// pairs to warps one after the other and re-read the query with every row: 3 round trips for the first 10 rows
// of a 1M x 4096 search, 11 us where the loads themselves need ~2.
template <int NQ>
__device__ __forceinline__ void rescore_pass(const K6Dev& p, const float* __restrict__ q, const u64* cand, int nr,
                                             int g_log2, float (*pd)[kMaxSegs], float (*pn)[kMaxSegs]) {
  constexpr int UR = NQ == 1 ? 4 : 8 / NQ;   // NQ == 1 (rows up to 2 KB): groups are narrow and many, 4 rows each is plenty
  const int tid = threadIdx.x, lane = tid & 31;
  const int G = 1 << g_log2, sub = tid & (G - 1), grp = tid >> g_log2, groups = (W6 * 32) >> g_log2;
  const int wig = (tid >> 5) & ((G >> 5) - 1);   // warp within the group
  const int full4 = p.dim >> 2;                  // float4s that lie entirely inside the descriptor
  const int span = G * NQ;
  const float4* q4 = reinterpret_cast<const float4*>(q);
  const bool one_chunk = full4 <= span;
  float4 qv[NQ];
#pragma unroll
  for (int j = 0; j < NQ; ++j) {
    const int c = sub + j * G;
    qv[j] = c < full4 ? __ldg(q4 + c) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  for (int i0 = 0; (i0 * groups + grp) < nr; i0 += UR) {
    const float4* rp[UR];
    bool ok[UR];
    float dot[UR], nn[UR];
#pragma unroll
    for (int u = 0; u < UR; ++u) {
      const int r = (i0 + u) * groups + grp;
      const uint32_t row = r < nr ? key_row(cand[r]) : 0xffffffffu;
      ok[u] = (long long)row < p.n;   // never follow a row number the database does not have
      rp[u] = reinterpret_cast<const float4*>(p.db + (size_t)(ok[u] ? row : 0u) * p.ld);
      dot[u] = 0.f;
      nn[u] = 0.f;
    }
    for (int c0 = 0; c0 < full4 || c0 == 0; c0 += span) {
      if (!one_chunk) {
#pragma unroll
        for (int j = 0; j < NQ; ++j) {
          const int c = c0 + sub + j * G;
          qv[j] = c < full4 ? __ldg(q4 + c) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
      float4 x[UR][NQ];
#pragma unroll
      for (int u = 0; u < UR; ++u)
#pragma unroll
        for (int j = 0; j < NQ; ++j) {
          const int c = c0 + sub + j * G;
          x[u][j] = (ok[u] && c < full4) ? __ldg(rp[u] + c) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
      for (int u = 0; u < UR; ++u)
#pragma unroll
        for (int j = 0; j < NQ; ++j) fma4x(x[u][j], qv[j], dot[u], nn[u]);
      if (full4 == 0) break;
    }
    if (sub == 0 && (p.dim & 3)) {  // the last, partial float4 of a dim that is not a multiple of 4
#pragma unroll
      for (int u = 0; u < UR; ++u)
        if (ok[u]) {
          const float* rf = reinterpret_cast<const float*>(rp[u]);
          for (int i = full4 * 4; i < p.dim; ++i) {
            const float xv = __ldg(rf + i);
            dot[u] = fmaf(xv, __ldg(q + i), dot[u]);
            nn[u] = fmaf(xv, xv, nn[u]);
          }
        }
    }
#pragma unroll
    for (int u = 0; u < UR; ++u) {
      float d = dot[u], n = nn[u];
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) {
        d += __shfl_xor_sync(kFull, d, off);
        n += __shfl_xor_sync(kFull, n, off);
      }
      const int r = (i0 + u) * groups + grp;
      if (lane == 0 && r < nr) {
        pd[r][wig] = d;
        pn[r][wig] = n;
      }
    }
  }
}

// One instance of the largest value held in v[0 .. N) of the warp's lanes is removed and returned (0 when the warp
// holds nothing any more).  Values are ordered-uint similarities, 0 = empty.
template <int N>
__device__ __forceinline__ unsigned warp_pop_max(unsigned (&v)[N], int lane) {
  unsigned lm = 0u;
#pragma unroll
  for (int u = 0; u < N; ++u) lm = v[u] > lm ? v[u] : lm;
  const unsigned m = __reduce_max_sync(kFull, lm);
  if (m == 0u) return 0u;
  const unsigned has = __ballot_sync(kFull, lm == m);
  if (lane == __ffs(has) - 1) {
    bool done = false;
#pragma unroll
    for (int u = 0; u < N; ++u)
      if (!done && v[u] == m) {
        v[u] = 0u;
        done = true;
      }
  }
  return m;
}

__device__ __forceinline__ unsigned long long k6_now() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

__global__ void __launch_bounds__(W6 * 32, 2) k6_kernel(const K6Dev p) {
  extern __shared__ __align__(16) unsigned char smem6[];
  u64* keys = reinterpret_cast<u64*>(smem6);  // the union of this query's lists: [parts * kl]
  __shared__ u64 band[kBandCap];              // candidates to re-score
  __shared__ u64 exact[kBandCap];             // their exact keys
  __shared__ u64 outk[32];
  __shared__ unsigned wtop[W6 * 32];          // [warp][r]: the warp's r-th largest screened similarity
  __shared__ float pd[kRowsPerPass][kMaxSegs], pn[kRowsPerPass][kMaxSegs];   // per (row, segment) partial sums
  __shared__ unsigned sel_prefix, sel_want, n_keys;
  __shared__ int band_cnt, bad;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int b = blockIdx.x, k = p.k, kl = p.kl;
  const int total = p.parts * kl;
  unsigned long long ts[6] = {0ull, 0ull, 0ull, 0ull, 0ull, 0ull};   // development trace (p.debug): phase time stamps
  const bool trace = p.debug != nullptr && tid == 0;
  if (trace) ts[0] = k6_now();
  if (tid == 0) {
    band_cnt = 0;
    bad = 0;
    n_keys = 0u;
    sel_prefix = 0u;
    sel_want = (unsigned)k;   // seed mode: k is the rank of the seed (the full pass's list length)
    if (b == 0 && p.stats) atomicAdd(p.stats, (unsigned long long)p.B);
  }
  if (tid < 32) outk[tid] = 0ull;
  wtop[tid] = 0u;
  __syncthreads();

  // ---- 1. the union of the lists -> shared memory ------------------------------------------------------------------
  {
    unsigned mine_n = 0u;
    int saw_nan = 0;
    for (int i = tid; i < total; i += W6 * 32) {
      const int part = i / kl, j = i - part * kl;
      const u64 key = __ldcg(p.part_keys + ((size_t)part * p.B + b) * kl + j);
      keys[i] = key;
      mine_n += key != 0ull;
      saw_nan |= ((unsigned)(key >> 32) == 0xffffffffu);
    }
    if (mine_n) atomicAdd(&n_keys, mine_n);
    if (saw_nan) bad = 1;  // NaN similarities sort first in the reference: left to the 3xTF32 pass
  }
  __syncthreads();
  const bool enough = n_keys >= (unsigned)k;
  if (trace) ts[1] = k6_now();

  // ---- 2. T = the k-th largest screened similarity of the union (as an ordered uint; equal values count once
  //         each).  Every warp pops the k largest of the <= 10 values per lane it holds (one REDUX per pop), warp 0
  //         pops the k largest of those 16 * k.  (Four radix passes over a shared-memory histogram took 7 us of
  //         a 15 us launch at 1.25M x 256.) -------------------------------------------------------------------------
  if (enough) {
    unsigned v[kKeysPerThread];
#pragma unroll
    for (int u = 0; u < kKeysPerThread; ++u) {
      const int i = tid + u * (W6 * 32);
      v[u] = i < total ? (unsigned)(keys[i] >> 32) : 0u;
    }
    const unsigned want = sel_want;   // k (seed mode: the rank of the seed)
    for (unsigned r = 0; r < want; ++r) {
      const unsigned m = warp_pop_max(v, lane);
      if (lane == 0) wtop[warp * 32 + r] = m;
      if (m == 0u) break;   // this warp holds nothing more (the rest of its column was zeroed below)
    }
    __syncthreads();
    if (warp == 0) {
      unsigned w[W6];
#pragma unroll
      for (int u = 0; u < W6; ++u) w[u] = wtop[u * 32 + lane];   // lane r holds every warp's r-th largest
      unsigned t = 0u;
      for (unsigned r = 0; r < want; ++r) t = warp_pop_max(w, lane);
      if (lane == 0) sel_prefix = t;   // 0 cannot happen: `enough` non-empty keys exist
    }
    __syncthreads();
  }
  if (p.seed) {
    // the kl-th largest screened similarity of a SAMPLE of the rows: kl real rows reach it, so no row below it can
    // be among any list's kl best of the full pass, which starts with this bound instead of an empty one
    if (tid == 0) {
      if (enough && sel_prefix != 0xffffffffu) p.gthr_w[b] = sel_prefix;
      p.gthr_clr[b] = 0u;
    }
    return;
  }
  if (trace) ts[2] = k6_now();
  const float E = p.ebound[b];
  const unsigned T = sel_prefix;
  const float* q = p.q + (size_t)b * p.q_ld;
  const float inv_qn = p.inv_qn[b];

  // Re-score band[from, to) into exact[from, to): rescore_pass leaves per-(row, warp) partial sums, one thread per
  // row adds them up in warp order, so the result does not depend on scheduling: the same formula as the
  // single-query kernel, x.q * rsqrt(max(|x|^2, eps^2)) * (1 / max(|q|, eps))  (:61).
  const int full4 = p.dim >> 2;
  int g_log2 = 5;
  while (g_log2 < 9 && (1 << g_log2) < full4) ++g_log2;        // threads per row: 32 .. 512
  const int G = 1 << g_log2;
  const int nq = full4 <= G ? 1 : (full4 <= 2 * G ? 2 : 4);   // float4 per thread and chunk (rows beyond 8192 floats: several chunks)
  const int wpr = G >> 5;                                      // warps per row
  auto rescore = [&](int from, int to) {
    for (int r0 = from; r0 < to; r0 += kRowsPerPass) {
      const int nr = to - r0 < kRowsPerPass ? to - r0 : kRowsPerPass;
      switch (nq) {
        case 1: rescore_pass<1>(p, q, band + r0, nr, g_log2, pd, pn); break;
        case 2: rescore_pass<2>(p, q, band + r0, nr, g_log2, pd, pn); break;
        default: rescore_pass<4>(p, q, band + r0, nr, g_log2, pd, pn); break;
      }
      __syncthreads();
      if (tid < nr) {
        const uint32_t row = key_row(band[r0 + tid]);
        if ((long long)row >= p.n) {  // cannot happen with lists written by the screening pass
          exact[r0 + tid] = 0ull;
          bad = 1;
          if (p.stats) atomicAdd(p.stats + 7, 1ull);
        } else {
          float dot = 0.f, nn = 0.f;
          for (int c = 0; c < wpr; ++c) {
            dot += pd[tid][c];
            nn += pn[tid][c];
          }
          const float s_ = dot * rsqrtf(fmaxf(nn, kEps * kEps)) * inv_qn;
          exact[r0 + tid] = make_key(s_, row);
          if (!(s_ == s_)) bad = 1;
        }
      }
      __syncthreads();
    }
  };

  // ---- 3. the (>= k) candidates at or above T, re-scored: the k-th largest of their exact similarities, L0, is a
  //         lower bound of the final k-th best ------------------------------------------------------------------------
  if (enough) {
    for (int i = tid; i < total; i += W6 * 32) {
      const u64 key = keys[i];
      if (key != 0ull && (unsigned)(key >> 32) >= T) {
        const int at = atomicAdd(&band_cnt, 1);
        if (at < kBandCap) band[at] = key;
      }
    }
  }
  __syncthreads();
  const int ntop = band_cnt < kBandCap ? band_cnt : kBandCap;   // > k only through ties at T
  rescore(0, ntop);
  __syncthreads();
  if (tid < 32) outk[tid] = 0ull;
  __syncthreads();
  block_rank_select(exact, ntop, k, outk);   // outk[k - 1] = the k-th largest exact key so far
  __syncthreads();
  if (trace) ts[3] = k6_now();
  const float l0 = ntop >= k ? key_sim(outk[k - 1]) : __int_as_float(0x7fc00000);
  const float lim = l0 - E;   // by (1) no row whose screened similarity is below lim can reach l0
  __syncthreads();            // everybody has read outk before it is cleared again below

  // ---- 4. the band: every OTHER candidate of the union whose screened similarity is >= lim, re-scored too --------
  if (enough && lim == lim && band_cnt <= kBandCap) {
    for (int i = tid; i < total; i += W6 * 32) {
      const u64 key = keys[i];
      if (key != 0ull && (unsigned)(key >> 32) < T && key_sim(key) >= lim) {
        const int at = atomicAdd(&band_cnt, 1);
        if (at < kBandCap) band[at] = key;
      }
    }
  }
  if (tid < 32) outk[tid] = 0ull;
  __syncthreads();
  const int nband = band_cnt;
  const int m = nband < kBandCap ? nband : kBandCap;
  rescore(ntop, m);
  __syncthreads();
  if (trace) ts[4] = k6_now();

  // ---- 5. the proof, or the fallback -------------------------------------------------------------------------------
  const unsigned g = __ldcg(p.gthr + b);
  // g == 0: no list ever filled up, nothing was dropped anywhere: the union is the whole database
  const bool outside_ok = g == 0u || (ord_to_sim(g) < lim);
  const bool proven = !bad && enough && nband >= k && nband <= kBandCap && outside_ok && (lim == lim);
  if (p.debug && tid == 0) {
    u64* d = p.debug + (size_t)b * 64;
    for (int j = 0; j < 24; ++j) d[j] = j < m ? band[j] : 0ull;
    for (int j = 0; j < 24; ++j) d[24 + j] = j < m ? exact[j] : 0ull;
    d[48] = (u64)n_keys;
    d[49] = (u64)nband;
    d[50] = (u64)sel_prefix;
    d[51] = (u64)__float_as_uint(lim);
    d[57] = (u64)ntop;
    d[58] = (u64)__float_as_uint(l0);
    d[52] = (u64)g;
    d[53] = (u64)__float_as_uint(E);
    d[54] = (u64)bad;
    d[55] = (u64)p.n;
    d[56] = (u64)total;
    for (int j = 0; j < 5; ++j) d[59 + j] = ts[j];
  }
  if (!proven) {
    if (tid == 0) {
      atomicOr(p.fail_mask + (b >> 5), 1u << (b & 31));
      atomicAdd(p.fail_count, 1u);
      if (p.stats) {
        atomicAdd(p.stats + 1, 1ull);
        // why (first reason that applies): 2 NaN / bad row, 3 fewer than k candidates, 4 band overflow,
        // 5 the bound of the rows outside the lists is too close, 6 NaN bound
        const int why = bad ? 2 : (!enough || nband < k) ? 3 : nband > kBandCap ? 4 : !outside_ok ? 5 : 6;
        atomicAdd(p.stats + why, 1ull);
      }
    }
    return;
  }
  block_rank_select(exact, m, k, outk);
  __syncthreads();
  if (tid < k) {
    const u64 key = outk[tid];
    const bool have = key != 0ull;
    const long long grow = (long long)key_row(key) * p.idx_stride + p.idx_offset;
    const size_t o = (size_t)b * k + tid;
    if (p.out_keys) p.out_keys[o] = have ? ((key & 0xffffffff00000000ull) | (u64)(~(uint32_t)grow)) : 0ull;
    if (p.out_idx) p.out_idx[o] = have ? grow : -1ll;
    if (p.out_sim) p.out_sim[o] = have ? key_sim(key) : -INFINITY;
  }
}

}  // namespace

cudaError_t k6_rescore(const K6Args& a, cudaStream_t stream) {
  if (a.B <= 0 || a.k <= 0 || a.k > 32 || a.kl <= 0 || a.kl > 32 || (!a.seed && a.kl < a.k)) return cudaErrorInvalidValue;
  K6Dev p;
  p.db = a.db;
  p.n = a.n;
  p.ld = a.ld;
  p.dim = a.dim;
  p.q = a.q;
  p.q_ld = a.q_ld;
  p.inv_qn = a.qpar + kQParInv * a.B;
  p.ebound = a.qpar + kQParEbound * a.B;
  p.gthr = reinterpret_cast<const unsigned*>(a.qpar) + kQParGthr * a.B;
  p.part_keys = a.part_keys;
  p.parts = a.parts;
  p.B = a.B;
  p.k = a.k;
  p.out_keys = a.out_keys;
  p.out_idx = a.out_idx;
  p.out_sim = a.out_sim;
  p.idx_stride = a.idx_stride;
  p.idx_offset = a.idx_offset;
  unsigned* tail = reinterpret_cast<unsigned*>(const_cast<float*>(a.qpar)) + kQParPlanes * a.B;
  p.fail_mask = tail;
  p.fail_count = tail + 8;
  p.stats = a.stats;
  p.kl = a.kl;
  p.debug = a.debug;
  p.seed = a.seed ? 1 : 0;
  p.gthr_w = reinterpret_cast<unsigned*>(const_cast<float*>(a.qpar)) + kQParGthr * a.B;
  p.gthr_clr = reinterpret_cast<unsigned*>(const_cast<float*>(a.qpar)) + kQParGthrFull * a.B;
  if (a.seed) p.gthr = p.gthr_clr;
  const size_t smem = (size_t)a.parts * a.kl * 8;
  if (smem > 40 * 1024) return cudaErrorInvalidConfiguration;
  // static (~16 KB) + dynamic shared memory exceeds the 48 KB a kernel gets without opting in once kl = 32
  static bool attr_set[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64 || !attr_set[dev]) {
    cudaError_t e = cudaFuncSetAttribute(k6_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    if (e != cudaSuccess) return e;
    if (dev >= 0 && dev < 64) attr_set[dev] = true;
  }
  k6_kernel<<<(unsigned)a.B, W6 * 32, smem, stream>>>(p);
  count_launch();
  return cudaGetLastError();
}

}  // namespace cslam
