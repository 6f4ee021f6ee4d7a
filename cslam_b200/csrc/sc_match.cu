// ScanContext matcher: the lidar sibling of the dense matcher, same three-method plugin
// interface (cslam/lidar_pr/scancontext_matching.py:23,46,89).  Everything is IEEE double like the
// reference's numpy arrays.
//
//   sc_rk_kernel    step 1 (scancontext_matching.py:58-63): brute-force Euclidean nearest ring keys instead
//                   of a KD-tree rebuilt per search; per-thread sorted shortlist -> CTA shortlist
//   sc_dist_kernel  one CTA per shortlisted candidate: merges the CTA shortlists (every CTA redundantly, it
//                   is a few thousand entries), then step 2 (scancontext_utils.py:78-113): column norms,
//                   the sectors x sectors table of column cosines for every shift, and per shift the
//                   SEQUENTIAL sum over columns in the reference's order (j = 0..sectors-1), mean over the
//                   engaged columns, first argmax
// Memory: descriptors [capacity][rings*sectors] and ring keys [capacity][rings] in two VMM arenas.
#include <stdlib.h>
#include <string.h>

#include <cmath>
#include <string>

#include "../../include/cslam_nns.h"
#include "../../include/cslam_scancontext.h"
#include "arena.h"
#include "kernels.h"

namespace cslam {
namespace {

constexpr int kScThreads = 256;
constexpr int kScMaxCand = 32;
constexpr int kScMaxGrid = 296;

struct ScKey {
  double d;
  int idx;
  int pos;
};

__device__ __forceinline__ bool sc_less(double d0, int i0, double d1, int i1) {
  return d0 < d1 || (d0 == d1 && i0 < i1);
}

// block-wide arg-min over (d, idx) pairs held one per thread; result valid in thread 0
__device__ __forceinline__ ScKey sc_block_min(ScKey v, ScKey* red) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    const double od = __shfl_xor_sync(0xffffffffu, v.d, off);
    const int oi = __shfl_xor_sync(0xffffffffu, v.idx, off);
    const int op = __shfl_xor_sync(0xffffffffu, v.pos, off);
    if (sc_less(od, oi, v.d, v.idx)) {
      v.d = od;
      v.idx = oi;
      v.pos = op;
    }
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < (int)(blockDim.x >> 5); ++w)
      if (sc_less(red[w].d, red[w].idx, v.d, v.idx)) v = red[w];
    red[0] = v;
  }
  __syncthreads();
  return red[0];
}

// C rounds of arg-min over the M (d, idx) pairs in shared memory; out[r] = r-th smallest (idx -1 if none)
__device__ void sc_select(double* sd, int* si, int M, int C, double* out_d, int* out_i, ScKey* red) {
  for (int r = 0; r < C; ++r) {
    ScKey best;
    best.d = INFINITY;
    best.idx = 0x7fffffff;
    best.pos = -1;
    for (int e = threadIdx.x; e < M; e += blockDim.x) {
      if (si[e] >= 0 && sc_less(sd[e], si[e], best.d, best.idx)) {
        best.d = sd[e];
        best.idx = si[e];
        best.pos = e;
      }
    }
    best = sc_block_min(best, red);
    if (threadIdx.x == 0) {
      out_d[r] = best.pos >= 0 ? best.d : INFINITY;
      out_i[r] = best.pos >= 0 ? best.idx : -1;
      if (best.pos >= 0) si[best.pos] = -1;  // taken
    }
    __syncthreads();
  }
}

__global__ void __launch_bounds__(kScThreads) sc_rk_kernel(const double* __restrict__ rk, long long n, int R,
                                                            const double* __restrict__ rq, int C,
                                                            double* __restrict__ part_d, int* __restrict__ part_i) {
  extern __shared__ __align__(16) unsigned char sc_smem[];
  double* sd = reinterpret_cast<double*>(sc_smem);            // [threads * C]
  int* si = reinterpret_cast<int*>(sd + kScThreads * C);      // [threads * C]
  ScKey* red = reinterpret_cast<ScKey*>(si + kScThreads * C);  // [8]
  double* q = reinterpret_cast<double*>(red + 8);              // [R]
  for (int r = threadIdx.x; r < R; r += blockDim.x) q[r] = rq[r];
  __syncthreads();
  double ld[kScMaxCand];
  int li[kScMaxCand];
  int cnt = 0;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const double* row = rk + i * R;
    double d2 = 0.0;
    for (int r = 0; r < R; ++r) {
      const double t = row[r] - q[r];
      d2 = fma(t, t, d2);
    }
    if (cnt < C || d2 < ld[cnt - 1]) {   // rows come in increasing i: an equal distance goes behind
      int j = cnt < C ? cnt : C - 1;
      while (j > 0 && ld[j - 1] > d2) {
        ld[j] = ld[j - 1];
        li[j] = li[j - 1];
        --j;
      }
      ld[j] = d2;
      li[j] = (int)i;
      if (cnt < C) ++cnt;
    }
  }
  for (int j = 0; j < C; ++j) {
    sd[threadIdx.x * C + j] = j < cnt ? ld[j] : INFINITY;
    si[threadIdx.x * C + j] = j < cnt ? li[j] : -1;
  }
  __syncthreads();
  sc_select(sd, si, kScThreads * C, C, part_d + (size_t)blockIdx.x * C, part_i + (size_t)blockIdx.x * C, red);
}

__global__ void __launch_bounds__(kScThreads) sc_dist_kernel(const double* __restrict__ sc, int R, int S,
                                                             const double* __restrict__ query, int C, int parts,
                                                             const double* __restrict__ part_d,
                                                             const int* __restrict__ part_i,
                                                             long long* __restrict__ out_row,
                                                             double* __restrict__ out_dist,
                                                             int* __restrict__ out_yaw) {
  extern __shared__ __align__(16) unsigned char sc_smem[];
  const int M = parts * C;
  double* sd = reinterpret_cast<double*>(sc_smem);   // [M]
  double* fin_d = sd + M;                            // [C]
  double* c1 = fin_d + C;                            // [R*S] candidate
  double* c2 = c1 + R * S;                           // [R*S] query
  double* n1 = c2 + R * S;                           // [S] column norms of the candidate
  double* n2 = n1 + S;                               // [S] of the query
  double* cs = n2 + S;                               // [S*S] cosine of (candidate column a, query column j)
  double* sim = cs + S * S;                          // [S] per shift
  ScKey* red = reinterpret_cast<ScKey*>(sim + S);    // [8]
  int* si = reinterpret_cast<int*>(red + 8);         // [M]
  int* fin_i = si + M;                               // [C]
  unsigned char* any1 = reinterpret_cast<unsigned char*>(fin_i + C);  // [S]
  unsigned char* any2 = any1 + S;                                      // [S]

  for (int e = threadIdx.x; e < M; e += blockDim.x) {
    sd[e] = part_d[e];
    si[e] = part_i[e];
  }
  __syncthreads();
  sc_select(sd, si, M, C, fin_d, fin_i, red);
  const int c = blockIdx.x;
  const int row = fin_i[c];
  if (row < 0) {
    if (threadIdx.x == 0) {
      out_row[c] = -1;
      out_dist[c] = 1.0;
      out_yaw[c] = 0;
    }
    return;
  }
  const double* cand = sc + (size_t)row * R * S;
  for (int e = threadIdx.x; e < R * S; e += blockDim.x) {
    c1[e] = cand[e];
    c2[e] = query[e];
  }
  __syncthreads();
  // column norms and "any non-zero" flags (scancontext_utils.py:94, np.any: NaN counts as non-zero)
  for (int j = threadIdx.x; j < 2 * S; j += blockDim.x) {
    const double* m = j < S ? c1 : c2;
    const int col = j < S ? j : j - S;
    double acc = 0.0;
    bool any = false;
    for (int r = 0; r < R; ++r) {
      const double v = m[r * S + col];
      acc = fma(v, v, acc);
      any = any || (v != 0.0);
    }
    if (j < S) {
      n1[col] = sqrt(acc);
      any1[col] = any;
    } else {
      n2[col] = sqrt(acc);
      any2[col] = any;
    }
  }
  __syncthreads();
  // cosine table: candidate column a against query column j (:99)
  for (int e = threadIdx.x; e < S * S; e += blockDim.x) {
    const int a = e / S, j = e - a * S;
    double dot = 0.0;
    for (int r = 0; r < R; ++r) dot = fma(c1[r * S + a], c2[r * S + j], dot);
    cs[e] = dot / (n1[a] * n2[j]);
  }
  __syncthreads();
  // shift i+1 (np.roll by one column per iteration, :87): rolled[:, j] = candidate[:, (j - (i+1)) mod S];
  // sum over j in the reference's order, mean over the engaged columns (:90-107)
  for (int i = threadIdx.x; i < S; i += blockDim.x) {
    double sum = 0.0;
    int engaged = 0;
    for (int j = 0; j < S; ++j) {
      int a = j - (i + 1);
      a %= S;
      if (a < 0) a += S;
      if (!any1[a] || !any2[j]) continue;
      sum = sum + cs[a * S + j];
      ++engaged;
    }
    sim[i] = engaged == 0 ? 0.0 : sum / (double)engaged;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    // np.argmax / np.max (:109-110): first maximum; a NaN entry wins like numpy's max propagates it
    int best = 0;
    double bv = sim[0];
    for (int i = 1; i < S; ++i) {
      if (bv != bv) break;
      if (sim[i] > bv || sim[i] != sim[i]) {
        bv = sim[i];
        best = i;
      }
    }
    out_row[c] = row;
    out_dist[c] = 1.0 - bv;
    out_yaw[c] = best + 1;
  }
}

struct HostBuf {
  void* p = nullptr;
  size_t bytes = 0;
  bool ensure(size_t want) {
    if (want <= bytes) return true;
    if (p) cudaFreeHost(p);
    p = nullptr;
    bytes = 0;
    if (cudaHostAlloc(&p, want, cudaHostAllocDefault) != cudaSuccess) {
      cudaGetLastError();
      return false;
    }
    bytes = want;
    return true;
  }
};

std::string g_sc_create_error;

}  // namespace
}  // namespace cslam

using namespace cslam;

struct cslam_sc_s {
  int device = 0;
  int R = 20, S = 60, C = 10;
  long long n = 0, capacity = 0, initial = 1000;
  Arena desc, keys;
  std::string err;
  void* dev_tmp = nullptr;   // query descriptor + ring key, partial shortlists, outputs
  size_t dev_tmp_bytes = 0;
  HostBuf pin;               // staging for add / query / results
};

namespace {

int sc_fail(cslam_sc_t h, int code, const std::string& msg) {
  if (h) h->err = msg;
  else g_sc_create_error = msg;
  return code;
}
#define SCK(h, call)                                                                                   \
  do {                                                                                                 \
    cudaError_t e__ = (call);                                                                          \
    if (e__ != cudaSuccess) {                                                                          \
      cudaGetLastError();                                                                              \
      return sc_fail(h, CSLAM_NNS_ECUDA, std::string(#call) + ": " + cudaGetErrorString(e__));         \
    }                                                                                                  \
  } while (0)

struct ScDeviceGuard {
  int prev = -1;
  explicit ScDeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    if (prev != dev) cudaSetDevice(dev);
    else prev = -1;
  }
  ~ScDeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

int sc_grow(cslam_sc_t h, long long rows, cudaStream_t stream) {
  long long cap = h->capacity > 0 ? h->capacity : h->initial;
  while (cap < rows) cap *= 2;  // scancontext_matching.py:33-37 doubles
  if (cap == h->capacity) return 0;
  std::string err;
  if (!h->desc.ensure(h->device, (size_t)cap * h->R * h->S * 8, stream, &err) ||
      !h->keys.ensure(h->device, (size_t)cap * h->R * 8, stream, &err))
    return sc_fail(h, CSLAM_NNS_ENOMEM, err);
  h->capacity = cap;
  return 0;
}

}  // namespace

#pragma GCC visibility push(default)
extern "C" {

const char* cslam_sc_last_error(cslam_sc_t h) { return h ? h->err.c_str() : g_sc_create_error.c_str(); }

int cslam_sc_create(int device, int rings, int sectors, int num_candidates, int64_t initial_capacity,
                    cslam_sc_t* out) {
  if (!out) return sc_fail(nullptr, CSLAM_NNS_EINVAL, "out is NULL");
  *out = nullptr;
  if (rings < 1 || sectors < 1 || rings > 256 || sectors > 128 || num_candidates < 1 ||
      num_candidates > kScMaxCand)
    return sc_fail(nullptr, CSLAM_NNS_EINVAL, "unsupported shape or shortlist size");
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || device < 0 || device >= count) {
    cudaGetLastError();
    return sc_fail(nullptr, CSLAM_NNS_ENODEV, "no usable CUDA device (the ScanContext matcher has no CPU fallback)");
  }
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess || prop.major != 10) {
    cudaGetLastError();
    return sc_fail(nullptr, CSLAM_NNS_ENODEV, "device is not sm_100 (this library is built for sm_100a only)");
  }
  cslam_sc_t h = new cslam_sc_s();
  h->device = device;
  h->R = rings;
  h->S = sectors;
  h->C = num_candidates;
  h->initial = initial_capacity > 0 ? initial_capacity : 1000;
  ScDeviceGuard g(device);
  int rc = sc_grow(h, h->initial, 0);
  if (rc != 0) {
    g_sc_create_error = h->err;
    delete h;
    return rc;
  }
  *out = h;
  return 0;
}

int cslam_sc_destroy(cslam_sc_t h) {
  if (!h) return 0;
  {
    ScDeviceGuard g(h->device);
    cudaDeviceSynchronize();
    if (h->dev_tmp) cudaFree(h->dev_tmp);
    if (h->pin.p) cudaFreeHost(h->pin.p);
    h->desc.release();
    h->keys.release();
  }
  delete h;
  return 0;
}

int64_t cslam_sc_size(cslam_sc_t h) { return h ? h->n : -1; }
void* cslam_sc_descriptors_ptr(cslam_sc_t h) { return h ? h->desc.base() : nullptr; }
void* cslam_sc_ringkeys_ptr(cslam_sc_t h) { return h ? h->keys.base() : nullptr; }

int cslam_sc_add(cslam_sc_t h, const double* descriptor, const double* ringkey, void* stream_, int64_t* row_out) {
  if (!h || !descriptor || !ringkey) return sc_fail(h, CSLAM_NNS_EINVAL, "NULL argument");
  cudaStream_t stream = (cudaStream_t)stream_;
  ScDeviceGuard g(h->device);
  int rc = sc_grow(h, h->n + 1, stream);
  if (rc) return rc;
  const size_t dbytes = (size_t)h->R * h->S * 8, kbytes = (size_t)h->R * 8;
  // the sources are pageable host memory: cudaMemcpyAsync stages them before returning
  SCK(h, cudaMemcpyAsync((char*)h->desc.base() + (size_t)h->n * dbytes, descriptor, dbytes, cudaMemcpyHostToDevice, stream));
  SCK(h, cudaMemcpyAsync((char*)h->keys.base() + (size_t)h->n * kbytes, ringkey, kbytes, cudaMemcpyHostToDevice, stream));
  if (row_out) *row_out = h->n;
  h->n += 1;
  return 0;
}

int cslam_sc_search(cslam_sc_t h, const double* query_descriptor, const double* query_ringkey,
                    int64_t* cand_row, double* cand_dist, int32_t* cand_yaw, void* stream_) {
  if (!h || !query_descriptor || !query_ringkey || !cand_row || !cand_dist || !cand_yaw)
    return sc_fail(h, CSLAM_NNS_EINVAL, "NULL argument");
  cudaStream_t stream = (cudaStream_t)stream_;
  ScDeviceGuard g(h->device);
  const int R = h->R, S = h->S, C = h->C;
  if (h->n == 0) {
    for (int i = 0; i < C; ++i) {
      cand_row[i] = -1;
      cand_dist[i] = 1.0;
      cand_yaw[i] = 0;
    }
    return 0;
  }
  long long grid_ll = (h->n + kScThreads - 1) / kScThreads;
  const int grid = (int)(grid_ll > kScMaxGrid ? kScMaxGrid : grid_ll);
  // device scratch: [query R*S][ringkey R][part_d grid*C][out_dist C][out_row C (i64)][part_i grid*C][out_yaw C]
  const size_t q_off = 0, rk_off = q_off + (size_t)R * S * 8, pd_off = rk_off + (size_t)R * 8;
  const size_t od_off = pd_off + (size_t)kScMaxGrid * C * 8, or_off = od_off + (size_t)C * 8;
  const size_t pi_off = or_off + (size_t)C * 8, oy_off = pi_off + (size_t)kScMaxGrid * C * 4;
  const size_t total = oy_off + (size_t)C * 4;
  if (h->dev_tmp_bytes < total) {
    if (h->dev_tmp) cudaFree(h->dev_tmp);
    h->dev_tmp = nullptr;
    h->dev_tmp_bytes = 0;
    SCK(h, cudaMalloc(&h->dev_tmp, total));
    h->dev_tmp_bytes = total;
  }
  const size_t in_bytes = (size_t)(R * S + R) * 8, out_bytes = (size_t)C * 20;
  if (!h->pin.ensure(in_bytes + out_bytes + 64)) return sc_fail(h, CSLAM_NNS_ENOMEM, "pinned staging");
  char* base = (char*)h->dev_tmp;
  memcpy(h->pin.p, query_descriptor, (size_t)R * S * 8);
  memcpy((char*)h->pin.p + (size_t)R * S * 8, query_ringkey, (size_t)R * 8);
  SCK(h, cudaMemcpyAsync(base + q_off, h->pin.p, in_bytes, cudaMemcpyHostToDevice, stream));  // q and rk are adjacent
  const size_t smem1 = (size_t)kScThreads * C * 12 + 8 * sizeof(ScKey) + (size_t)R * 8 + 16;
  static bool attr1[64] = {}, attr2[64] = {};
  if (smem1 > 48 * 1024 && h->device < 64 && !attr1[h->device]) {
    SCK(h, cudaFuncSetAttribute(sc_rk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    attr1[h->device] = true;
  }
  sc_rk_kernel<<<grid, kScThreads, smem1, stream>>>((const double*)h->keys.base(), h->n, R, (const double*)(base + rk_off), C,
                                                    (double*)(base + pd_off), (int*)(base + pi_off));
  count_launch();
  SCK(h, cudaGetLastError());
  const size_t M = (size_t)grid * C;
  const size_t smem2 = (M + C + 2 * (size_t)R * S + 2 * S + (size_t)S * S + S) * 8 + 8 * sizeof(ScKey) + (M + C) * 4 + 2 * S + 64;
  if (smem2 > 48 * 1024 && h->device < 64 && !attr2[h->device]) {
    SCK(h, cudaFuncSetAttribute(sc_dist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    attr2[h->device] = true;
  }
  if (smem2 > 200 * 1024) return sc_fail(h, CSLAM_NNS_EINVAL, "descriptor shape too large for the distance kernel");
  sc_dist_kernel<<<C, kScThreads, smem2, stream>>>((const double*)h->desc.base(), R, S, (const double*)(base + q_off), C, grid,
                                                   (const double*)(base + pd_off), (const int*)(base + pi_off),
                                                   (long long*)(base + or_off), (double*)(base + od_off), (int*)(base + oy_off));
  count_launch();
  SCK(h, cudaGetLastError());
  char* hp = (char*)h->pin.p + in_bytes;
  hp += (8 - ((uintptr_t)hp & 7)) & 7;
  SCK(h, cudaMemcpyAsync(hp, base + od_off, (size_t)C * 16, cudaMemcpyDeviceToHost, stream));           // dist, row
  SCK(h, cudaMemcpyAsync(hp + (size_t)C * 16, base + oy_off, (size_t)C * 4, cudaMemcpyDeviceToHost, stream));
  SCK(h, cudaStreamSynchronize(stream));
  memcpy(cand_dist, hp, (size_t)C * 8);
  memcpy(cand_row, hp + (size_t)C * 8, (size_t)C * 8);
  memcpy(cand_yaw, hp + (size_t)C * 16, (size_t)C * 4);
  return 0;
}

}  // extern "C"
#pragma GCC visibility pop
