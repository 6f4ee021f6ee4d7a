// cslam_nns_search_static (include/cslam_nns.h): the single-query search for callers that may not allocate —
// code under CUDA graph capture, serving loops and harnesses that forbid cudaMalloc.  It has NO counterpart in the
// reference.
//
// The kernel is the single-query kernel K1 (k1_single.cu) unchanged; what differs is where its scratch lives.  K1
// needs, per launch, one 16-byte ticket block (arrival counter + the k == 1 path's best key; zero before the
// launch) and grid x k packed keys of per-CTA survivors.  cslam_nns_search takes both from buffers the handle
// allocates; here they are one file-scope __device__ array of a fixed, modest size, so this entry point calls no
// cudaMalloc / cudaFree / allocating library routine at all.  Consequences, stated in the header as well:
//   * limits: k <= 32, at most 512 CTAs (any GPU with <= 512 SMs), rows < 2^32 - 1;
//   * every GPU that loads the library holds the ~129 KB array until the process exits;
//   * EVERY call on a device shares the array: two calls whose kernels overlap on the GPU (different streams,
//     or programmatic dependent launch) race on it.  Calls on ONE stream are ordered and safe.  The array keeps
//     the previous call's contents; the ticket block is cleared on the stream before every launch (the kernel
//     also leaves it at zero, but a launch that was aborted would not), the survivor keys are fully rewritten by
//     the CTAs of the launch before its last CTA reads them.
#include <stdio.h>
#include <string.h>

#include <mutex>

#include "../../include/cslam_nns.h"
#include "kernels.h"

namespace cslam {
namespace {

constexpr int kStaticMaxK = CSLAM_NNS_STATIC_MAX_K;        // 32
constexpr int kStaticMaxCtas = CSLAM_NNS_STATIC_MAX_CTAS;  // 512
constexpr int kTicketWords = 32;                           // 256 bytes: ticket block, padded to the alignment
constexpr size_t kScratchWords = kTicketWords + (size_t)kStaticMaxCtas * kStaticMaxK;   // u64 words

// The scratch itself.  In HOST code this name is only a stand-in for the symbol: the device address comes from
// cudaGetSymbolAddress (device_scratch below), once per device, with that device current.
static __device__ __align__(256) unsigned long long g_static_scratch[kScratchWords];

__global__ void fill_tail_kernel(long long* idx, float* sim, int from, int k) {
  const int i = from + (int)threadIdx.x;
  if (i < k) {
    if (idx) idx[i] = -1ll;
    if (sim) sim[i] = -INFINITY;
  }
}

struct PerDevice {
  unsigned long long* scratch = nullptr;   // device address of g_static_scratch on this device
  int num_sms = 0;
  bool ready = false;
};
PerDevice g_dev[64];
std::mutex g_dev_mu;

int set_err(char* err, int64_t err_len, int code, const char* msg) {
  if (err && err_len > 0) snprintf(err, (size_t)err_len, "cslam_nns_search_static: %s", msg);
  return code;
}

// Set-up, once per device, with `device` current.
cudaError_t device_scratch(int device, PerDevice** out, const char** what) {
  std::lock_guard<std::mutex> lock(g_dev_mu);
  PerDevice& d = g_dev[device];
  if (!d.ready) {
    int major = 0, sms = 0;
    cudaError_t e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    if (e != cudaSuccess) { *what = "cudaDeviceGetAttribute"; return e; }
    if (major != 10) { *what = "not an sm_100 device"; return cudaErrorInvalidDevice; }
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess || !p) { *what = "cudaGetSymbolAddress(g_static_scratch)"; return e != cudaSuccess ? e : cudaErrorInvalidSymbol; }
    d.scratch = static_cast<unsigned long long*>(p);
    d.num_sms = sms;
    d.ready = true;
  }
  *out = &d;
  return cudaSuccess;
}

}  // namespace
}  // namespace cslam

using namespace cslam;

#pragma GCC visibility push(default)
extern "C" int cslam_nns_search_static(int device, const float* db_dev, int64_t n, int64_t ld, int64_t dim,
                                        const float* query_dev, int k, int64_t* out_idx_dev, float* out_sim_dev,
                                        void* stream, char* err, int64_t err_len) {
  if (err && err_len > 0) err[0] = 0;
  char msg[256];
  if (device < 0 || device >= 64) return set_err(err, err_len, CSLAM_NNS_EINVAL, "device must be in [0, 64)");
  if (!db_dev || !query_dev || (!out_idx_dev && !out_sim_dev))
    return set_err(err, err_len, CSLAM_NNS_EINVAL, "db_dev, query_dev and one of the outputs must be non-NULL");
  if (k < 1 || k > kStaticMaxK) {
    snprintf(msg, sizeof msg, "k = %d is outside [1, %d] (CSLAM_NNS_STATIC_MAX_K: the static scratch holds %d keys "
             "per CTA); use cslam_nns_search for larger k", k, kStaticMaxK, kStaticMaxK);
    return set_err(err, err_len, CSLAM_NNS_EINVAL, msg);
  }
  if (dim < 1 || dim > (1 << 20)) return set_err(err, err_len, CSLAM_NNS_EINVAL, "dim must be in [1, 2^20]");
  if (ld == 0) ld = dim;
  if (ld < dim || (ld & 3) || ld > (1 << 20))
    return set_err(err, err_len, CSLAM_NNS_EINVAL, "row pitch ld must be >= dim, a multiple of 4 floats and <= 2^20");
  if ((reinterpret_cast<uintptr_t>(db_dev) & 15) || (reinterpret_cast<uintptr_t>(query_dev) & 15))
    return set_err(err, err_len, CSLAM_NNS_EINVAL, "db_dev and query_dev must be 16-byte aligned");
  if (n < 1 || n >= 0xffffffffll) {
    snprintf(msg, sizeof msg, "n = %lld is outside [1, 2^32 - 2] (rows are packed into 32 bits of the key)",
             (long long)n);
    return set_err(err, err_len, CSLAM_NNS_EINVAL, msg);
  }

  int prev = -1;
  cudaError_t e = cudaGetDevice(&prev);
  if (e == cudaSuccess && prev != device) e = cudaSetDevice(device);
  if (e != cudaSuccess) {
    snprintf(msg, sizeof msg, "cudaSetDevice(%d): %s", device, cudaGetErrorString(e));
    cudaGetLastError();
    return set_err(err, err_len, CSLAM_NNS_ENODEV, msg);
  }
  int rc = 0;
  do {
    PerDevice* d = nullptr;
    const char* what = "";
    e = device_scratch(device, &d, &what);
    if (e != cudaSuccess) {
      snprintf(msg, sizeof msg, "%s: %s", what, cudaGetErrorString(e));
      rc = set_err(err, err_len, e == cudaErrorInvalidDevice ? CSLAM_NNS_ENODEV : CSLAM_NNS_ECUDA, msg);
      break;
    }
    if (d->num_sms > kStaticMaxCtas) {
      snprintf(msg, sizeof msg, "device has %d SMs, the static scratch is sized for %d CTAs "
               "(CSLAM_NNS_STATIC_MAX_CTAS)", d->num_sms, kStaticMaxCtas);
      rc = set_err(err, err_len, CSLAM_NNS_EINVAL, msg);
      break;
    }
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int k_run = (int)(n < k ? n : k);
    // what must start at zero: the ticket block (arrival counter, k == 1 best key)
    e = cudaMemsetAsync(d->scratch, 0, kTicketWords * sizeof(unsigned long long), s);
    if (e != cudaSuccess) {
      snprintf(msg, sizeof msg, "cudaMemsetAsync(ticket): %s", cudaGetErrorString(e));
      rc = set_err(err, err_len, CSLAM_NNS_ECUDA, msg);
      break;
    }
    K1Args a;
    a.db = db_dev;
    a.n = n;
    a.ld = (int)ld;
    a.dim = (int)dim;
    a.q = query_dev;
    a.k = k_run;
    a.scratch = d->scratch + kTicketWords;
    a.ticket = reinterpret_cast<unsigned*>(d->scratch);
    a.out_keys = nullptr;
    a.out_idx = reinterpret_cast<long long*>(out_idx_dev);
    a.out_sim = out_sim_dev;
    a.idx_stride = 1;
    a.idx_offset = 0;
    a.all_sims = nullptr;
    K1Tuning t;
    t.stage_bytes = 8192;   // the defaults of cslam_nns_create (abi.cu)
    t.stages = 2;
    t.ctas = d->num_sms;    // <= kStaticMaxCtas: k1_launch never uses more CTAs than this
    e = k1_launch(a, t, d->num_sms, s);
    if (e == cudaSuccess && k_run < k) {
      fill_tail_kernel<<<1, kStaticMaxK, 0, s>>>(a.out_idx, a.out_sim, k_run, k);
      count_launch();
      e = cudaGetLastError();
    }
    if (e != cudaSuccess) {
      snprintf(msg, sizeof msg, "kernel launch: %s", cudaGetErrorString(e));
      cudaGetLastError();
      rc = set_err(err, err_len, CSLAM_NNS_ECUDA, msg);
    }
  } while (0);
  if (prev >= 0 && prev != device) cudaSetDevice(prev);
  return rc;
}
#pragma GCC visibility pop
