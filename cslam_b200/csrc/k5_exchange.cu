// K5 — the exchange step of the row-sharded search as ONE kernel over NVLink peer memory.
//
// After the local search every rank holds, per query, its k best packed keys (global row in the low word).
// The final answer is the k largest keys over all ranks.  Instead of "all-gather (NCCL) -> merge kernel",
// block b of this kernel
//   1. stores query b's k local keys straight into EVERY rank's exchange buffer (peer stores; the buffers
//      are opened once with CUDA IPC), then publishes them with a release store of the call's epoch into
//      that rank's flag word for (sender, query);
//   2. waits (acquire loads on ITS OWN flags) until all `world` senders have published query b;
//   3. folds the world x k keys with the warp register top-k of common.cuh, decodes and writes (row, sim).
// Every block pushes before it waits and a block only ever waits for the SAME block index of the other
// ranks, so there is no ordering between blocks that could deadlock; all B <= 4096 one-warp blocks are
// co-resident anyway.  Buffers are double-buffered by epoch parity: a rank cannot start call e+2 before
// every peer has finished reading call e (it needs their flags of call e+1 first).
// A bounded spin (globaltimer) turns a dead peer into an error word in mapped host memory, not a hung GPU.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "../../include/cslam_nns.h"
#include "common.cuh"
#include "kernels.h"

namespace cslam {
namespace {

constexpr int kMaxWorld = 8;
constexpr long long kXchgCap = 131072;  // keys per (parity, sender): B * k of one call
constexpr int kXchgMaxB = 4096;

struct XchgDev {
  u64* keys[kMaxWorld];        // rank p's key buffer  [2][world][kXchgCap], in THIS process' address space
  unsigned* flags[kMaxWorld];  // rank p's flag buffer [2][world][kXchgMaxB]
  int world, rank;
  unsigned* err;  // mapped host word: set to the epoch of a call whose wait timed out
};

__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ u64 globaltimer_ns() {
  u64 t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// phase bit 1: push, bit 2: wait + merge (3 = the production kernel; the single-GPU emulation in the tests
// runs the two phases as separate launches so that no launch ever waits for another)
__global__ void __launch_bounds__(32) k5_kernel(const XchgDev x, const u64* __restrict__ local, int k,
                                                unsigned epoch, int phase, long long* out_idx,
                                                float* out_sim, u64 timeout_ns) {
  const int lane = threadIdx.x;
  const long long b = blockIdx.x;
  const unsigned parity = epoch & 1u;
  if (phase & 1) {
    if (lane < k) {
      const u64 key = local[b * k + lane];
      const size_t slot = ((size_t)(parity * x.world + x.rank)) * kXchgCap + (size_t)b * k + lane;
      for (int p = 0; p < x.world; ++p) x.keys[p][slot] = key;
    }
    // the warp barrier orders the other lanes' key stores before this lane's release store (release is
    // cumulative), so no separate system-scope fence is needed
    __syncwarp();
    if (lane < x.world)
      st_release_sys(x.flags[lane] + ((size_t)(parity * x.world + x.rank)) * kXchgMaxB + b, epoch);
  }
  if (!(phase & 2)) return;
  bool ok = true;
  if (lane < x.world) {
    const unsigned* f = x.flags[x.rank] + ((size_t)(parity * x.world + lane)) * kXchgMaxB + b;
    const u64 t0 = globaltimer_ns();
    while (ld_acquire_sys(f) != epoch) {
      if (globaltimer_ns() - t0 > timeout_ns) {
        ok = false;
        break;
      }
      __nanosleep(64);
    }
  }
  if (!__all_sync(kFull, ok)) {
    // a peer never published this query: report the call (mapped host word, read by cslam_nns_xchg_status
    // right after the caller's synchronisation) and return "no result" for the query instead of a merge of
    // whatever the buffers hold
    if (lane == 0) *x.err = epoch;
    if (lane < k) {
      if (out_idx) out_idx[(size_t)b * k + lane] = -1ll;
      if (out_sim) out_sim[(size_t)b * k + lane] = -INFINITY;
    }
    return;
  }
  u64 thr = 0ull, mine = 0ull;
  const u64* mykeys = x.keys[x.rank] + (size_t)parity * x.world * kXchgCap + (size_t)b * k;
  for (int r = 0; r < x.world; ++r) {
    const u64 key = lane < k ? __ldcv(mykeys + (size_t)r * kXchgCap + lane) : 0ull;
    offer(nullptr, k, key, key != 0ull, thr, mine, lane);
  }
  if (k == 1) mine = thr;  // offer() keeps the k == 1 winner in `thr`
  if (lane < k) {
    const bool have = mine != 0ull;
    const size_t o = (size_t)b * k + lane;
    if (out_idx) out_idx[o] = have ? (long long)key_row(mine) : -1ll;
    if (out_sim) out_sim[o] = have ? key_sim(mine) : -INFINITY;
  }
}

}  // namespace
}  // namespace cslam

using namespace cslam;

struct cslam_nns_xchg_s {
  int device = 0, world = 0, rank = 0;
  u64* keys = nullptr;        // this rank's buffers (cudaMalloc: IPC-exportable)
  unsigned* flags = nullptr;
  unsigned* err_host = nullptr;  // mapped pinned
  unsigned* err_dev = nullptr;
  void* opened[2 * kMaxWorld] = {};  // IPC mappings to close
  XchgDev dev{};
  bool connected = false;
  unsigned epoch = 0;
  double timeout_s = 20.0;
  std::string error;
};

namespace {
std::string g_xchg_error;

int xfail(cslam_nns_xchg_t x, int code, const std::string& msg) {
  (x ? x->error : g_xchg_error) = msg;
  return code;
}
int xcuda(cslam_nns_xchg_t x, const char* what, cudaError_t e) {
  return xfail(x, CSLAM_NNS_ECUDA, std::string(what) + ": " + cudaGetErrorString(e));
}
struct DevGuard {
  int prev = -1;
  explicit DevGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    if (prev != dev) cudaSetDevice(dev);
    else prev = -1;
  }
  ~DevGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};
size_t keys_bytes(int world) { return (size_t)2 * world * kXchgCap * sizeof(u64); }
size_t flags_bytes(int world) { return (size_t)2 * world * kXchgMaxB * sizeof(unsigned); }
}  // namespace

#pragma GCC visibility push(default)
extern "C" {

const char* cslam_nns_xchg_last_error(cslam_nns_xchg_t x) { return (x ? x->error : g_xchg_error).c_str(); }

int cslam_nns_xchg_create(int device, int world, int rank, cslam_nns_xchg_t* out) {
  if (!out || world < 1 || world > kMaxWorld || rank < 0 || rank >= world)
    return xfail(nullptr, CSLAM_NNS_EINVAL, "bad arguments to cslam_nns_xchg_create");
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || device < 0 || device >= count)
    return xfail(nullptr, CSLAM_NNS_ENODEV, "no such CUDA device");
  DevGuard g(device);
  cslam_nns_xchg_t x = new cslam_nns_xchg_s();
  x->device = device;
  x->world = world;
  x->rank = rank;
  cudaError_t e = cudaMalloc(&x->keys, keys_bytes(world));
  if (e == cudaSuccess) e = cudaMalloc(&x->flags, flags_bytes(world));
  if (e == cudaSuccess) e = cudaMemset(x->keys, 0, keys_bytes(world));
  if (e == cudaSuccess) e = cudaMemset(x->flags, 0, flags_bytes(world));
  if (e == cudaSuccess) e = cudaHostAlloc(&x->err_host, sizeof(unsigned), cudaHostAllocMapped);
  if (e == cudaSuccess) {
    *x->err_host = 0u;
    e = cudaHostGetDevicePointer(&x->err_dev, x->err_host, 0);
  }
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    int rc = xcuda(nullptr, "exchange buffers", e);
    cslam_nns_xchg_destroy(x);
    return rc;
  }
  const char* t = getenv("CSLAM_NNS_XCHG_TIMEOUT_S");
  if (t && atof(t) > 0) x->timeout_s = atof(t);
  *out = x;
  return 0;
}

int cslam_nns_xchg_destroy(cslam_nns_xchg_t x) {
  if (!x) return 0;
  DevGuard g(x->device);
  cudaDeviceSynchronize();
  for (void* p : x->opened)
    if (p) cudaIpcCloseMemHandle(p);
  if (x->keys) cudaFree(x->keys);
  if (x->flags) cudaFree(x->flags);
  if (x->err_host) cudaFreeHost(x->err_host);
  delete x;
  return 0;
}

int cslam_nns_xchg_export(cslam_nns_xchg_t x, void* handle128) {
  if (!x || !handle128) return xfail(x, CSLAM_NNS_EINVAL, "bad arguments to cslam_nns_xchg_export");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  DevGuard g(x->device);
  cudaIpcMemHandle_t hk, hf;
  cudaError_t e = cudaIpcGetMemHandle(&hk, x->keys);
  if (e == cudaSuccess) e = cudaIpcGetMemHandle(&hf, x->flags);
  if (e != cudaSuccess) return xcuda(x, "cudaIpcGetMemHandle", e);
  memcpy(handle128, &hk, 64);
  memcpy((char*)handle128 + 64, &hf, 64);
  return 0;
}

static int finish_connect(cslam_nns_xchg_t x) {
  x->dev.world = x->world;
  x->dev.rank = x->rank;
  x->dev.err = x->err_dev;
  x->connected = true;
  return 0;
}

int cslam_nns_xchg_connect(cslam_nns_xchg_t x, const void* handles) {
  if (!x || !handles) return xfail(x, CSLAM_NNS_EINVAL, "bad arguments to cslam_nns_xchg_connect");
  if (x->connected) return xfail(x, CSLAM_NNS_ESTATE, "exchange already connected");
  DevGuard g(x->device);
  for (int p = 0; p < x->world; ++p) {
    if (p == x->rank) {
      x->dev.keys[p] = x->keys;
      x->dev.flags[p] = x->flags;
      continue;
    }
    cudaIpcMemHandle_t hk, hf;
    memcpy(&hk, (const char*)handles + (size_t)p * 128, 64);
    memcpy(&hf, (const char*)handles + (size_t)p * 128 + 64, 64);
    void *pk = nullptr, *pf = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&pk, hk, cudaIpcMemLazyEnablePeerAccess);
    if (e == cudaSuccess) {
      x->opened[2 * p] = pk;
      e = cudaIpcOpenMemHandle(&pf, hf, cudaIpcMemLazyEnablePeerAccess);
    }
    if (e != cudaSuccess) return xcuda(x, "cudaIpcOpenMemHandle", e);
    x->opened[2 * p + 1] = pf;
    x->dev.keys[p] = (u64*)pk;
    x->dev.flags[p] = (unsigned*)pf;
  }
  return finish_connect(x);
}

int cslam_nns_xchg_connect_local(cslam_nns_xchg_t x, const cslam_nns_xchg_t* all) {
  if (!x || !all) return xfail(x, CSLAM_NNS_EINVAL, "bad arguments to cslam_nns_xchg_connect_local");
  if (x->connected) return xfail(x, CSLAM_NNS_ESTATE, "exchange already connected");
  for (int p = 0; p < x->world; ++p) {
    if (!all[p] || all[p]->world != x->world || all[p]->rank != p)
      return xfail(x, CSLAM_NNS_EINVAL, "exchange objects do not form one group");
    x->dev.keys[p] = all[p]->keys;
    x->dev.flags[p] = all[p]->flags;
  }
  return finish_connect(x);
}

int cslam_nns_xchg_merge(cslam_nns_xchg_t x, const uint64_t* local_keys_dev, int64_t B, int k, int phase,
                         int64_t* out_idx_dev, float* out_sim_dev, void* stream_) {
  if (!x) return CSLAM_NNS_EINVAL;
  if (!x->connected) return xfail(x, CSLAM_NNS_ESTATE, "exchange not connected");
  if (!local_keys_dev || B <= 0 || k <= 0 || phase < 1 || phase > 3)
    return xfail(x, CSLAM_NNS_EINVAL, "bad arguments to cslam_nns_xchg_merge");
  if (k > 32 || B > kXchgMaxB || B * k > kXchgCap)
    return xfail(x, CSLAM_NNS_EINVAL, "exchange holds k <= 32, B <= 4096, B * k <= 131072 per call");
  if (*x->err_host != 0u) {
    char msg[128];
    snprintf(msg, sizeof msg, "a peer did not publish its keys within %.0f s (call %u)", x->timeout_s,
             *x->err_host);
    return xfail(x, CSLAM_NNS_ECUDA, msg);
  }
  DevGuard g(x->device);
  if (phase & 1) ++x->epoch;  // a call is identified by the epoch of its push
  k5_kernel<<<(unsigned)B, 32, 0, (cudaStream_t)stream_>>>(x->dev, (const u64*)local_keys_dev, k, x->epoch,
                                                            phase, (long long*)out_idx_dev, out_sim_dev,
                                                            (u64)(x->timeout_s * 1e9));
  count_launch();
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return xcuda(x, "k5_kernel", e);
  return 0;
}

int cslam_nns_xchg_status(cslam_nns_xchg_t x) {
  if (!x) return CSLAM_NNS_EINVAL;
  return *x->err_host != 0u ? xfail(x, CSLAM_NNS_ECUDA, "a peer did not publish its keys in time") : 0;
}

}  // extern "C"
#pragma GCC visibility pop
