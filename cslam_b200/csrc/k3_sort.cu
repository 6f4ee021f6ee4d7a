// Two-pass select for k larger than the fused kernels hold (k > 128, e.g. "rank the whole DB"):
// K1 writes every row's similarity, then a device radix sort orders (orderable(sim), row) pairs.
// This is the same ordering torch.argsort(descending=True) gives at cslam/nns_matching.py:62, made
// deterministic: the sort is stable and rows enter in ascending order, so ties keep the lower row.
// GPU only — there is no host fallback.
#include <cub/device/device_radix_sort.cuh>

#include "common.cuh"
#include "kernels.h"

namespace cslam {
namespace {

__global__ void sort_prep_kernel(const float* __restrict__ sims, long long n, uint32_t* keys,
                                 uint32_t* idx) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long step = (long long)gridDim.x * blockDim.x;
  for (; i < n; i += step) {
    keys[i] = sim_to_ord(sims[i]);
    idx[i] = (uint32_t)i;
  }
}

__global__ void sort_emit_kernel(const uint32_t* __restrict__ keys, const uint32_t* __restrict__ idx,
                                 int k, u64* out_keys, long long* out_idx, float* out_sim,
                                 long long idx_stride, long long idx_offset) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= k) return;
  const long long grow = (long long)idx[i] * idx_stride + idx_offset;
  if (out_keys) out_keys[i] = ((u64)keys[i] << 32) | (u64)(~(uint32_t)grow);
  if (out_idx) out_idx[i] = grow;
  if (out_sim) out_sim[i] = ord_to_sim(keys[i]);
}

}  // namespace

size_t sort_temp_bytes(long long n) {
  size_t bytes = 0;
  cub::DeviceRadixSort::SortPairsDescending(nullptr, bytes, (const uint32_t*)nullptr,
                                            (uint32_t*)nullptr, (const uint32_t*)nullptr,
                                            (uint32_t*)nullptr, (int)n);
  return bytes;
}

cudaError_t sort_select(const float* all_sims, long long n, int k, void* temp, size_t temp_bytes,
                        uint32_t* keys_a, uint32_t* keys_b, uint32_t* idx_a, uint32_t* idx_b,
                        u64* out_keys, long long* out_idx, float* out_sim, long long idx_stride,
                        long long idx_offset, cudaStream_t stream) {
  if (n <= 0 || k <= 0) return cudaSuccess;
  if (n > 0x7fffffffll) return cudaErrorInvalidValue;
  int blocks = (int)((n + 255) / 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  sort_prep_kernel<<<blocks, 256, 0, stream>>>(all_sims, n, keys_a, idx_a);
  count_launch();
  cudaError_t e = cub::DeviceRadixSort::SortPairsDescending(temp, temp_bytes, keys_a, keys_b,
                                                            idx_a, idx_b, (int)n, 0, 32, stream);
  if (e != cudaSuccess) return e;
  count_launch(3);
  if (k > n) k = (int)n;
  sort_emit_kernel<<<(k + 127) / 128, 128, 0, stream>>>(keys_b, idx_b, k, out_keys, out_idx,
                                                       out_sim, idx_stride, idx_offset);
  count_launch();
  return cudaGetLastError();
}

}  // namespace cslam
