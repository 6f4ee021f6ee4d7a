// K4 — merge `parts` sorted key lists per query into one top-k list and decode it.
// Used (a) after the batched kernel K2 (parts = its grid), (b) after the all-gather of the
// row-sharded search (parts = world size) and (c) by the looped single-query path.
// The merge is a thresholded stream into a warp-held sorted list (common.cuh), so the result is the
// k largest 64-bit keys regardless of how the candidates were partitioned: identical output for any
// shard count or grid size.
#include "common.cuh"
#include "kernels.h"

namespace cslam {
namespace {

constexpr int W4 = 4;

// One block per query.  k_out <= 32: every warp keeps its running top-k in registers (one key per lane) and
// folds its share of the parts with loads batched 4 deep; warp 0 folds the 4 warp lists.  `floor_ord`
// (optional, [B]): a proven lower bound of each query's k-th best similarity as an ordered uint — the
// batched kernel's shared threshold — below which a key is dropped before it is offered; with it almost no
// insertion is left.  k_out > 32: sorted lists in shared memory (warp_list_stream).
__global__ void __launch_bounds__(W4 * 32) k4_kernel(const u64* __restrict__ keys, int parts,
                                                     long long B, int k_in, int k_out,
                                                     const unsigned* __restrict__ floor_ord,
                                                     const unsigned* __restrict__ active_mask,
                                                     u64* out_keys, long long* out_idx,
                                                     float* out_sim, long long idx_stride,
                                                     long long idx_offset) {
  // merge behind the masked fallback pass: only the queries that pass served have lists (and results to write)
  if (active_mask != nullptr && !((__ldcg(active_mask + (blockIdx.x >> 5)) >> (blockIdx.x & 31)) & 1u)) return;
  extern __shared__ __align__(16) unsigned char smem4[];
  u64* lists = reinterpret_cast<u64*>(smem4);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long b = blockIdx.x;
  u64* my_list = lists + (size_t)warp * k_out;
  u64 thr = 0ull, mine = 0ull;
  if (k_out <= 32) {
    const u64 floor = floor_ord ? ((u64)floor_ord[b] << 32) : 0ull;
    // this warp's parts are warp, warp + 4, ...: element e of the warp's virtual key stream
    const int my_parts = parts > warp ? (parts - warp + W4 - 1) / W4 : 0;
    const int total = my_parts * k_in;
    for (int base = 0; base < total; base += 128) {
      u64 kk[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int e = base + u * 32 + lane;
        u64 v = 0ull;
        if (e < total) {
          const int pi = e / k_in, j = e - pi * k_in;
          v = keys[((size_t)(warp + pi * W4) * B + b) * k_in + j];
        }
        kk[u] = v;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) offer(my_list, k_out, kk[u], kk[u] != 0ull, thr, mine, lane, floor);
    }
    if (k_out == 1) mine = thr;  // offer() keeps the k == 1 winner in `thr`
    if (lane < k_out) my_list[lane] = mine;
    __syncthreads();
    if (warp != 0) return;
    for (int base = k_out; base < W4 * k_out; base += 32) {
      const int i = base + lane;
      const u64 key = i < W4 * k_out ? lists[i] : 0ull;
      offer(my_list, k_out, key, key != 0ull, thr, mine, lane);
    }
    if (k_out == 1) mine = thr;  // offer() keeps the k == 1 winner in `thr`
    if (lane < k_out) my_list[lane] = mine;
    __syncwarp();
  } else {
    warp_list_clear(my_list, k_out, lane);
    for (int p = warp; p < parts; p += W4) {
      const u64* src = keys + ((size_t)p * B + b) * k_in;
      warp_list_stream(my_list, k_out, thr, src, k_in, lane);
    }
    __syncthreads();
    if (warp != 0) return;
    for (int w = 1; w < W4; ++w)
      warp_list_stream(my_list, k_out, thr, lists + (size_t)w * k_out, k_out, lane);
  }
  for (int i = lane; i < k_out; i += 32) {
    const u64 key = my_list[i];
    const bool have = key != 0ull;
    const long long grow = (long long)key_row(key) * idx_stride + idx_offset;
    const size_t o = (size_t)b * k_out + i;
    if (out_keys)
      out_keys[o] = have ? ((key & 0xffffffff00000000ull) | (u64)(~(uint32_t)grow)) : 0ull;
    if (out_idx) out_idx[o] = have ? grow : -1ll;
    if (out_sim) out_sim[o] = have ? key_sim(key) : -INFINITY;
  }
}

}  // namespace

cudaError_t k4_merge(const u64* keys, int parts, long long B, int k_in, int k_out, u64* out_keys,
                     long long* out_idx, float* out_sim, long long idx_stride,
                     long long idx_offset, cudaStream_t stream, const unsigned* floor_ord,
                     const unsigned* active_mask) {
  if (B <= 0 || k_out <= 0) return cudaSuccess;
  const size_t smem = (size_t)W4 * k_out * 8;
  k4_kernel<<<(unsigned)B, W4 * 32, smem, stream>>>(keys, parts, B, k_in, k_out, floor_ord, active_mask, out_keys,
                                                    out_idx, out_sim, idx_stride, idx_offset);
  count_launch();
  return cudaGetLastError();
}

}  // namespace cslam
