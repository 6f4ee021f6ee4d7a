// TMA throughput probe (development tool, not part of the library).
// One thread per CTA keeps S tensor-tile loads in flight and re-issues a slot as soon as it lands; nothing
// reads the data.  Reports bytes/clk/SM and TB/s for a given box shape, source pattern and ring depth.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/tma_probe tools/tma_probe.cu -lcuda
//   tools/tma_probe <bk floats> <box rows> <panels> <stages> <hot 0|1> <ctas> <ld floats> [iters]
// hot=1: every CTA re-reads the same 2048-row region (L2 resident); hot=0: CTAs stream disjoint row tiles.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(256, 1)
probe(const __grid_constant__ CUtensorMap pmap, const CUtensorMap* gmap, int bk, int rows, int panels, int S, int hot, long long total_rows,
      int kblocks, int iters, long long* cycles) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* smem = smem_raw + ((1024u - (s32(smem_raw) & 1023u)) & 1023u);
  __shared__ uint64_t bars_all[8][16];
  const uint32_t box_bytes = (uint32_t)bk * 4u * rows * panels;
  const int w = threadIdx.x >> 5, W = blockDim.x >> 5;
  uint64_t* bars = bars_all[w];
  smem += (size_t)w * S * box_bytes;
  const CUtensorMap* mapp = gmap ? gmap : &pmap;
  if ((threadIdx.x & 31) == 0) {
    for (int s = 0; s < S; ++s)
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(&bars[s])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if ((threadIdx.x & 31) != 0) return;
  const long long tiles = total_rows / rows;
  long long tile = (long long)blockIdx.x * W + w;
  int kb = 0;
  auto issue = [&](int s) {
    // tiles and 2048/rows are powers of two: masks, no 64-bit division in the issuing thread
    const long long row0 = hot ? (long long)((tile * 7 + blockIdx.x) & (2048 / rows - 1)) * rows : (tile & (tiles - 1)) * rows;
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(&bars[s])), "r"(box_bytes) : "memory");
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(s32(smem + (size_t)s * box_bytes)), "l"(mapp), "r"(s32(&bars[s])), "r"(0), "r"((int)row0), "r"(kb * panels)
        : "memory");
    kb += 1;
    if (kb * panels >= kblocks) { kb = 0; tile += (long long)gridDim.x * W; }
  };
  for (int s = 0; s < S; ++s) issue(s);
  const long long t0 = clock64();
  int s = 0;
  uint32_t ph = 0;
  for (int i = 0; i < iters; ++i) {
    uint32_t ok;
    do {
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                   : "=r"(ok) : "r"(s32(&bars[s])), "r"(ph) : "memory");
    } while (!ok);
    issue(s);
    if (++s == S) { s = 0; ph ^= 1; }
  }
  const long long t1 = clock64();
  for (int j = 0; j < S; ++j) {  // drain
    uint32_t ok;
    do {
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                   : "=r"(ok) : "r"(s32(&bars[s])), "r"(ph) : "memory");
    } while (!ok);
    if (++s == S) { s = 0; ph ^= 1; }
  }
  if (w == 0) cycles[blockIdx.x] = t1 - t0;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main(int argc, char** argv) {
  if (argc < 8) { printf("usage: bk rows panels stages hot ctas ld [iters]\n"); return 1; }
  const int bk = atoi(argv[1]), rows = atoi(argv[2]), panels = atoi(argv[3]), S = atoi(argv[4]);
  const int hot = atoi(argv[5]), ctas = atoi(argv[6]), ld = atoi(argv[7]);
  const int iters = argc > 8 ? atoi(argv[8]) : 2000;
  const int W = argc > 9 ? atoi(argv[9]) : 1;         // issuing warps per CTA (each with its own ring)
  const int gdesc = argc > 10 ? atoi(argv[10]) : 0;   // 1: tensor map read from global memory
  const int dim = 4096 < ld ? 4096 : ld;
  const long long total_rows = hot ? 2048 : 262144;  // cold: 4.3 GB at ld 4096
  float* db;
  CK(cudaMalloc(&db, (size_t)total_rows * ld * 4));
  CK(cudaMemset(db, 0, (size_t)total_rows * ld * 4));
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qr;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr));
  CUtensorMap map;
  const cuuint64_t gdim[3] = {(cuuint64_t)bk, (cuuint64_t)total_rows, (cuuint64_t)(dim / bk)};
  const cuuint64_t gstr[2] = {(cuuint64_t)ld * 4, (cuuint64_t)bk * 4};
  const cuuint32_t box[3] = {(cuuint32_t)bk, (cuuint32_t)rows, (cuuint32_t)panels};
  const cuuint32_t estr[3] = {1, 1, 1};
  const CUtensorMapSwizzle sw = bk == 32 ? CU_TENSOR_MAP_SWIZZLE_128B : bk == 16 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B;
  CUresult r = ((EncodeTiledFn)fn)(&map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, db, gdim, gstr, box, estr,
                                   CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); return 1; }
  const size_t smem = (size_t)W * S * bk * 4 * rows * panels + 1024;
  if (smem > 227 * 1024) { printf("smem %zu too large\n", smem); return 1; }
  CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  long long* cyc;
  CK(cudaMalloc(&cyc, ctas * 8));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  CUtensorMap* gmap = nullptr;
  if (gdesc) {
    CK(cudaMalloc(&gmap, sizeof(CUtensorMap)));
    CK(cudaMemcpy(gmap, &map, sizeof(CUtensorMap), cudaMemcpyHostToDevice));
  }
  probe<<<ctas, 32 * W, smem>>>(map, gmap, bk, rows, panels, S, hot, total_rows, dim / bk, 200, cyc);  // warm
  CK(cudaDeviceSynchronize());
  cudaEventRecord(e0);
  probe<<<ctas, 32 * W, smem>>>(map, gmap, bk, rows, panels, S, hot, total_rows, dim / bk, iters, cyc);
  cudaEventRecord(e1);
  CK(cudaDeviceSynchronize());
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  long long* h = (long long*)malloc(ctas * 8);
  CK(cudaMemcpy(h, cyc, ctas * 8, cudaMemcpyDeviceToHost));
  double avg = 0;
  for (int i = 0; i < ctas; ++i) avg += (double)h[i];
  avg /= ctas;
  const double box_bytes = (double)bk * 4 * rows * panels;
  printf("bk=%d rows=%d panels=%d S=%d hot=%d ctas=%d ld=%d W=%d gdesc=%d: %.1f cyc/box/warp  %.1f B/clk/SM  %.2f TB/s (event)\n",
         bk, rows, panels, S, hot, ctas, ld, W, gdesc, avg / iters, W * box_bytes * iters / avg,
         W * box_bytes * (iters + S) * ctas / (ms * 1e-3) / 1e12);
  return 0;
}
