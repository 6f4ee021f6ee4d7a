// Internal launch interface between the C-ABI (abi.cu) and the kernels.  Not installed.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace cslam {

typedef unsigned long long u64;

// ---- K1: fused single-query cosine GEMV + warp top-k (k1_single.cu) ---------------------------
static constexpr int kK1Warps = 8;
static constexpr int kK1MaxK = 128;  // largest k the fused select handles

struct K1Tuning {
  int stage_bytes;  // bytes per bulk copy (per warp, per stage)
  int stages;       // ring depth per warp
  int ctas;         // grid size cap (0 = number of SMs)
};

struct K1Args {
  const float* db;
  long long n;
  int ld;  // row pitch in floats, multiple of 4
  int dim;
  const float* q;  // device, dim floats
  int k;
  u64* scratch;      // [grid][k]
  unsigned* ticket;  // zero before launch; reset by the kernel
  u64* out_keys;     // [k] or null
  long long* out_idx;
  float* out_sim;
  long long idx_stride, idx_offset;
  float* all_sims;  // non-null: write every row's similarity, skip the select
  // true: launch with programmatic stream serialisation, i.e. this search may start on SMs the PREVIOUS K1
  // launch in the stream has vacated.  Only valid when the previous launch is a K1 search that shares no
  // scratch, ticket or output with this one (another database, cslam_nns_search_multi).
  bool overlap_prev = false;
  // with overlap_prev: the previous launch searched the SAME database (same handle): the kernel then waits for
  // it (griddepcontrol.wait) after streaming its share of the rows and before it touches its ticket / scratch
  // slot, the output or the done flag.  false = independent databases (cslam_nns_search_multi): no wait at all.
  bool dep_wait = false;
  // host destination: mapped pinned word that receives done_value once the results are written (k <= 32 only)
  unsigned* done_flag = nullptr;
  unsigned done_value = 0;
};
static constexpr int kK1Slots = 4;  // ring of ticket / scratch sets used by chained searches of one database

// Maximum grid K1 will ever use on `device` (sizes the scratch buffer).
int k1_max_grid(int num_sms);
cudaError_t k1_launch(const K1Args& a, const K1Tuning& t, int num_sms, cudaStream_t stream);

// ---- two-pass select for k > kK1MaxK (k3_sort.cu) ----------------------------------------------
// all_sims[n] -> out (k best), using a device radix sort.  temp storage is owned by the caller.
size_t sort_temp_bytes(long long n);
cudaError_t sort_select(const float* all_sims, long long n, int k, void* temp, size_t temp_bytes,
                        uint32_t* keys_a, uint32_t* keys_b, uint32_t* idx_a, uint32_t* idx_b,
                        u64* out_keys, long long* out_idx, float* out_sim, long long idx_stride,
                        long long idx_offset, cudaStream_t stream);

// ---- K4: merge `parts` key lists per query and decode (k4_merge.cu) -----------------------------
// keys laid out [parts][B][k_in]; writes B x k_out (k_out <= k_in is fine).
cudaError_t k4_merge(const u64* keys, int parts, long long B, int k_in, int k_out, u64* out_keys,
                     long long* out_idx, float* out_sim, long long idx_stride,
                     long long idx_offset, cudaStream_t stream, const unsigned* floor_ord = nullptr,
                     const unsigned* active_mask = nullptr);
// whether the batched search of B queries may take the screening route (query-major tiling, k2_qm.cu)
bool k2_screen_supported(int B, int k, int dim);
// capacity of the screening pass's per-thread lists for a top-k search (>= k, <= 32)
int k2_screen_list_len(int k);

// ---- K6: exact fp32 re-scoring of the screening pass's candidates, with the completeness proof (k6_rescore.cu)
struct K6Args {
  const float* db;
  long long n;
  int ld, dim;
  const float* q;        // raw fp32 queries [B][q_ld]
  int q_ld;
  const float* qpar;     // the k2_prepare parameter block
  const u64* part_keys;  // [parts][B][kl] lists of the screening pass
  int parts, B, k;
  int kl;                // capacity of those lists, k <= kl <= 32
  u64* out_keys;
  long long* out_idx;
  float* out_sim;
  long long idx_stride, idx_offset;
  unsigned long long* stats;  // [0] += queries screened, [1] += queries handed to the fallback
  u64* debug = nullptr;       // development: 64 words per query
  // seed mode (the lists come from a screening pass over every S-th row): write the kl-th largest screened
  // similarity of the union — a lower bound of the query's kl-th best over the WHOLE database, since the sample's
  // rows are real rows — into the first pass's bound plane, clear the sample pass's plane, and do nothing else
  bool seed = false;
};
cudaError_t k6_rescore(const K6Args& a, cudaStream_t stream);
// the per-query shared threshold plane the query-major batched kernel leaves behind ([B] ordered uints inside
// the inv_qn buffer), or null when the batch was served by the row-major kernel
const unsigned* k2_floor_plane(const float* inv_qn, int B, int dim);

// ---- K2: batched tcgen05 3xTF32 GEMM + fused per-query top-k (k2_batched.cu) --------------------
static constexpr int kK2MaxBatch = 256;  // queries per pass (UMMA N)
// row-major kernel: two accumulators (2 * padded batch columns) and >= 2 operand slots (64 columns each)
// share the 512 columns of tensor memory
static constexpr int kRowsMaxBatch = 192;
static constexpr int kK2ListBudget = 2560;  // k * padded batch must not exceed this (smem lists)

// Per-batch query parameter block written by k2_prepare: kQParPlanes planes of B 32-bit words, then a tail.
static constexpr int kQParInv = 0;       // 1 / max(|q|, eps)
static constexpr int kQParMean = 1;      // mean(q) (the centring constant)
static constexpr int kQParGthr = 2;      // shared k-th-best lower bound of the first pass (ordered uint)
static constexpr int kQParGthrFull = 3;  // the same for the masked 3xTF32 fallback pass
static constexpr int kQParEbound = 4;    // E_q: bound on |screened - exact| similarity
static constexpr int kQParPlanes = 5;
static constexpr int kQParTailWords = 16;  // [0..7] mask of queries the screening could not prove, [8] their count
static inline size_t k2_qpar_bytes(int B) { return ((size_t)kQParPlanes * B + kQParTailWords) * 4; }

struct K2Args {
  const float* db;
  long long n;
  int ld, dim;
  const float* q_hi;   // device [B][q_ld] raw fp32 queries
  const float* q_lo;   // device [B][q_ld] tf32 residuals (written by k2_prepare)
  const float* inv_qn; // device [B]
  int q_ld;
  int B;               // <= kK2MaxBatch
  int k;
  u64* part_keys;      // [grid][B][k] scratch
  void* encode_fn;     // cuTensorMapEncodeTiled entry point
  // query-major kernel only.  screen: ONE tf32 product (q_hi plane, raw rows), read-only converter — the
  // similarities are approximate and the lists feed k6_rescore.  active_mask / active_count (device): the
  // masked fallback — only the queries whose bit is set are searched, and the whole grid exits at once when
  // *active_count == 0 (the launch is queued unconditionally, without a host round trip).
  bool screen = false;
  const unsigned* active_mask = nullptr;
  const unsigned* active_count = nullptr;
  int gthr_plane = kQParGthr;  // which plane of the parameter block holds this pass's shared k-th-best bounds
};
bool k2_supported(int B, int k, int dim);
// number of partial key lists [parts][B][k] a launch for batch B writes (sizes part_keys, feeds k4_merge)
int k2_parts(long long n, int B, int dim, int num_sms);
// queries (device, pitch q_ld) -> q_lo, inv_qn
cudaError_t k2_prepare(const float* q, int B, int dim, int q_ld, float* q_lo, float* inv_qn,
                       cudaStream_t stream);
cudaError_t k2_launch(const K2Args& a, int num_sms, cudaStream_t stream, const char** err);
// row-major one-CTA kernel (k2_batched.cu): database rows on the MMA's M side, N = batch padded to 16;
// serves the small batches (k2_qm.cu routes)
bool k2s_supported(int B, int k);
int k2s_grid(long long n, int num_sms);
cudaError_t k2s_launch(const K2Args& a, int num_sms, cudaStream_t stream, const char** err);

// launch counter shared by all kernels (reported through cslam_nns_launch_count)
void count_launch(int n = 1);

}  // namespace cslam
