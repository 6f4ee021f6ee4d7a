/*
 * cslam_nns.h — C-ABI of the B200-native global-descriptor nearest-neighbour matcher.
 *
 * This is the drop-in boundary for ONE hot path of nhahn/cslam:
 *   cslam/nns_matching.py  NearestNeighborsMatching.{__init__,add_item,search,search_best}
 * The reference has no FFI for this path (it calls torch/ATen from Python, nns_matching.py:16-63);
 * the entry points below are what a ctypes binding placed in that file would call
 * (INTEGRATION.md shows the stub).  Each entry point cites the reference line(s) it replaces.
 *
 * Conventions
 *   - plain C, handle based, no global state; one handle == one descriptor database (one robot's DB),
 *     exactly like one NearestNeighborsMatching instance (loop_closure_sparse_matching.py:25,32).
 *   - every function returns 0 on success, a negative CSLAM_NNS_E* code otherwise;
 *     cslam_nns_last_error(handle) gives the message (handle may be NULL for create-time errors).
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).  The Python host
 *     passes torch.cuda.current_stream().cuda_stream.
 *   - host pointers are ordinary pageable or pinned memory; `*_is_device` says which side a pointer
 *     lives on.  Calls with host destinations are synchronous on return; calls with device
 *     destinations are stream-ordered and return immediately.
 *   - similarities are cosine similarities, fp32, computed as  x.q / (max(|x|,1e-8) * max(|q|,1e-8))
 *     (torch.nn.functional.cosine_similarity as called at nns_matching.py:61); results are ordered by
 *     similarity descending (nns_matching.py:62), ties broken by the lower row index, NaN first.
 *   - row indices are 0-based insertion order (the `self.n` counter of nns_matching.py:40-42).
 *   - there is NO CPU fallback: every compute entry point fails with CSLAM_NNS_ENODEV when no sm_100
 *     device is usable.
 */
#ifndef CSLAM_NNS_H_
#define CSLAM_NNS_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CSLAM_NNS_OK 0
#define CSLAM_NNS_EINVAL (-1)   /* bad argument (wrong dim, k<0, NULL pointer, ...) */
#define CSLAM_NNS_ENODEV (-2)   /* no usable CUDA device / wrong architecture */
#define CSLAM_NNS_ENOMEM (-3)   /* device or pinned allocation failed, VA reservation exhausted */
#define CSLAM_NNS_ECUDA (-4)    /* a CUDA runtime/driver call failed; message has the detail */
#define CSLAM_NNS_ESTATE (-5)   /* call not valid in this state (e.g. search before dim is known) */

typedef struct cslam_nns_s* cslam_nns_t;

/* ABI version of this header (bumped on any signature change). */
int cslam_nns_abi_version(void);

/* Number of kernels of this library launched so far in this process (all handles). bench.py reads it
 * before/after the timed region to report `gpu_launches`. */
int64_t cslam_nns_launch_count(void);

/* Message for the last failing call on `h` (or the last create failure when h == NULL). */
const char* cslam_nns_last_error(cslam_nns_t h);

/* -- lifetime ----------------------------------------------------------------------------------- */

/* Replaces NearestNeighborsMatching.__init__ (nns_matching.py:9-21).
 * device: CUDA ordinal.  dim: descriptor size, 0 = decide at first add (the reference's dim=None).
 * initial_capacity: rows of logical capacity to make available at once (reference: 1000, :21,:36);
 * <=0 selects 1000.  Capacity doubles like the reference (:38) but growth never copies the DB:
 * storage is a reserved virtual range into which physical chunks are mapped on demand. */
int cslam_nns_create(int device, int64_t dim, int64_t initial_capacity, cslam_nns_t* out);
int cslam_nns_destroy(cslam_nns_t h);

/* Fix the descriptor size of a lazily-dimensioned matcher (nns_matching.py:34-36). */
int cslam_nns_set_dim(cslam_nns_t h, int64_t dim);
/* Make room for at least `rows` rows (logical capacity keeps the 1000*2^j sequence). */
int cslam_nns_reserve(cslam_nns_t h, int64_t rows);

/* -- introspection (backs the public attributes n / dim / data, nns_matching.py:15-21) ---------- */
int64_t cslam_nns_size(cslam_nns_t h);      /* n */
int64_t cslam_nns_capacity(cslam_nns_t h);  /* len(self.data) */
int64_t cslam_nns_dim(cslam_nns_t h);       /* 0 while undecided */
int64_t cslam_nns_ld(cslam_nns_t h);        /* row pitch in floats (dim rounded up to 4) */
void* cslam_nns_data_ptr(cslam_nns_t h);    /* device pointer of row 0; changes only when the reserved virtual
                                               range itself is outgrown (re-map, no copy): re-read after adds */
int cslam_nns_device(cslam_nns_t h);

/* -- add_item (nns_matching.py:23-42) ----------------------------------------------------------- */

/* Append `count` rows of `dim` fp32 values (row pitch src_ld floats, >= dim; 0 means dim).
 * Stream-ordered with respect to later calls on the same stream.  Source lifetime:
 *   - HOST source, pageable or pinned: may be reused / freed as soon as the call returns (the semantics of the
 *     reference's synchronous `vector.to(device)`, nns_matching.py:41).  For a pinned (page-locked or
 *     cudaHostRegister'ed) source the copy engine reads the buffer asynchronously, so the call waits for THAT
 *     copy (an event, not the whole stream) before returning;
 *   - DEVICE source: stream-ordered — it may be overwritten by work queued on the same stream after the call;
 *     work on other streams (or the host) must first synchronise with `stream`.
 * *new_n (optional) receives the row count after the append. */
int cslam_nns_add(cslam_nns_t h, const float* rows, int64_t count, int64_t src_ld,
                  int src_is_device, void* stream, int64_t* new_n);

/* -- search / search_best (nns_matching.py:44-78) ----------------------------------------------- */

/* For each of the B queries (B x dim fp32, pitch q_ld floats, 0 = dim) return the kk = min(k, n)
 * most similar rows.  out_idx / out_sim are B x k (pitch k); entries [kk, k) of each row are
 * filled with -1 / -inf.  B == 1 runs the fused single-query kernel (HBM-bound); B >= 2 runs the
 * tcgen05 3xTF32 batched kernel when k fits its fused epilogue, otherwise one single-query pass per
 * query.  *out_kk (optional, host) receives kk. */
int cslam_nns_search(cslam_nns_t h, const float* queries, int64_t B, int64_t q_ld,
                     int src_is_device, int k, int64_t* out_idx, float* out_sim, int dst_is_device,
                     void* stream, int64_t* out_kk);

/* -- one keyframe against several databases ---------------------------------------- */

/* Replaces the per-robot loop of LoopClosureSparseMatching.add_local_global_descriptor
 * (loop_closure_sparse_matching.py:47-55): `count` databases (one per other robot) are searched with
 * the SAME query.  The query is staged once, one single-query kernel per non-empty database is
 * launched back to back on `stream`, all of them write into one mapped pinned result block, and the
 * host synchronises ONCE (the reference pays R-1 searches with two host syncs each).
 * hs[i] may be NULL or empty (kk 0).  All non-empty handles must live on one device and share dim.
 * query: q_len >= dim fp32 values.  out_idx / out_sim: count x k HOST arrays, entries past kk[i]
 * filled with -1 / -inf; out_kk (optional): count entries.  Synchronous on return. */
int cslam_nns_search_multi(const cslam_nns_t* hs, int count, const float* query, int64_t q_len,
                           int src_is_device, int k, int64_t* out_idx, float* out_sim,
                           int64_t* out_kk, void* stream);

/* -- candidate filter on the device ----------------------------------------------- */

/* Integer identities (keyframe ids) of rows [first_row, first_row + count): the `item` argument of
 * add_item when it is an int (nns_matching.py:40; loop_closure_sparse_matching.py:45,65 pass keyframe
 * ids).  Host-side bookkeeping only; the ids reach the device lazily, as one copy of the not yet
 * uploaded tail, at the next cslam_nns_search_filtered.  first_row must not leave a gap. */
int cslam_nns_set_row_ids(cslam_nns_t h, int64_t first_row, int64_t count, const int64_t* ids);

#define CSLAM_NNS_FILTER_BEST 0   /* loop_closure_sparse_matching.py:51,71: best hit iff sim >= threshold */
#define CSLAM_NNS_FILTER_INTRA 1  /* :83-97: drop a leading self hit, skip |id - query_id| < min_gap,
                                     skip sim < threshold, first survivor wins */
/* cslam_nns_search + the caller's candidate filter in one stream-ordered sequence: the top-k goes to
 * device memory, a filter kernel looks the rows up in the device id table, applies the rule of `mode`
 * and writes rows, similarities and the position of the accepted candidate (-1 = none) to mapped pinned
 * memory; one host synchronisation.  query_ids: B host values (the new keyframe's id per query) or NULL
 * (no self-hit / gap test).  threshold is compared in fp32 like numpy compares an np.float32 with a
 * python float.  Needs ids for every row (ESTATE otherwise).  out_idx/out_sim: B x k host arrays,
 * out_match_pos: B host ints.  k <= 128. */
int cslam_nns_search_filtered(cslam_nns_t h, const float* queries, int64_t B, int64_t q_ld,
                              int src_is_device, int k, int mode, const int64_t* query_ids,
                              int64_t min_gap, float threshold, int64_t* out_idx, float* out_sim,
                              int32_t* out_match_pos, void* stream, int64_t* out_kk);

/* -- row-sharded building blocks (no reference counterpart) -------------------------- */

/* Local top-k of this shard as packed 64-bit keys, written to DEVICE memory out_keys[B*k]
 * (stream-ordered).  key = orderable(sim) << 32 | ~global_row, global_row = local_row*row_stride +
 * row_offset (striped sharding: stride = world size, offset = rank).  Empty slots hold 0. */
int cslam_nns_search_keys(cslam_nns_t h, const float* queries, int64_t B, int64_t q_ld,
                          int src_is_device, int k, int64_t row_stride, int64_t row_offset,
                          uint64_t* out_keys_dev, void* stream);

/* Merge `parts` key lists per query (keys_dev laid out [parts][B][k], device memory, e.g. the
 * output of an all-gather of cslam_nns_search_keys results) into the global top-k.
 * Same output conventions as cslam_nns_search; valid entries per query = number of non-empty keys,
 * capped at k; *out_kk gets the count for query 0 when the destination is the host. */
int cslam_nns_merge_keys(int device, const uint64_t* keys_dev, int parts, int64_t B, int k,
                         int64_t* out_idx, float* out_sim, int dst_is_device, void* stream,
                         int64_t* out_kk);

/* -- exchange step of the sharded search as one kernel over NVLink peer memory -------- */
/* Replaces "all-gather of the packed keys (NCCL) + cslam_nns_merge_keys" by ONE kernel: every rank stores its
 * per-query keys directly into every rank's exchange buffer (peer memory opened once with CUDA IPC),
 * publishes them with release stores, waits for the other ranks' flags and merges.  The reference has no
 * multi-GPU path; the caller is cslam_b200/sharded.py.  One process per GPU, all ranks make the same calls. */
typedef struct cslam_nns_xchg_s* cslam_nns_xchg_t;
const char* cslam_nns_xchg_last_error(cslam_nns_xchg_t x);
/* Allocate this rank's exchange buffers on `device` (world <= 8). */
int cslam_nns_xchg_create(int device, int world, int rank, cslam_nns_xchg_t* out);
int cslam_nns_xchg_destroy(cslam_nns_xchg_t x);
/* 128 opaque bytes (two CUDA IPC memory handles) that the other ranks need to reach this rank's buffers;
 * the caller exchanges them (e.g. one all_gather at start-up). */
int cslam_nns_xchg_export(cslam_nns_xchg_t x, void* handle128);
/* `handles`: world x 128 bytes, entry r exported by rank r (the own entry is ignored). */
int cslam_nns_xchg_connect(cslam_nns_xchg_t x, const void* handles);
/* Same, for `world` exchange objects living in ONE process (single-GPU emulation in the tests). */
int cslam_nns_xchg_connect_local(cslam_nns_xchg_t x, const cslam_nns_xchg_t* all);
/* local_keys_dev: this rank's [B][k] packed keys (cslam_nns_search_keys).  Writes the merged global rows and
 * similarities [B][k] to device memory (-1 / -inf past the end).  k <= 32, B <= 4096, B * k <= 131072.
 * phase: 3 = push + wait + merge (production); 1 = push only, 2 = wait + merge only (tests, where one GPU
 * plays every rank and no launch may wait for another).  Stream-ordered, no host synchronisation.  A peer
 * that does not arrive within CSLAM_NNS_XCHG_TIMEOUT_S (20 s) makes the NEXT call fail with
 * CSLAM_NNS_ECUDA instead of hanging the GPU. */
int cslam_nns_xchg_merge(cslam_nns_xchg_t x, const uint64_t* local_keys_dev, int64_t B, int k, int phase,
                         int64_t* out_idx_dev, float* out_sim_dev, void* stream);
/* 0, or CSLAM_NNS_ECUDA when an earlier call timed out waiting for a peer (read after a stream sync). */
int cslam_nns_xchg_status(cslam_nns_xchg_t x);

/* This rank's part of ONE row-sharded search in a single call (SURVEY §8 b2 / e): the local top-k of shard `h`
 * as packed keys carrying the GLOBAL row (local_row * world + rank, the striped layout), then the exchange +
 * merge kernel over peer memory.  Equivalent to cslam_nns_search_keys into an internal buffer followed by
 * cslam_nns_xchg_merge; every rank makes the same call with the same queries.  Outputs are DEVICE arrays
 * [B][k] (global rows, similarities; -1 / -inf past the end), identical on every rank and bit-identical to an
 * unsharded search of the concatenated database.  Stream-ordered, no host synchronisation.  `h` may hold an
 * empty shard.  phase as in cslam_nns_xchg_merge: 3 in production; 1 = search + push, 2 = wait + merge (tests).
 * Limits of the exchange kernel apply (k <= 32, B <= 4096); beyond them use cslam_nns_search_keys, an
 * all-gather of the keys (NCCL) and cslam_nns_merge_keys. */
int cslam_nns_search_sharded(cslam_nns_t h, cslam_nns_xchg_t x, const float* queries, int64_t B, int64_t q_ld,
                             int src_is_device, int k, int rank, int world, int phase,
                             int64_t* out_idx_dev, float* out_sim_dev, void* stream);

/* -- single-query search that allocates nothing (NO counterpart in the reference)
 *    ---------------------------------------------------------------------------------------- */
#define CSLAM_NNS_STATIC_MAX_K 32      /* largest k of cslam_nns_search_static */
#define CSLAM_NNS_STATIC_MAX_CTAS 512  /* CTAs (= SMs of the device) its scratch is sized for */
/* For callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it): the k most similar rows of a CALLER-OWNED device database for one query, by the same single-query
 * kernel as cslam_nns_search(B = 1), with the kernel's scratch (a 256-byte ticket block + 512 x 32 packed keys,
 * ~129 KB) in a file-scope `static __device__` array instead of handle-owned buffers.  No handle; the call makes no
 * cudaMalloc / cudaFree / allocating library call.  Every GPU that loads the library holds the array until exit.
 *   db_dev: n rows of dim fp32 values, row pitch ld floats (0 = dim; >= dim, multiple of 4; the floats
 *   [dim, ld) of every row must be zero), 16-byte aligned, on `device`; query_dev: dim fp32 values, 16-byte aligned;
 *   out_idx_dev / out_sim_dev: k entries each in device memory (either may be NULL); entries [min(k, n), k) get
 *   -1 / -inf.  Stream-ordered on `stream`, no host synchronisation.
 * Limits (CSLAM_NNS_EINVAL with a message naming the limit otherwise): 1 <= k <= CSLAM_NNS_STATIC_MAX_K,
 * 1 <= n <= 2^32 - 2, a device with at most CSLAM_NNS_STATIC_MAX_CTAS SMs.
 * The array is SHARED by every call on the device and keeps the last call's contents (the call clears the part
 * that must start at zero on `stream` first): calls whose kernels can overlap on the GPU — different streams, or
 * concurrent host threads on different streams — RACE on it; order them on one stream or with events.
 * err / err_len (optional): receives a NUL-terminated message on failure. */
int cslam_nns_search_static(int device, const float* db_dev, int64_t n, int64_t ld, int64_t dim,
                             const float* query_dev, int k, int64_t* out_idx_dev, float* out_sim_dev,
                             void* stream, char* err, int64_t err_len);

/* -- tuning / diagnostics ----------------------------------------------------------------------- */

/* Force a search variant for tests and profiling: 0 = automatic, 1 = single-query kernel looped,
 * 2 = tcgen05 batched kernel (error if k does not fit), 3 = two-pass (all sims + radix sort). */
int cslam_nns_set_variant(cslam_nns_t h, int variant);
/* Variant actually used by the last search on this handle (1, 2 or 3; 0 = none yet). */
int cslam_nns_last_variant(cslam_nns_t h);
/* Screening route of the batched search (variant 2, batches the query-major kernel serves): one tf32 product
 * over the database, exact fp32 re-scoring of each query's candidates with a completeness proof, the 3xTF32 pass
 * only for the queries the proof does not cover.  Results obey the same parity rule either way.  enable = 0
 * makes every batch take the 3xTF32 pass; min_rows (>= 0) is the smallest database that is screened (-1 keeps
 * the current value, default 8192).  Environment: CSLAM_NNS_SCREEN=0, CSLAM_NNS_SCREEN_MIN_ROWS. */
int cslam_nns_set_screen(cslam_nns_t h, int enable, int64_t min_rows);
/* Counters since the handle was created: queries that went through the screening pass, and how many of them the
 * proof did not cover (recomputed by the 3xTF32 pass).  Synchronises the device. */
int cslam_nns_screen_stats(cslam_nns_t h, int64_t* screened, int64_t* fallback);
/* Development aid: copy an internal device buffer of the last batched search to the host (what = 0: the partial
 * key lists, 1: the per-query parameter block).  Synchronises the device. */
int cslam_nns_debug_copy(cslam_nns_t h, int what, void* dst, int64_t bytes);
/* Time in milliseconds (CUDA events on `stream`) of `iters` back-to-back device-resident searches of
 * the same query block: the kernel-only figure the roofline is computed from.  queries_dev is device
 * memory; results are discarded.  Returns the per-search average in *ms_per_search. */
int cslam_nns_time_search(cslam_nns_t h, const float* queries_dev, int64_t B, int64_t q_ld, int k,
                          int iters, void* stream, float* ms_per_search);

/* Per-kernel timing for the roofline: while enabled, every search on `h` brackets its DOMINANT kernel
 * (the single-query kernel, or the batched tcgen05 kernel) with CUDA events on the launch stream.
 * cslam_nns_profile_read synchronises, returns the summed duration of those launches in
 * milliseconds and their count, and resets the accumulators. */
int cslam_nns_profile(cslam_nns_t h, int enable);
int cslam_nns_profile_read(cslam_nns_t h, float* ms_total, int64_t* launches);

#ifdef __cplusplus
}
#endif
#endif /* CSLAM_NNS_H_ */
