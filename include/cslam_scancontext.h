/*
 * cslam_scancontext.h — C-ABI of the B200-native ScanContext matcher.
 *
 * Drop-in boundary for the lidar sibling of the dense matcher,
 *   cslam/lidar_pr/scancontext_matching.py  ScanContextMatching.{__init__,add_item,search,search_best}
 * which LoopClosureSparseMatching instantiates for frontend.sensor_type == "lidar"
 * (cslam/loop_closure_sparse_matching.py:22-23,29-30).  The reference is numpy + scipy (a KD-tree over the
 * ring keys rebuilt on EVERY search, scancontext_matching.py:58-59, and a Python triple loop for the
 * column-shift distance, scancontext_utils.py:78-113); there is no FFI.  The entry points below are what a
 * ctypes binding placed in scancontext_matching.py would call (INTEGRATION.md shows the stub).
 *
 * Conventions: as cslam_nns.h — plain C, handle based, 0 on success, negative CSLAM_NNS_E* codes otherwise
 * (same values), message through cslam_sc_last_error; `stream` is a cudaStream_t passed as void*; no CPU
 * fallback.  All arithmetic is IEEE double, like the reference's numpy arrays
 * (scancontext_matching.py:18-19 allocate float64).
 */
#ifndef CSLAM_SCANCONTEXT_H_
#define CSLAM_SCANCONTEXT_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct cslam_sc_s* cslam_sc_t;

/* Message for the last failing call on `h` (or the last create failure when h == NULL). */
const char* cslam_sc_last_error(cslam_sc_t h);

/* Replaces ScanContextMatching.__init__ (scancontext_matching.py:9-22).
 * rings x sectors is the descriptor shape (reference default 20 x 60, :9), num_candidates the size of the
 * ring-key shortlist (:9, 10).  Storage starts at 1000 rows (:18-19) and doubles (:33-37) without copying
 * (same VMM arena as the dense matcher). */
int cslam_sc_create(int device, int rings, int sectors, int num_candidates, int64_t initial_capacity,
                    cslam_sc_t* out);
int cslam_sc_destroy(cslam_sc_t h);

/* Number of stored descriptors (`nb_items`, :22,:45). */
int64_t cslam_sc_size(cslam_sc_t h);

/* Replaces add_item's array writes (:31-43): append one descriptor (rings*sectors doubles, row-major
 * [ring][sector] like descriptor.reshape(shape), :31) and its ring key (rings doubles; the host computes
 * it exactly as the reference does, sc2rk = np.mean(sc, axis=1), scancontext_utils.py:75-76).  Host
 * pointers.  *row_out receives the row index (the reference's nb_items before the increment). */
int cslam_sc_add(cslam_sc_t h, const double* descriptor, const double* ringkey, void* stream,
                 int64_t* row_out);

/* Replaces search steps 1 and 2 (:57-76).
 * step 1: the num_candidates rows whose ring keys are nearest (Euclidean) to `query_ringkey`, nearest
 *         first — what KDTree(ringkeys).query(k=num_candidates) returns (:58-63) — by brute force;
 * step 2: for each, the column-shift distance of scancontext_utils.distance_sc (:78-113): for every
 *         shift 1..sectors, the mean cosine similarity of the non-zero column pairs; dist = 1 - max.
 * Host outputs, each num_candidates long, in shortlist order; missing candidates (fewer rows than
 * num_candidates) have cand_row = -1, dist = 1:
 *   cand_row[i]  row index;  cand_dist[i]  distance_sc;  cand_yaw[i]  argmax shift (1-based, :109)
 * The caller applies the reference's selection loop (:68-75, first strictly smaller distance wins).
 * Synchronous on return. */
int cslam_sc_search(cslam_sc_t h, const double* query_descriptor, const double* query_ringkey,
                    int64_t* cand_row, double* cand_dist, int32_t* cand_yaw, void* stream);

/* Device pointers of the stored descriptors [capacity][rings*sectors] and ring keys [capacity][rings]
 * (for tests and for callers that want to read the arrays the reference exposes as attributes). */
void* cslam_sc_descriptors_ptr(cslam_sc_t h);
void* cslam_sc_ringkeys_ptr(cslam_sc_t h);

#ifdef __cplusplus
}
#endif
#endif /* CSLAM_SCANCONTEXT_H_ */
