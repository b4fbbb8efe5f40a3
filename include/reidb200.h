/* reidb200.h -- C ABI of libreidb200.so, the B200 (sm_100a) ReID matching hot path.
 *
 * Drop-in boundary for the reference SuperbTUM/real-time-ReID-tracking @ bb113fd.  The reference
 * has no FFI of its own (its hot path is plain Python calling NumPy / torch / faiss); every entry
 * point below names the reference function (file:line) whose arithmetic it replaces, and the
 * Python shims in real-time-reid-tracking_b200/ keep the reference call signatures on top of it.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless its name ends in _host; the caller owns all buffers;
 *   - every call is asynchronous on `stream` (a cudaStream_t passed as void*), takes no locks and
 *     keeps no global state except the last-error string (thread local);
 *   - return value: 0 = ok, negative = REIDB200_E* ; reidb200_last_error() describes the failure;
 *   - scratch memory is passed in by the caller, sized by the matching *_workspace_bytes();
 *   - there is no CPU fallback anywhere behind this interface.
 */
#ifndef REIDB200_H_
#define REIDB200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define REIDB200_OK 0
#define REIDB200_EINVAL (-1)  /* bad argument (shape, alignment, enum, workspace too small) */
#define REIDB200_ECUDA (-2)   /* CUDA runtime / driver error */
#define REIDB200_ELIMIT (-3)  /* size outside what this build supports */

const char* reidb200_last_error(void);
/* ABI version of this header (checked by the Python loader). */
int reidb200_abi_version(void);
/* 0 when the current device is sm_100 class with tcgen05/TMA, REIDB200_ECUDA otherwise. */
int reidb200_check_device(void);

/* Per-kernel timing for bench.py: when enabled every kernel launch of the library is bracketed by a CUDA event
 * pair on its launch stream; profile_read sums the spans recorded under `name` (kernel names as in csrc/). */
int reidb200_profile_enable(int on);
int reidb200_profile_read(const char* name, int* launches, double* total_ms);
int reidb200_profile_names(char* buf, int cap);

/* ------------------------------------------------------------------------------------------
 * (1) crop + resize + normalise
 *   CV_LINEAR     modification_deepsort/feature_extractor.py:31-46  Extractor._preprocess
 *                 (u8 -> f32 /255, cv2.resize INTER_LINEAR half-pixel no antialias, HWC->CHW,
 *                 (x-mean)/std)
 *   PIL_ANTIALIAS yolov8_tracking/boxmot/appearance/reid_multibackend.py:58-66,210-220
 *                 ReIDDetectMultiBackend._preprocess (PIL two-pass fixed-point bilinear on uint8,
 *                 /255, ImageNet normalise)
 *   crop boxes    yolov8_tracking/boxmot/trackers/strongsort/strong_sort.py:115-121,149-159
 *                 (img[y1:y2, x1:x2]; the int-truncated clamped xyxy is computed by the shim)
 * src        u8 bytes holding HWC (3-channel interleaved) pixels
 * crop_off   i64[M]  byte offset of pixel (0,0) of crop i inside src
 * crop_geom  i32[M,3] = (h, w, row_stride_bytes) of crop i (h,w >= 1)
 * out        f32[M,3,out_h,out_w]
 * ------------------------------------------------------------------------------------------ */
enum { REIDB200_RESIZE_CV_LINEAR = 0, REIDB200_RESIZE_PIL_ANTIALIAS = 1 };

int reidb200_crop_resize_norm(const uint8_t* src, const int64_t* crop_off, const int32_t* crop_geom,
                              int M, int out_h, int out_w, const float mean_host[3],
                              const float std_host[3], int mode, float* out, void* stream);
/* Same with a scratch buffer; required for PIL_ANTIALIAS (uint8 intermediate of the horizontal pass,
 * sized by reidb200_resize_pil_workspace_bytes), ignored for CV_LINEAR. */
int reidb200_crop_resize_norm_ws(const uint8_t* src, const int64_t* crop_off, const int32_t* crop_geom,
                                 int M, int out_h, int out_w, const float mean_host[3],
                                 const float std_host[3], int mode, float* out, void* workspace,
                                 size_t workspace_bytes, void* stream);
/* PIL mode only: the bit-exact uint8 resample (what PIL.Image.resize returns), u8[M,out_h,out_w,3]. */
int reidb200_resize_pil_u8(const uint8_t* src, const int64_t* crop_off, const int32_t* crop_geom,
                           int M, int out_h, int out_w, uint8_t* out_u8, void* workspace,
                           size_t workspace_bytes, void* stream);
size_t reidb200_resize_pil_workspace_bytes(int M, int max_crop_h, int out_w);

/* ------------------------------------------------------------------------------------------
 * (2) distance matrix, tcgen05 / TMA tensor-core GEMM with fp32-accurate (scaled fp16 hi|lo split, 3 UMMAs) operands
 *   EUCLID       reid/losses/utils.py:21-35   euclidean_dist: sqrt(clamp(xx+yy-2xy, 1e-12))
 *   COSINE_HALF  reid/losses/utils.py:12-18   cosine_dist: (1 - xy/(|x||y|))/2
 *   DOT          reid/evaluate.py:58          score = gf . q
 *   SQL2         reid/faiss_utils.py:97-111   faiss METRIC_L2 (squared L2): xx+yy-2xy
 * x f32[m,d] (row stride ldx), y f32[n,d] (ldy) -> out f32[m,n] (row stride ldo, in elements).
 * ------------------------------------------------------------------------------------------ */
enum { REIDB200_DIST_EUCLID = 0, REIDB200_DIST_COSINE_HALF = 1, REIDB200_DIST_DOT = 2,
       REIDB200_DIST_SQL2 = 3 };

size_t reidb200_distmat_workspace_bytes(int m, int n, int d);
int reidb200_distmat(const float* x, int m, int64_t ldx, const float* y, int n, int64_t ldy, int d,
                     int kind, float* out, int64_t ldo, void* workspace, size_t workspace_bytes,
                     void* stream);
/* Plain fp32 SIMT kernel with the same epilogues: the on-device cross-check of the tensor-core
 * path used by the GPU tests.  Not called by any product path. */
int reidb200_distmat_simt(const float* x, int m, int64_t ldx, const float* y, int n, int64_t ldy,
                          int d, int kind, float* out, int64_t ldo, void* stream);

/* Operand packing used by distmat/knn: f32[rows,d] -> fp16 hi|lo planes [rows, 2*dp] of the row scaled by a
 * power of two (dp = d rounded up to 64, zero padded; 22 of fp32's 24 significand bits survive), the fp32
 * squared norms of the unscaled rows (sqnorm f32[rows], may be NULL) and the inverse scales (iscale f32[rows]).
 * Exposed so a gallery can be packed once and searched many times (reidb200_knn_packed). */
size_t reidb200_pack_bytes(int rows, int d);
int reidb200_pack_rows(const float* x, int rows, int64_t ldx, int d, void* packed, float* sqnorm,
                       float* iscale, void* stream);

/* ------------------------------------------------------------------------------------------
 * (2b) brute-force k-NN (reid/faiss_utils.py:30-139: faiss bfKnn / GpuIndexFlatL2 / IndexFlatL2 /
 *      IndexShards stand-in; exact, ascending, ties -> lower index).  metric: SQL2 (ascending),
 *      DOT (descending inner product).
 * xq f32[nq,d], xb f32[nb,d] -> out_d f32[nq,k], out_i i64[nq,k] or i32[nq,k] (idx64 flag);
 * index_base is added to every returned index (gallery shard offset).  k <= 1024.
 * ------------------------------------------------------------------------------------------ */
size_t reidb200_knn_workspace_bytes(int nq, int nb, int d, int k);
int reidb200_knn(const float* xq, int nq, int64_t ldq, const float* xb, int nb, int64_t ldb, int d,
                 int k, int metric, int64_t index_base, float* out_d, void* out_i, int idx64,
                 void* workspace, size_t workspace_bytes, void* stream);
/* Same search against a gallery packed once with reidb200_pack_rows (packed_b = fp16 hi|lo planes, sqnorm_b and
 * iscale_b f32[nb], 16-byte aligned):
 * what a resident index does (GpuIndexFlatL2.add once, search many times, reid/faiss_utils.py:121-139) and what every
 * rank of the gallery-sharded search calls on its own shard.  No fp32 copy of the gallery is needed. */
size_t reidb200_knn_packed_workspace_bytes(int nq, int nb, int d, int k);
int reidb200_knn_packed(const float* xq, int nq, int64_t ldq, const void* packed_b, const float* sqnorm_b,
                        const float* iscale_b, int nb, int d, int k, int metric, int64_t index_base, float* out_d, void* out_i, int idx64,
                        void* workspace, size_t workspace_bytes, void* stream);
/* Row-wise top-k of a materialised matrix (ascending unless descending != 0); optional explicit
 * candidate ids (cand_i, i64[rows, cols]) for merging per-shard results, else id = column +
 * index_base.  Used for the per-shard merge after the NCCL all-gather. */
int reidb200_topk_rows(const float* vals, int rows, int cols, int64_t ldv, const int64_t* cand_i,
                       int k, int descending, int64_t index_base, float* out_d, void* out_i,
                       int idx64, void* stream);

/* ------------------------------------------------------------------------------------------
 * (4) CMC / mAP  (reid/evaluate.py:33-105  evaluate_all / evaluate / compute_mAP)
 * mat  f32[Q,G] (row stride ldm): similarity scores (descending = 1, the reference's gf.q) or
 *      distances (descending = 0, re-ranked evaluation)
 * ql,qc i32[Q]; gl,gc i32[G]
 * out_cmc f32[G] = sum_q cmc_q / Q ; out_ap f64[1] = sum_q ap_q / Q  (all Q, reid/evaluate.py:49-50)
 * out_ap_q f64[Q] (per-query AP, 0 for skipped) and out_first i32[Q] (first-hit rank, -1 =
 * no valid gallery match, the cmc[0] == -1 sentinel of reid/evaluate.py:80-82) may be NULL.
 * ------------------------------------------------------------------------------------------ */
size_t reidb200_eval_workspace_bytes(int Q, int G);
int reidb200_eval_cmc_map(const float* mat, int Q, int G, int64_t ldm, int descending,
                          const int32_t* ql, const int32_t* qc, const int32_t* gl, const int32_t* gc,
                          float* out_cmc, double* out_ap, double* out_ap_q, int32_t* out_first,
                          void* workspace, size_t workspace_bytes, void* stream);

/* Allocation-free variant of reidb200_eval_cmc_map.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same arguments without the workspace: the scratch (first-hit histogram, per-query AP /
 * first hit when out_ap_q / out_first are NULL, status words; 229 632 bytes) is a file-scope static __device__ array
 * inside the library, held by every GPU that loads it until exit.  Limits: Q <= REIDB200_EVAL_STATIC_MAX_Q and
 * G <= REIDB200_EVAL_STATIC_MAX_G, larger sizes are rejected with REIDB200_EINVAL (use reidb200_eval_cmc_map).  The call
 * allocates nothing (no cudaMalloc / cudaFree, no allocating library routine).  All calls on a device share the array
 * and it keeps the last call's contents: each call clears what must start at zero on `stream`, so calls on one stream
 * are safe back to back, while calls that overlap on the GPU (different streams, no ordering) race. */
#define REIDB200_EVAL_STATIC_MAX_Q 8192
#define REIDB200_EVAL_STATIC_MAX_G 32768
int reidb200_eval_cmc_map_static(const float* mat, int Q, int G, int64_t ldm, int descending,
                                 const int32_t* ql, const int32_t* qc, const int32_t* gl, const int32_t* gc,
                                 float* out_cmc, double* out_ap, double* out_ap_q, int32_t* out_first,
                                 void* stream);

/* ------------------------------------------------------------------------------------------
 * (3) k-reciprocal re-ranking (reid/faiss_utils.py:142-244 compute_jaccard_distance,
 *     k_reciprocal_neigh) on ONE merged set of N rows.
 * feats f32[N,d].  Output rows [row0, row0+nrows) x columns [col0, col0+ncols) of the N x N
 * Jaccard distance, f32 (or f16 when use_float16, the reference's mat_type) with row stride ldo.
 * lambda_value > 0 blends the classic way (not in the reference):
 *      out = (1-lambda) * jaccard + lambda * sqrt(max(sql2, 1e-12))
 * rank_out (optional) i32[N,k1]: the initial ranking (stage D2).
 * ------------------------------------------------------------------------------------------ */
size_t reidb200_jaccard_workspace_bytes(int N, int d, int k1, int k2);
/* The N x N squared-L2 matrix is never materialised (the initial ranking is selected inside the GEMM epilogue); only the
 * rows the lambda blend reads are kept.  blend_rows = nrows of the output region when lambda_value > 0, else 0: the
 * workspace for exactly that call (reidb200_jaccard_workspace_bytes covers any region: blend_rows = N). */
size_t reidb200_jaccard_region_workspace_bytes(int N, int d, int k1, int k2, int blend_rows);
int reidb200_jaccard(const float* feats, int N, int64_t ldf, int d, int k1, int k2, int use_float16,
                     float lambda_value, int row0, int nrows, int col0, int ncols, void* out,
                     int64_t ldo, int32_t* rank_out, void* workspace, size_t workspace_bytes,
                     void* stream);
/* Stages D3-D7 only, from a caller-supplied initial ranking i32[N,k1] and the materialised
 * squared-L2 matrix f32[N,N] (row stride lds). */
int reidb200_jaccard_from_rank(const int32_t* rank, const float* sql2, int64_t lds,
                               const float* sqnorm, int N, int k1, int k2, int use_float16,
                               float lambda_value, int row0, int nrows, int col0, int ncols,
                               void* out, int64_t ldo, void* workspace, size_t workspace_bytes,
                               void* stream);
/* Row-sharded re-ranking (one rank of a multi-GPU run; also reid/faiss_utils.py:172-181 search_option 2 / 3,
 * where the initial ranking comes from an index object): stages D3-D8 for the output rows [row0, row0 + nrows) x columns
 * [col0, col0 + ncols) from a GIVEN global initial ranking rank_i i32[N,k1] with its squared-L2 values rank_d f32[N,k1].
 * No squared-L2 matrix of the merged set is needed: the expansion's pair values come from the rows' own top-k1 lists or a
 * direct dot product of the fp32 feature rows.  blend_sql2 f32[nrows, ldb] = squared L2 of exactly the output block (only
 * read when lambda_value > 0).  Either `out` (the block of the Jaccard / re-ranked distance, f32 or f16) or the fused
 * evaluation (hist != NULL: the output rows are queries, the columns the gallery; ql / qc i32[nrows], gl / gc
 * i32[ncols]; hist i32[ncols] first-hit histogram is ADDED to, ap_q f64[nrows], first i32[nrows], status i32[64]) or both. */
size_t reidb200_rerank_rows_workspace_bytes(int N, int d, int k1, int k2);
int reidb200_rerank_rows_from_rank(const float* feats, int N, int64_t ldf, int d, const int32_t* rank_i,
                                   const float* rank_d, int k1, int k2, int use_float16, float lambda_value,
                                   int row0, int nrows, int col0, int ncols, const float* blend_sql2, int64_t ldb,
                                   void* out, int64_t ldo, const int32_t* ql, const int32_t* qc, const int32_t* gl,
                                   const int32_t* gc, int32_t* hist, double* ap_q, int32_t* first, int32_t* status,
                                   void* workspace, size_t workspace_bytes, void* stream);
/* R(i,k) lists for inspection/tests: out_idx i32[N,k+1] (-1 padded), out_cnt i32[N]
 * (reid/faiss_utils.py:142-146). */
int reidb200_k_reciprocal(const int32_t* rank, int N, int k1_cols, int k, int32_t* out_idx,
                          int32_t* out_cnt, void* stream);

/* ------------------------------------------------------------------------------------------
 * (5) DeepSort / StrongSort association cost, one fused kernel per frame.
 *   gallery     boxmot/utils/matching.py:363-409  NearestNeighborDistanceMetric
 *               (partial_fit keeps the last `budget` features per track id; distance = per-track
 *               min over samples of the cosine distance, rows re-normalised, :278-339)
 *   gate        boxmot/motion/adapters/strongsort_kf_adapter.py:137-169 -> kalman_filter.py:577-586
 *               sqrt(y^T SI y), y = z - x[:4]  (stale SI)
 *               boxmot/trackers/strongsort/sort/linear_assignment.py:147-198 gate_cost_matrix,
 *               :58-59 clamp of min_cost_matching
 *   iou         boxmot/trackers/strongsort/sort/iou_matching.py:9-86 (iou, iou_cost) and the DIoU of
 *               modification_deepsort/iou_matching.py:5-47
 * The gallery is an opaque device-resident ring buffer [slots, budget, d].
 * ------------------------------------------------------------------------------------------ */
typedef struct reidb200_gallery reidb200_gallery;

int reidb200_gallery_create(int slots, int budget, int d, reidb200_gallery** out);
int reidb200_gallery_destroy(reidb200_gallery* g);
/* Grow a gallery (more slots and / or a longer ring per slot; never shrinks), contents preserved bit for bit: the
 * reference's sample store is unbounded (budget=None keeps every sample, any number of targets, matching.py:363-389).
 * Synchronises `stream`. */
int reidb200_gallery_resize(reidb200_gallery* g, int new_slots, int new_budget, void* stream);
/* Append n feature rows f32[n,d]; row i goes to slot slot_of_row[i] (i32[n], device).  Rows for one
 * slot are appended in row order; only the last `budget` survive (matching.py:385-386). */
int reidb200_gallery_partial_fit(reidb200_gallery* g, const float* feats, int64_t ldf,
                                 const int32_t* slot_of_row, int n, void* stream);
/* Forget slots (tracks that are no longer active, matching.py:387-389). slots i32[n] device. */
int reidb200_gallery_reset_slots(reidb200_gallery* g, const int32_t* slots, int n, void* stream);
/* counts i32[slots] (device copy out), for tests. */
int reidb200_gallery_counts(reidb200_gallery* g, int32_t* out_counts, void* stream);
/* sample rows of a slot as stored (L2-normalised), f32[budget,d], oldest first; for tests. */
int reidb200_gallery_read_slot(reidb200_gallery* g, int slot, float* out_rows, void* stream);

enum { REIDB200_IOU_PLAIN = 0, REIDB200_IOU_DIOU = 1 };
enum { REIDB200_ASSOC_CLAMP_APP = 1, REIDB200_ASSOC_CLAMP_IOU = 2, REIDB200_ASSOC_NO_GATE = 4,
       REIDB200_ASSOC_SIMT = 8 /* plain fp32 SIMT kernel instead of the tcgen05 one: on-device cross-check for the tests */ };

/* One frame:
 *   det_feat f32[M,d]; det_tlwh f32[M,4];
 *   trk_slot i32[T] gallery slot per track row; trk_mean f64[T,8] KF mean (xyah + velocities);
 *   trk_SI f64[T,16] inverse innovation covariance as stored on the track (stale, F6);
 *   trk_tsu i32[T] time_since_update.
 * Outputs (f64[T,M] each, row-major):
 *   out_app   gated+blended appearance cost (what gate_cost_matrix returns; with
 *             REIDB200_ASSOC_CLAMP_APP also `cost > max_dist -> max_dist + 1e-5`)
 *   out_raw   (optional) NearestNeighborDistanceMetric.distance before gating
 *   out_iou   iou_cost for every (track,det) pair (rows with tsu > 1 = 1e5; with
 *             REIDB200_ASSOC_CLAMP_IOU also clamped at max_iou_dist + 1e-5)
 */
int reidb200_assoc_frame(reidb200_gallery* g, const float* det_feat, int64_t ldf,
                         const float* det_tlwh, int M, const int32_t* trk_slot,
                         const double* trk_mean, const double* trk_SI, const int32_t* trk_tsu, int T,
                         double mc_lambda, double max_dist, double max_iou_dist, int iou_mode,
                         int flags, double* out_app, double* out_raw, double* out_iou, void* stream);

/* ------------------------------------------------------------------------------------------
 * One tracking frame FROM HOST ARRAYS in two calls (crops before the backbone, costs after it), the way the reference's
 * per-frame code hands over its data (strong_sort.py:149-159 _get_features; sort/tracker.py:175-233 _match): plain host
 * pointers in, each call ONE pinned host->device copy of the small per-frame tables + the kernels (for large frames the
 * cost tables go up on the context's own side stream while `stream` is still busy with the crops / the embedding network);
 * the association kernel writes both float64 cost matrices straight into mapped pinned memory (they cross PCIe while the
 * kernel runs: no device->host copy behind it), frame_costs ends with a stream synchronisation and returns pointers into
 * that pinned result (valid until the call after the next).  No per-frame allocation.
 *   frame      u8[H,W,3] on the DEVICE (where the detector left it); xyxy_host i32[n,4] crop boxes img[y1:y2, x1:x2]
 *              (slice ends are clipped at the frame border as numpy does; an empty crop is REIDB200_EINVAL)
 *   det_feat   f32[M,d] on the DEVICE (the backbone's output); everything else is host memory:
 *   det_tlwh_host f32[M,4], trk_slot_host i32[T], trk_mean_host f64[T,8], trk_SI_host f64[T,16], trk_tsu_host i32[T]
 * out_app_host / out_iou_host receive f64[T,M] (row-major) as reidb200_assoc_frame defines them. */
typedef struct reidb200_frame reidb200_frame;
int reidb200_frame_create(reidb200_gallery* g, int max_dets, int max_tracks, reidb200_frame** out);
int reidb200_frame_destroy(reidb200_frame* f);
int reidb200_frame_crops(reidb200_frame* f, const uint8_t* frame, int H, int W, const int32_t* xyxy_host, int n,
                         int out_h, int out_w, const float mean_host[3], const float std_host[3], float* out,
                         void* stream);
int reidb200_frame_costs(reidb200_frame* f, const float* det_feat, int64_t ldf, const float* det_tlwh_host, int M,
                         const int32_t* trk_slot_host, const double* trk_mean_host, const double* trk_SI_host,
                         const int32_t* trk_tsu_host, int T, double mc_lambda, double max_dist, double max_iou_dist,
                         int iou_mode, int flags, const double** out_app_host, const double** out_iou_host,
                         void* stream);

/* Stand-alone pieces of the same kernel family (the reference's free functions):
 *   iou / DIoU of T boxes (f64 tlwh) against M candidates (f32 or f64 tlwh) -> f64[T,M] */
int reidb200_iou_matrix(const double* bbox_tlwh, int T, const void* cand_tlwh, int cand_is_f64, int M,
                        int iou_mode, double* out, void* stream);
/*   Mahalanobis gate distances f64[T,M] from x4 f64[T,4], SI f64[T,16], measurements f64[M,4]
 *   (the values are whatever the caller passes: Detection.to_xyah() is float32-rounded) */
int reidb200_gating_distance(const double* x4, const double* SI, const double* meas, int T, int M,
                             double* out, void* stream);
/*   cosine distance 1 - a^.b^T (both re-normalised unless normalized != 0), f32[na,nb]
 *   (boxmot/utils/matching.py:278-298) */
size_t reidb200_cosine_distance_workspace_bytes(int na, int nb, int d);
int reidb200_cosine_distance(const float* a, int na, int64_t lda, const float* b, int nb, int64_t ldb,
                             int d, int normalized, float* out, void* workspace,
                             size_t workspace_bytes, void* stream);

/* Re-ranked evaluation in one call (stages 3 + 4 fused): feats f32[G+Q, d] = gallery rows first, then query rows
 * (the merged set of image_reid_inference.py:270); the Q x G block of the lambda-blended Jaccard distance is reduced
 * to CMC f32[G] / mAP f64[1] by the CTA that produces each query row, so the re-ranked matrix is not re-read and, with
 * out_dist == NULL, never written.  G <= 16384 (wider galleries: reidb200_jaccard + reidb200_eval_cmc_map).
 * The N x N squared-L2 matrix inside the workspace is kept in TRIANGLE mode when ceil(N / 32) >= max(64, 4 k1) and
 * N <= 32768 (N >= 2529 for k1 = 20; above 32768 rows there is no matrix at all) (every element stored once,
 * the row select reads the other half from two words per 32-column chunk): same bits in every output
 * as with the full matrix, which the environment variable REIDB200_TRI=0 restores.
 * replaces reid/faiss_utils.py:149-244 followed by reid/evaluate.py:33-105 on the re-ranked distance. */
size_t reidb200_rerank_eval_workspace_bytes(int N, int d, int k1, int k2, int Q, int G);
int reidb200_rerank_eval(const float* feats, int G, int Q, int64_t ldf, int d, int k1, int k2, float lambda_value,
                         const int32_t* ql, const int32_t* qc, const int32_t* gl, const int32_t* gc,
                         float* out_dist, int64_t ldo, float* out_cmc, double* out_ap, void* workspace,
                         size_t workspace_bytes, void* stream);
/* Same, also returning what reid/evaluate.py:55-75 evaluate() returns per query: out_ap_q f64[Q] (0 for a skipped query)
 * and out_first i32[Q] (first-hit rank in the junk-free ranking, -1 = the cmc[0] == -1 sentinel); either may be NULL. */
int reidb200_rerank_eval_pq(const float* feats, int G, int Q, int64_t ldf, int d, int k1, int k2, float lambda_value,
                            const int32_t* ql, const int32_t* qc, const int32_t* gl, const int32_t* gc,
                            float* out_dist, int64_t ldo, float* out_cmc, double* out_ap, double* out_ap_q,
                            int32_t* out_first, void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------
 * N2: pairwise box costs of the other boxmot trackers.
 *   modes 0-3  iou_batch / giou_batch / diou_batch / ciou_batch of xyxy boxes, computed and returned in the
 *              inputs' dtype (float32 or float64)      replaces boxmot/utils/iou.py:4-164
 *   mode 4     bbox_ious: the "+1" pixel IoU on float32 tlbr boxes -> float32
 *   mode 5     iou_distance = 1 - bbox_ious                                replaces boxmot/utils/matching.py:71-114,216-248
 * boxes_a [n,4], boxes_b [k,4], out [n,k] (row-major, dense). */
enum {
  REIDB200_BOX_IOU = 0, REIDB200_BOX_GIOU = 1, REIDB200_BOX_DIOU = 2, REIDB200_BOX_CIOU = 3,
  REIDB200_BOX_PLUS_ONE = 4, REIDB200_BOX_IOU_DISTANCE = 5
};
int reidb200_box_pairs(const void* boxes_a, int n, const void* boxes_b, int k, int is_f64, int mode, void* out,
                       void* stream);
/*   fuse_score (det_scores != NULL): out f64[n,k] = 1 - (1 - cost) * score[col]
 *   fuse_iou   (tlbr_a, tlbr_b != NULL): out (cost's dtype) = 1 - (1 - cost) * (1 + iou_sim) / 2, iou_sim from bbox_ious
 *   cost f32 or f64 [n,k]                                                  replaces boxmot/utils/matching.py:189-213 */
int reidb200_fuse_cost(const void* cost, int cost_is_f64, int n, int k, const double* det_scores,
                       const float* tlbr_a, const float* tlbr_b, void* out, void* stream);

/*   OC-SORT / DeepOCSORT association cost, one launch per frame: out_iou f64[D,T] = iou_batch(detections, trackers),
 *   out_angle f64[D,T] = the velocity-direction-consistency term (valid * (pi/2 - |acos|) / pi * vdc_weight * score).
 *   detections f64[D,5] (x1,y1,x2,y2,score), trackers f64[T,5], velocities f64[T,2] (dy,dx), previous_obs f64[T,5]
 *   (score < 0 = no previous observation)                replaces boxmot/utils/association.py:6-16,107-140 */
int reidb200_ocsort_cost(const double* detections, int D, const double* trackers, int T, const double* velocities,
                         const double* previous_obs, double vdc_weight, double* out_iou, double* out_angle,
                         void* stream);

/* ------------------------------------------------------------------------------------------
 * N3: StrongSORT's Kalman filter for many tracks at once, device resident.
 * State arrays (caller-owned, indexed by track slot): x f64[slots,8], P f64[slots,64], SI f64[slots,16]; x and SI feed
 * reidb200_assoc_frame (trk_mean / trk_SI) without leaving the device.  `slots` i32[n] lists the tracks touched;
 * measurements f64[n,4] are xyah (Detection.to_xyah() values promoted to double), confidences f64[n] or NULL (= 0).
 *   initiate  replaces StrongSortKalmanFilterAdapter.initiate (strongsort_kf_adapter.py:25-58); SI := 0
 *   predict   replaces .predict (:60-94) + KalmanFilter.predict (kalman_filter.py:340-379); slots may be NULL (= 0..n-1)
 *   update    replaces .update (:96-135) + KalmanFilter.update (kalman_filter.py:486-575), writes SI */
int reidb200_kf_initiate(double* x, double* P, double* SI, const int32_t* slots, const double* measurements, int n,
                         double w_pos, double w_vel, void* stream);
int reidb200_kf_predict(double* x, double* P, const int32_t* slots, int n, double dt, double w_pos, double w_vel,
                        void* stream);
int reidb200_kf_update(double* x, double* P, double* SI, const int32_t* slots, const double* measurements,
                       const double* confidences, int n, double w_pos, void* stream);

/* ------------------------------------------------------------------------------------------
 * N4: the steps around compute_jaccard_distance in reid/image_reid_inference.py, on the device-resident
 * N x N matrix and embeddings.
 *   tta_merge_normalize  out = F.normalize((a + b) / 2, dim=1) (image_reid_inference.py:252-253, 268-269); b may be NULL
 *                        (plain row normalisation)
 *   clamp_add            dist[dist < 0] = 0 ; dist += other (:287-290); other may be NULL
 *   dbscan_precomputed   sklearn.cluster.DBSCAN(eps, min_samples, metric="precomputed").fit_predict(dist) (:291-301) on a
 *                        symmetric f32[N, N] matrix: labels i32[N] (-1 = noise, clusters numbered by their lowest core
 *                        index as sklearn's search does), n_clusters i32[1] (device, may be NULL), core_out i32[N] 0/1
 *                        (may be NULL).  N <= 57984 (one CTA numbers the clusters).
 *   smooth_groups        smooth_tracklets (reid/inference_utils.py:19-27): rows order[seg_start[g] .. seg_start[g+1]) of
 *                        emb become emb * keep + mean(rows) * avg_w (reference: 0.1 / 0.9) */
int reidb200_tta_merge_normalize(const float* a, size_t lda, const float* b, size_t ldb, int rows, int D, float* out,
                                 size_t ldo, void* stream);
int reidb200_clamp_add(float* dist, int rows, int cols, size_t ld, const float* other, size_t ldo, void* stream);
size_t reidb200_dbscan_workspace_bytes(int N);
int reidb200_dbscan_precomputed(const float* dist, int N, size_t ld, float eps, int min_samples, int32_t* labels,
                                int32_t* n_clusters, int32_t* core_out, void* ws, size_t ws_bytes, void* stream);
int reidb200_smooth_groups(float* emb, int D, size_t lde, const int32_t* order, const int32_t* seg_start, int n_groups,
                           float keep, float avg_w, void* stream);
/* out = x - column mean(x), mean_out f32[D] optional: the centring step of diminish_camera_bias
 * (reid/inference_utils.py:9-10); its Gram matrix and projection go through reidb200_distmat(kind DOT). */
int reidb200_center_columns(const float* x, int rows, int D, size_t ldx, float* out, size_t ldo, float* mean_out,
                            void* stream);

/* ------------------------------------------------------------------------------------------
 * N1: min_cost_matching's assignment on the device-resident cost matrix
 * (boxmot/trackers/strongsort/sort/linear_assignment.py:58-77: clamp to max_distance + 1e-5,
 * scipy.optimize.linear_sum_assignment, pairs above max_distance dropped).
 * cost f64[T, M] (row stride ld); the solve takes the rows row_cand[0..nrc) whose match_col_of_row entry is still -1 and
 * the columns whose det_matched entry is still 0, and records its matches in both arrays (match_col_of_row[row] = col,
 * det_matched[col] = 1), so that the stages of Tracker._match chain on the device.  The admissible entries
 * (cost <= max_distance) split into connected components, each solved by one warp with scipy's shortest-augmenting-path
 * algorithm in float64: the same matches as scipy whenever the optimum is unique.  status i32[1] (device) is set to 1 if a
 * solve could not finish (non-finite costs).  At most 2048 rows / columns per solve. */
size_t reidb200_lap_workspace_bytes(int T, int M);
int reidb200_lap_gated(const double* cost, int T, int M, size_t ld, const int32_t* row_cand, int nrc, double max_distance,
                       int32_t* match_col_of_row, int32_t* det_matched, int32_t* status, void* ws, size_t ws_bytes, void* stream);
/* Exact mode: the assignment scipy.optimize.linear_sum_assignment returns for the clamped matrix of the candidate rows x
 * candidate columns, TIES INCLUDED (one warp runs _lsap.cpp statement by statement), and with it the reference's
 * `unmatched_detections` list in the reference's order (linear_assignment.py:61-76: the pairs scipy forms through gated
 * entries decide where a detection lands in that list, hence which id a new track gets, tracker.py:175-233).
 *   row_cand i32[nrc]: rows in the reference's track_indices order (rows already matched are skipped)
 *   col_cand i32[<= max_cols], col_count i32[1] or NULL: columns in the reference's detection_indices order (device list +
 *   device count, so that the list one solve writes feeds the next without a host round trip)
 *   um_dets i32[>= max_cols], um_count i32[1]: out.  um_rows i32[>= nrc], um_rows_count i32[1]: out, the reference's
 *   `unmatched_tracks` list in its order (both NULL: not wanted).  Matches go to match_col_of_row / det_matched like
 *   lap_gated. */
int reidb200_lap_exact(const double* cost, int T, int M, size_t ld, const int32_t* row_cand, int nrc,
                       const int32_t* col_cand, const int32_t* col_count, int max_cols, double max_distance,
                       int32_t* match_col_of_row, int32_t* det_matched, int32_t* um_dets, int32_t* um_count,
                       int32_t* um_rows, int32_t* um_rows_count, int32_t* status, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* REIDB200_H_ */
