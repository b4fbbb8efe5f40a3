// Internal interface of the split-fp16 tensor-core GEMM (gemm.cu), shared by distmat / knn /
// jaccard / association.
#pragma once
#include <cuda.h>

#include "common.cuh"

namespace reidb200 {

// K (feature dimension) rounded up to the 64-element k-block of the GEMM.
inline int padded_dim(int d) { return (d + 63) / 64 * 64; }
// bytes of the packed hi|lo fp16 planes of `rows` rows
inline size_t packed_bytes(int rows, int d) { return (size_t)rows * 2 * padded_dim(d) * sizeof(__half); }

// f32[rows,d] -> fp16 [rows, 2*dp] (hi plane then lo plane per row, the row scaled by a power of two) + squared norms of
// the unscaled rows (may be null) + the inverse scale per row.
int pack_rows_launch(const float* x, int rows, int64_t ldx, int d, void* packed, float* sqnorm, float* iscale,
                     cudaStream_t st);

// out[m,n] = epilogue(kind, isa[i]*isb[j]*(A.B^T), xx, yy) with A,B packed operands.  xx/yy may be null for DOT.
// chunk_red (optional) f32[m, ld_chunk_red]: min (max when chunk_red_max) of the outputs over every 32-column chunk,
// consumed by topk_rows_launch to skip the chunks that cannot hold a top-k element.
int gemm_split_launch(const void* packedA, const float* xx, const float* isa, int m, const void* packedB, const float* yy,
                      const float* isb, int n, int d, int kind, float* out, int64_t ldo, cudaStream_t st,
                      float* chunk_red = nullptr, int64_t ld_chunk_red = 0, int chunk_red_max = 0, int blocked = 0);

// Self squared-L2 product in TRIANGLE mode (gemm.cu GemmArgs::lower): element (i, c) of out[n, n] is stored iff
// (i >> 8) >= (c >> 8) (256-row tile blocks), bit-identical to gemm_split_launch's value; chunk_red f32[n, ld] holds the
// minimum of EVERY (row, 32-column chunk) and chunk_red2 u32[n, ld], for the chunks of the missing half,
// (ordered key of the chunk's second smallest value & ~31) | position of its minimum.  Needs n > 128 and 16-byte aligned
// rows (tri_mode_ok); consumed by topk_rows_launch(..., chunk_red2) and PairSrc::tri in rerank.cu.
int gemm_split_tri_launch(const void* packed, const float* sq, const float* isc, int n, int d, float* out, int64_t ldo,
                          cudaStream_t st, float* chunk_red, uint32_t* chunk_red2, int64_t ld_chunk_red);
// whether an n x n self product with a k-NN of k can run in triangle mode (the CTA-pair kernel and the chunk-guided selection)
inline bool tri_mode_ok(int n, int k) { return n > 128 && n <= 32768 && (n + 31) / 32 >= (4 * k > 64 ? 4 * k : 64) && k <= 128; }  // (n <= 32768: 32-bit element offsets in topk.cu)
// element (i, c) of a triangle-mode matrix is read at (c, i) when its own place was not stored
__host__ __device__ inline bool tri_stored(int i, int c) { return (i >> 8) >= (c >> 8); }

// epilogue formulas of the four kinds (shared by the GEMM, its SIMT cross-check and the selection redo)
#ifdef __CUDACC__
__device__ __forceinline__ float dist_epilogue(int kind, float acc, float xx, float yy) {
  switch (kind) {
    case REIDB200_DIST_DOT:
      return acc;
    case REIDB200_DIST_SQL2:
      return fmaf(-2.0f, acc, __fadd_rn(xx, yy));
    case REIDB200_DIST_EUCLID:
      return sqrtf(fmaxf(fmaf(-2.0f, acc, __fadd_rn(xx, yy)), 1e-12f));
    default: {  // COSINE_HALF: (1 - xy / (|x| |y|)) / 2
      const float den = __fmul_rn(sqrtf(xx), sqrtf(yy));
      return __fmul_rn(__fsub_rn(1.0f, __fdiv_rn(acc, den)), 0.5f);
    }
  }
}

#endif

// ---- selection fused into the GEMM epilogue (no distance block): per output row an append-only candidate list, a small
// SKETCH and a threshold that only tightens.
//   tau    u32[m]      ordered key of an upper bound of the row's k-th best value (init 0xFFFFFFFF).  Invariant: at least k
//                      elements of the row have key <= tau, so every element of the true top-k passes `key <= tau` whenever
//                      it is produced, and the exact top-k is the k best entries of the list (topk.cu select_lists_kernel).
//   list   u64[m, cap] (key << 32) | column, reserved through resv u32[m]; a row whose resv exceeds cap is redone exactly.
//   sketch u64[m, W]   slot (column mod W) holds the smallest (key << 32 | column) appended for that residue class
//                      (atomicMin, init all ones).  Occupied slots are DISTINCT elements of the row, so the k-th smallest
//                      slot is a valid threshold; with W ~ 3k it sits near the rank -W ln(1 - k / W) ~ 1.2 k element of
//                      everything produced so far, however much that is -- no ordering or commit protocol between the
//                      appending warps is needed (an earlier version read back the list and had to skip whenever an append
//                      was in flight: under the first wave's contention the thresholds never left +inf).
// Bootstrap: a threshold starts at +inf, so the first tile a row meets appends all of its columns.  To keep that flood to
// ONE tile per row, the first tile of every row block in the kernel's tile order (the diagonal tile of a symmetric
// product, column tile 0 otherwise) is its bootstrap tile: epilogue warps of every other tile wait (gate[]) until the
// bootstrap tiles of their row block -- and, for a mirrored tile, of their column block -- have finished and tightened.
// Bootstrap tiles never wait and come first in every CTA's work list, so the wait cannot deadlock in the persistent grid.
// Symmetric products (A == B) use the same arrays for the mirrored element (row = column index, entry = row index).
struct SelArgs {
  uint32_t* tau;
  uint32_t* resv;
  unsigned long long* list;
  unsigned long long* sketch;
  uint32_t* gate;                     // u32[row blocks], zero-init: epilogue warps that finished the row block's BOOTSTRAP tile
  int cap, k, W, window, descending;  // W: sketch slots (multiple of 32, >= k); window: appends between two tightenings
  int debug;                          // timing experiments only (REIDB200_SEL_DEBUG): 1 no appends, 2 no tightening, 4 no gates, 8 no boot work, 16 no slow path
  int st_r0, st_r1, st_c0, st_c1;     // block of the output that is still STORED (rows [r0, r1) x columns [c0, c1)); empty = none
};

int gemm_split_select_launch(const void* packedA, const float* xx, const float* isa, int m, const void* packedB, const float* yy,
                             const float* isb, int n, int d, int kind, float* out, int64_t ldo, const SelArgs& sel, cudaStream_t st);

// candidate-list capacity / tightening window the selection uses for a given k (0: k not supported, use the materialising path)
inline int sel_sketch(int k) { return k <= 21 ? 64 : (k <= 42 ? 128 : (k <= 100 ? 256 : (k <= 128 ? 512 : 0))); }  // slots W >= 2.5 k, a power of two <= 512
inline int sel_cap(int k) { return k <= 25 ? 1024 : (k <= 128 ? 4096 : 0); }
size_t sel_workspace_bytes(int m, int n, int k);
// carve + initialise (stream-ordered) the selection state of m rows inside `ws`
int sel_prepare(void* ws, size_t ws_bytes, int m, int n, int k, int descending, SelArgs* out, float** scratch_rows, cudaStream_t st);
// the k best entries of every row's list -> out_d f32[m, k], out_i (i64 | i32)[m, k]; rows whose list overflowed (massive
// ties) are redone exactly by a small persistent kernel from the packed operands (scratch: sel_redo_ctas(n) * n floats,
// carved by sel_prepare); kind SQL2 or DOT
// CTAs of the redo kernel: each owns a scratch row of n floats; about 256 MB of scratch in total, 8..296 CTAs
inline int sel_redo_ctas(int n) {
  long long c = (256ll << 20) / ((long long)(n > 0 ? n : 1) * 4);
  return (int)(c < 8 ? 8 : (c > 296 ? 296 : c));
}
int knn_select_finish(const SelArgs& sel, int m, int n, int k, int kind, const void* packedA, const float* xx, const float* isa,
                      const void* packedB, const float* yy, const float* isb, int d, int64_t index_base, float* out_d, void* out_i,
                      int idx64, float* scratch_rows, cudaStream_t st);

// TMA descriptor of a packed operand [rows, 2*dp] fp16: box = {64 k-elements, box_rows}, SWIZZLE_128B, OOB reads as zero.
int make_operand_map(CUtensorMap* tm, const void* packed, int rows, int dp, int box_rows);

// row-wise exact top-k (topk.cu); out_i is i64 when idx64 else i32; ids = cand_i[col] or col + index_base
// chunk_red (optional): the per-32-column reductions written by gemm_split_launch for this matrix (consumed: rows the
// warp-per-row selection hands to the exact kernel are flagged in it).  chunk_red2 != null: `vals` is a triangle-mode
// matrix (gemm_split_tri_launch; square, ascending).
int topk_rows_launch(const float* vals, int rows, int cols, int64_t ldv, const int64_t* cand_i, int k, int descending,
                     int64_t index_base, float* out_d, void* out_i, int idx64, cudaStream_t st,
                     float* chunk_red = nullptr, int64_t ld_chunk_red = 0, const uint32_t* chunk_red2 = nullptr);

}  // namespace reidb200
