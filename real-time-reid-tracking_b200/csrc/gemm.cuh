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

// TMA descriptor of a packed operand [rows, 2*dp] fp16: box = {64 k-elements, box_rows}, SWIZZLE_128B, OOB reads as zero.
int make_operand_map(CUtensorMap* tm, const void* packed, int rows, int dp, int box_rows);

// row-wise exact top-k (topk.cu); out_i is i64 when idx64 else i32; ids = cand_i[col] or col + index_base
// chunk_red (optional): the per-32-column reductions written by gemm_split_launch for this matrix (consumed: rows the
// warp-per-row selection hands to the exact kernel are flagged in it).
int topk_rows_launch(const float* vals, int rows, int cols, int64_t ldv, const int64_t* cand_i, int k, int descending,
                     int64_t index_base, float* out_d, void* out_i, int idx64, cudaStream_t st,
                     float* chunk_red = nullptr, int64_t ld_chunk_red = 0);

}  // namespace reidb200
