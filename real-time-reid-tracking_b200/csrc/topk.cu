// (2b) exact row-wise top-k selection and the brute-force k-NN built on it
// (reid/faiss_utils.py:30-139: faiss bfKnn / GpuIndexFlatL2 / IndexFlatL2 / IndexShards stand-in).
//
// Selection, one CTA per row, no sorting of the row and no shared-memory atomics in the hot loop:
//   pass 1  every thread keeps the minimum key of its strided sub-stream(s): G group minima
//           (G = 512 or 1024 >= k); the ordered keys are cached in shared memory when the row fits.  The k-th smallest group minimum t0 is an upper bound of the k-th
//           smallest element (k distinct elements are <= t0) and is nearly tight because k << G.
//   pass 2  re-read the row (L1/L2 hit) and append every element with key <= t0 to a small candidate
//           list (warp-aggregated append).
//   finish  rank-sort the candidates by (key, id) and emit the first k.
// If the candidate list overflows (heavy ties / adversarial order) the CTA falls back to an exact
// bisection on the 32-bit ordered key.  Ties are broken by the lower id, like a stable argsort.
#include <stdlib.h>
#include <string.h>

#include "gemm.cuh"

namespace reidb200 {

constexpr int TOPK_CAP = 2048;   // candidate capacity (key + column in shared memory)
constexpr int TOPK_MAX_K = 1024;
constexpr int TOPK_SMEM_ROW_MAX = 36 * 1024;  // columns whose keys are cached in shared memory (144 KB)
constexpr uint32_t TOPK_REDO_FLAG = 0x7fc0deadu;  // NaN payload written over a row's first chunk reduction: "redo this row"
constexpr int TW_CAP = 512;                       // candidates per warp in the chunk-guided kernel

__device__ __forceinline__ uint32_t sel_key(float v, int descending) {
  const uint32_t u = float_to_ordered(v);
  return descending ? ~u : u;
}
__device__ __forceinline__ float sel_val(uint32_t key, int descending) {
  uint32_t u = descending ? ~key : key;
  u = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
  return __uint_as_float(u);
}

template <int NT>
__device__ __forceinline__ int block_sum_int(int v, int* red) {
  v = warp_sum_i(v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  int t = 0;
#pragma unroll
  for (int w = 0; w < NT / 32; ++w) t += red[w];
  return t;
}

// warp-aggregated append of (key, col) to the candidate list
__device__ __forceinline__ void cand_append(bool take, uint32_t key, int col, int lane, int* s_count, uint32_t* ckey,
                                            int32_t* ccol) {
  const uint32_t ballot = __ballot_sync(0xffffffffu, take);
  if (ballot) {
    int base = 0;
    if (lane == 0) base = atomicAdd(s_count, __popc(ballot));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (take) {
      const int pos = base + __popc(ballot & ((1u << lane) - 1));
      if (pos < TOPK_CAP) { ckey[pos] = key; ccol[pos] = col; }
    }
  }
}

// MODE 0: two passes over the row in global memory (second pass hits L2).
// MODE 1: the ordered keys of the row are cached in dynamic shared memory by pass 1 (row read from HBM once).
// (The chunk-guided selection over GEMM outputs is topk_chunk_warp_kernel below; rows it could not finish are flagged
//  in their chunk-reduction row and redone here: `flag_rows` != null makes every other CTA exit at once.)
template <int MODE, int NT>
__device__ __forceinline__ void topk_row_body(const int row, const float* __restrict__ vals, int cols, int64_t ldv,
                                              const int64_t* __restrict__ cand_i, int k, int descending, int64_t index_base,
                                              float* __restrict__ out_d, void* __restrict__ out_i, int idx64, bool tri = false) {
  extern __shared__ __align__(16) uint32_t rowkey[];  // MODE 1: keys of the row
  __shared__ uint32_t ckey[TOPK_CAP];
  __shared__ int32_t ccol[TOPK_CAP];
  __shared__ int red[NT / 32];
  __shared__ int s_count;

  const float* v = vals + (size_t)row * ldv;
  const int64_t* ci = cand_i ? cand_i + (size_t)row * cols : nullptr;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int kk = min(k, cols);
  float* od = out_d + (size_t)row * k;
  // `tri` (MODE 0 only): the row lives in a triangle-mode matrix (gemm.cuh tri_stored) -- scalar loads, element c of the
  // row read at its transposed place when its own was not stored
  const bool vec = !tri && (reinterpret_cast<uintptr_t>(v) & 15) == 0;
  const int ncol4 = vec ? (cols >> 2) : 0;
  const float4* v4 = reinterpret_cast<const float4*>(v);
  auto ldat = [&](int c) -> float { return (tri && !tri_stored(row, c)) ? __ldg(vals + (size_t)c * ldv + row) : __ldg(v + c); };
  if (tid == 0) s_count = 0;
  uint32_t t0 = 0xffffffffu;

  {
    // ---- pass 1: group minima (128-bit loads when the row is 16-byte aligned)
    const int groups = (kk <= NT) ? NT : 2 * NT;
    const int per_thread = groups / NT;  // 1 or 2
    uint32_t m[2] = {0xffffffffu, 0xffffffffu};
#pragma unroll 4
    for (int i = tid; i < ncol4; i += NT) {
      const float4 x = __ldg(v4 + i);
      const uint32_t k0 = sel_key(x.x, descending), k1 = sel_key(x.y, descending);
      const uint32_t k2 = sel_key(x.z, descending), k3 = sel_key(x.w, descending);
      if (MODE == 1) *reinterpret_cast<uint4*>(rowkey + 4 * i) = make_uint4(k0, k1, k2, k3);
      const int g = per_thread == 1 ? 0 : ((i / NT) & 1);
      m[g] = min(m[g], min(min(k0, k1), min(k2, k3)));
    }
    for (int c = ncol4 * 4 + tid; c < cols; c += NT) {
      const uint32_t kc = sel_key(ldat(c), descending);
      if (MODE == 1) rowkey[c] = kc;
      const int g = per_thread == 1 ? 0 : ((c / NT) & 1);
      m[g] = min(m[g], kc);
    }
    // k-th smallest group minimum by bisection on the ordered key (32 barrier-counts): an upper bound of the k-th
    // smallest element, nearly tight because k << number of groups.
    if (kk > 0 && kk <= groups) {
      uint32_t lo = 0, hi = 0xffffffffu;
      while (lo < hi) {
        const uint32_t mid = lo + ((hi - lo) >> 1);
        int c = __syncthreads_count(m[0] <= mid);
        if (per_thread == 2) c += __syncthreads_count(m[1] <= mid);
        if (c >= kk) hi = mid; else lo = mid + 1;
      }
      t0 = lo;
    }
    __syncthreads();

    // ---- pass 2: collect candidates <= t0
    if (MODE == 1) {
      const int nq4 = cols >> 2;
      for (int i0 = 0; i0 < nq4; i0 += NT) {
        const int i = i0 + tid;
        uint4 k4 = make_uint4(0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu);
        const bool in = i < nq4;
        if (in) k4 = *reinterpret_cast<const uint4*>(rowkey + 4 * i);
        const bool any = in && min(min(k4.x, k4.y), min(k4.z, k4.w)) <= t0;
        if (__ballot_sync(0xffffffffu, any) == 0) continue;  // warp-uniform: nothing to take in these 128 columns
        cand_append(in && k4.x <= t0, k4.x, 4 * i + 0, lane, &s_count, ckey, ccol);
        cand_append(in && k4.y <= t0, k4.y, 4 * i + 1, lane, &s_count, ckey, ccol);
        cand_append(in && k4.z <= t0, k4.z, 4 * i + 2, lane, &s_count, ckey, ccol);
        cand_append(in && k4.w <= t0, k4.w, 4 * i + 3, lane, &s_count, ckey, ccol);
      }
      for (int c0 = nq4 * 4; c0 < cols; c0 += NT) {
        const int c = c0 + tid;
        const uint32_t key = c < cols ? rowkey[c] : 0xffffffffu;
        cand_append(c < cols && key <= t0, key, c, lane, &s_count, ckey, ccol);
      }
    } else {
      for (int i0 = 0; i0 < ncol4; i0 += NT) {
        const int i = i0 + tid;
        uint32_t key4[4] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu};
        const bool in = i < ncol4;
        if (in) {
          const float4 x = __ldg(v4 + i);
          key4[0] = sel_key(x.x, descending); key4[1] = sel_key(x.y, descending);
          key4[2] = sel_key(x.z, descending); key4[3] = sel_key(x.w, descending);
        }
        const bool any = in && (min(min(key4[0], key4[1]), min(key4[2], key4[3])) <= t0);
        if (__ballot_sync(0xffffffffu, any) == 0) continue;  // warp-uniform
#pragma unroll
        for (int u = 0; u < 4; ++u) cand_append(in && key4[u] <= t0, key4[u], 4 * i + u, lane, &s_count, ckey, ccol);
      }
      for (int c0 = ncol4 * 4; c0 < cols; c0 += NT) {
        const int c = c0 + tid;
        const uint32_t key = c < cols ? sel_key(ldat(c), descending) : 0xffffffffu;
        cand_append(c < cols && key <= t0, key, c, lane, &s_count, ckey, ccol);
      }
    }
  }
  __syncthreads();
  int count = s_count;

  if (count > TOPK_CAP) {
    // ---- fallback (heavy ties / adversarial order): exact k-th key by bisection over the row, ordered collection
    auto key_at = [&](int c) -> uint32_t { return MODE == 1 ? rowkey[c] : sel_key(ldat(c), descending); };
    uint32_t lo = 0, hi = 0xffffffffu;  // invariant: count(key <= hi) >= kk, count(key <= lo-1) < kk
    while (lo < hi) {
      const uint32_t mid = lo + ((hi - lo) >> 1);
      int cnt = 0;
      for (int c = tid; c < cols; c += NT) cnt += key_at(c) <= mid;
      cnt = block_sum_int<NT>(cnt, red);
      if (cnt >= kk) hi = mid; else lo = mid + 1;
    }
    t0 = lo;  // the kk-th smallest key
    __syncthreads();
    if (tid == 0) s_count = 0;
    __syncthreads();
    // strictly smaller keys (fewer than kk of them) in any order
    for (int c0 = 0; c0 < cols; c0 += NT) {
      const int c = c0 + tid;
      const uint32_t key = c < cols ? key_at(c) : 0xffffffffu;
      cand_append(c < cols && key < t0, key, c, lane, &s_count, ckey, ccol);
    }
    __syncthreads();
    int have = s_count;
    // ties == t0 in column order until kk are held (chunk by chunk, ordered prefix over the CTA)
    for (int c0 = 0; c0 < cols && have < kk; c0 += NT) {
      const int c = c0 + tid;
      const bool take = c < cols && key_at(c) == t0;
      const uint32_t ballot = __ballot_sync(0xffffffffu, take);
      __syncthreads();
      if (lane == 0) red[warp] = __popc(ballot);
      __syncthreads();
      int before = 0, total = 0;
      for (int w = 0; w < NT / 32; ++w) { if (w < warp) before += red[w]; total += red[w]; }
      if (take) {
        const int pos = have + before + __popc(ballot & ((1u << lane) - 1));
        if (pos < kk) { ckey[pos] = t0; ccol[pos] = c; }
      }
      have = min(kk, have + total);
    }
    __syncthreads();
    count = kk;
  }

  // ---- finish: rank-sort candidates by (key, id) and emit the first kk
  for (int e = tid; e < count; e += NT) {
    const uint32_t key = ckey[e];
    const int c = ccol[e];
    const int64_t id = ci ? ci[c] : (int64_t)c;
    int rank = 0;
    for (int o = 0; o < count; ++o) {
      const uint32_t ok = ckey[o];
      if (ok < key) { ++rank; continue; }
      if (ok == key && o != e) {
        const int oc = ccol[o];
        const int64_t oid = ci ? ci[oc] : (int64_t)oc;
        rank += (oid < id) || (oid == id && o < e);
      }
    }
    if (rank < kk) {
      od[rank] = sel_val(key, descending);
      const int64_t gid = ci ? id : id + index_base;
      if (idx64) reinterpret_cast<int64_t*>(out_i)[(size_t)row * k + rank] = gid;
      else reinterpret_cast<int32_t*>(out_i)[(size_t)row * k + rank] = (int32_t)gid;
    }
  }
  // fewer columns than k: pad like faiss (distance = +/-inf, id = -1)
  for (int e = kk + tid; e < k; e += NT) {
    od[e] = descending ? -INFINITY : INFINITY;
    if (idx64) reinterpret_cast<int64_t*>(out_i)[(size_t)row * k + e] = -1;
    else reinterpret_cast<int32_t*>(out_i)[(size_t)row * k + e] = -1;
  }
}


template <int MODE, int NT>
__global__ void __launch_bounds__(NT)
topk_rows_kernel(const float* __restrict__ vals, int cols, int64_t ldv, const int64_t* __restrict__ cand_i, int k,
                 int descending, int64_t index_base, float* __restrict__ out_d, void* __restrict__ out_i, int idx64) {
  topk_row_body<MODE, NT>(blockIdx.x, vals, cols, ldv, cand_i, k, descending, index_base, out_d, out_i, idx64);
}

// Redo of the rows the chunk-guided kernel flagged (first chunk reduction overwritten with TOPK_REDO_FLAG): a small
// persistent grid scans the flags; normally no row is flagged and the kernel ends after one pass over them.
template <int NT>
__global__ void __launch_bounds__(NT)
topk_rows_redo_kernel(const float* __restrict__ vals, int rows, int cols, int64_t ldv, int k, int descending,
                      int64_t index_base, float* __restrict__ out_d, void* __restrict__ out_i, int idx64,
                      const float* __restrict__ flag_rows, int64_t ldcr, int tri) {
  // every thread checks one row per pass (a contiguous block of NT rows per CTA); flagged rows are queued in shared memory
  __shared__ int todo[NT];
  __shared__ int ntodo;
  for (int base = blockIdx.x * NT; base < rows; base += gridDim.x * NT) {
    if (threadIdx.x == 0) ntodo = 0;
    __syncthreads();
    const int row = base + threadIdx.x;
    if (row < rows && __float_as_uint(__ldg(flag_rows + (size_t)row * ldcr)) == TOPK_REDO_FLAG) todo[atomicAdd(&ntodo, 1)] = row;
    __syncthreads();
    const int n = ntodo;
    for (int i = 0; i < n; ++i) {
      topk_row_body<0, NT>(todo[i], vals, cols, ldv, nullptr, k, descending, index_base, out_d, out_i, idx64, tri != 0);
      __syncthreads();
    }
  }
}

// Chunk-guided selection, one WARP per row: the producing GEMM already reduced every 32-column chunk of the row (min, or
// max for descending).  Lane-local minima of 32*S interleaved groups of chunk keys give, by a register-only bisection, an
// upper bound t0 of the k-th smallest element (k distinct elements are <= the k-th smallest group minimum); only chunks
// whose reduction reaches t0 (about k..2k of them) are read from the matrix, 128 bytes each, and their elements <= t0
// become candidates (warp-ballot compaction, no atomics, no block barriers).  A rank sort of the candidates by
// (key, column) emits the k best; ties -> lower column.  Rows with more than TW_CAP candidates (massive ties) are
// flagged and redone by the exact CTA-per-row kernel.
// TRI: `vals` is a triangle-mode matrix (gemm.cuh gemm_split_tri_launch; ascending only).  A chunk of the half that was
// not stored (tile block of the chunk > tile block of the row) whose second smallest value cannot pass t0 contributes its
// minimum alone, taken from the two chunk words (value from chunk_red, column from chunk_red2) without touching the
// matrix; the few chunks with two or more passing elements are read at their transposed place (32 sectors).
// PIPE: the loads of the chunk words are issued in batches (two groups of S in the first pass, four 32-chunk groups -- and
// their second words -- in the second) before any of them is used: the kernel was bound by one dependent L2 / DRAM round
// trip per 32 chunks in each pass (ncu: long scoreboard 5.7 warps per issue at 40 resident warps).
template <int S, bool TRI, bool PIPE>
__global__ void __launch_bounds__(128)
topk_chunk_warp_kernel(const float* __restrict__ vals, int rows, int cols, int64_t ldv, int k, int descending,
                       int64_t index_base, float* __restrict__ out_d, void* __restrict__ out_i, int idx64,
                       float* __restrict__ chunk_red, int64_t ldcr, const uint32_t* __restrict__ chunk_red2) {
  __shared__ uint32_t ckey[4][TW_CAP];
  __shared__ int32_t ccol[4][TW_CAP];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * 4 + warp;
  if (row >= rows) return;
  const float* v = vals + (size_t)row * ldv;
  float* cr = chunk_red + (size_t)row * ldcr;
  const uint32_t* cr2 = TRI ? chunk_red2 + (size_t)row * ldcr : nullptr;
  const int rt = row >> 8;  // TRI: chunks c with (c >> 3) > rt belong to the half that was not stored
  const int nchunks = (cols + 31) >> 5;
  const int kk = min(k, cols);

  uint32_t m[S];
#pragma unroll
  for (int s = 0; s < S; ++s) m[s] = 0xffffffffu;
  if (PIPE) {
    for (int c0 = 0; c0 < nchunks; c0 += 64 * S) {
      uint32_t t[2 * S];
#pragma unroll
      for (int j = 0; j < 2 * S; ++j) {
        const int c = c0 + j * 32 + lane;
        t[j] = c < nchunks ? sel_key(__ldg(cr + c), descending) : 0xffffffffu;
      }
#pragma unroll
      for (int j = 0; j < 2 * S; ++j) m[j % S] = min(m[j % S], t[j]);
    }
  } else {
    for (int c0 = 0; c0 < nchunks; c0 += 32 * S) {
#pragma unroll
      for (int s = 0; s < S; ++s) {
        const int c = c0 + s * 32 + lane;
        if (c < nchunks) m[s] = min(m[s], sel_key(__ldg(cr + c), descending));
      }
    }
  }
  uint32_t t0;
  if (PIPE && S == 2) {
    // k <= 32 of 64 group minima: a register bitonic sort (21 exchange steps) instead of 32 bisection rounds of a warp
    // reduction each -- the bisection was 17 % of this kernel's instructions; the k-th smallest key is the same number
    uint32_t ms[S], none[S];
#pragma unroll
    for (int s = 0; s < S; ++s) { ms[s] = m[s]; none[s] = 0; }
    warp_bitonic_sort<S, false>(ms, none, lane);
    const int e = max(kk, 1) - 1;  // sorted element e lives in lane e / S, slot e % S
    const uint32_t pick = (e % S) == 0 ? ms[0] : ms[S - 1];
    t0 = __shfl_sync(0xffffffffu, pick, e / S);
  } else {
    uint32_t lo = 0, hi = 0xffffffffu;
    while (lo < hi) {
      const uint32_t mid = lo + ((hi - lo) >> 1);
      int cnt = 0;
#pragma unroll
      for (int s = 0; s < S; ++s) cnt += m[s] <= mid;
      cnt = __reduce_add_sync(0xffffffffu, cnt);
      if (cnt >= kk) hi = mid; else lo = mid + 1;
    }
    t0 = lo;
  }

  int count = 0;
  uint32_t* mk = ckey[warp];
  int32_t* mc = ccol[warp];
  // The chunks whose minimum passes are first COLLECTED (lane u of `pending` holds the u-th of them), then fetched eight at
  // a time: the chunk values come from DRAM (the matrix was just written and is larger than L2), and one dependent round
  // trip per chunk -- ~25 per row -- was most of this kernel's time.
  int pending = 0, np = 0;
  auto flush = [&]() {
    for (int u0 = 0; u0 < np; u0 += 8) {
      uint32_t key8[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        key8[u] = 0xffffffffu;
        if (u0 + u < np) {  // warp-uniform
          const int ch = __shfl_sync(0xffffffffu, pending, (u0 + u) & 31);
          const int col = ch * 32 + lane;
          if (TRI) {
            // square matrix of at most 32768 rows (tri_mode_ok): element offsets fit 32 bits
            const bool up = (ch >> 3) > rt;
            const uint32_t off = (uint32_t)(up ? col : row) * (uint32_t)ldv + (uint32_t)(up ? row : col);
            if (col < cols) key8[u] = sel_key(__ldg(vals + off), descending);
          } else {
            if (col < cols) key8[u] = sel_key(__ldg(v + col), descending);
          }
        }
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        if (u0 + u >= np) break;  // warp-uniform
        const int ch = __shfl_sync(0xffffffffu, pending, (u0 + u) & 31);
        const int col = ch * 32 + lane;
        const bool take = col < cols && key8[u] <= t0;
        const uint32_t tb = __ballot_sync(0xffffffffu, take);
        const int pos = count + __popc(tb & ((1u << lane) - 1));
        if (take && pos < TW_CAP) { mk[pos] = key8[u]; mc[pos] = col; }
        count += __popc(tb);
      }
    }
    np = 0;
  };
  // one group of 32 chunks: lane's chunk c0 + lane with its minimum's key and (TRI, missing half) its second word
  auto chunk_group = [&](int c0, uint32_t ckeyc, uint32_t w2) {
    const int c = c0 + lane;
    bool q = c < nchunks && ckeyc <= t0;
    if (TRI) {
      // (a lower bound of) the second smallest key does not pass: the minimum is the chunk's only candidate
      const bool single = q && (c >> 3) > rt && (w2 & ~31u) > t0 && c * 32 + (int)(w2 & 31u) < cols;
      const uint32_t sb = __ballot_sync(0xffffffffu, single);
      if (sb) {
        const int pos = count + __popc(sb & ((1u << lane) - 1));
        if (single && pos < TW_CAP) { mk[pos] = ckeyc; mc[pos] = c * 32 + (int)(w2 & 31u); }
        count += __popc(sb);
      }
      q = q && !single;
    }
    uint32_t todo = __ballot_sync(0xffffffffu, q);
    while (todo) {
      const int b = __ffs(todo) - 1;
      todo &= todo - 1;
      if (lane == np) pending = c0 + b;
      if (++np == 32) flush();
    }
  };
  if (PIPE) {
    for (int c0 = 0; c0 < nchunks; c0 += 128) {
      uint32_t kx[4], wx[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int c = c0 + u * 32 + lane;
        kx[u] = c < nchunks ? sel_key(__ldg(cr + c), descending) : 0xffffffffu;
        wx[u] = (TRI && c < nchunks && (c >> 3) > rt) ? __ldg(cr2 + c) : 0u;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (c0 + u * 32 < nchunks) chunk_group(c0 + u * 32, kx[u], wx[u]);  // warp-uniform
    }
  } else {
    for (int c0 = 0; c0 < nchunks; c0 += 32) {
      const int c = c0 + lane;
      const uint32_t ckeyc = c < nchunks ? sel_key(__ldg(cr + c), descending) : 0xffffffffu;
      uint32_t w2 = 0;
      if (TRI && c < nchunks && ckeyc <= t0 && (c >> 3) > rt) w2 = __ldg(cr2 + c);
      chunk_group(c0, ckeyc, w2);
    }
  }
  flush();
  if (count > TW_CAP) {  // hand the row to the exact CTA-per-row kernel
    if (lane == 0) cr[0] = __uint_as_float(TOPK_REDO_FLAG);
    return;
  }
  __syncwarp();
  float* od = out_d + (size_t)row * k;
  for (int e = lane; e < count; e += 32) {
    const uint32_t key = mk[e];
    const int col = mc[e];
    int rank = 0;
    for (int o = 0; o < count; ++o) {
      const uint32_t ok = mk[o];
      rank += (ok < key) || (ok == key && mc[o] < col);
    }
    if (rank < kk) {
      od[rank] = sel_val(key, descending);
      const int64_t gid = (int64_t)col + index_base;
      if (idx64) reinterpret_cast<int64_t*>(out_i)[(size_t)row * k + rank] = gid;
      else reinterpret_cast<int32_t*>(out_i)[(size_t)row * k + rank] = (int32_t)gid;
    }
  }
  for (int e = kk + lane; e < k; e += 32) {
    od[e] = descending ? -INFINITY : INFINITY;
    if (idx64) reinterpret_cast<int64_t*>(out_i)[(size_t)row * k + e] = -1;
    else reinterpret_cast<int32_t*>(out_i)[(size_t)row * k + e] = -1;
  }
}

int topk_rows_launch(const float* vals, int rows, int cols, int64_t ldv, const int64_t* cand_i, int k, int descending,
                     int64_t index_base, float* out_d, void* out_i, int idx64, cudaStream_t st, float* chunk_red,
                     int64_t ld_chunk_red, const uint32_t* chunk_red2) {
  if (rows == 0 || k == 0) return REIDB200_OK;
  RB_REQUIRE(k > 0 && k <= TOPK_MAX_K, "topk: k=%d outside [1,%d]", k, TOPK_MAX_K);
  RB_REQUIRE(cols >= 0 && ldv >= cols, "topk: bad cols/ld");
  const int nchunks = (cols + 31) / 32;
  static DeviceOnce attr;
  if (attr.pending()) {
    RB_CUDA(cudaFuncSetAttribute(topk_rows_kernel<1, 512>, cudaFuncAttributeMaxDynamicSharedMemorySize, TOPK_SMEM_ROW_MAX * 4));
  }
  const bool chunk_guided = chunk_red != nullptr && cand_i == nullptr && nchunks >= 4 * k && nchunks >= 64 && k <= 128;
  RB_REQUIRE(!chunk_red2 || (chunk_guided && rows == cols && rows <= 32768 && !descending),
             "topk: a triangle-mode matrix needs the chunk-guided selection (square, ascending, n/32 >= max(64, 4k), k <= 128)");
  if (chunk_guided) {
    // groups = 32*S >= 2k lane-local minima (S = 2, 4 or 8)
    const int grid = ceil_div(rows, 4);
    const bool tri = chunk_red2 != nullptr;
    {
      ProfScope _ps("topk_rows_kernel", st);
#define RB_TOPK_CHUNK(S, T, P) topk_chunk_warp_kernel<S, T, P><<<grid, 128, 0, st>>>(vals, rows, cols, ldv, k, descending, index_base, out_d, out_i, idx64, chunk_red, ld_chunk_red, chunk_red2)
#define RB_TOPK_CHUNK_S(T, P) do { if (k <= 32) RB_TOPK_CHUNK(2, T, P); else if (k <= 64) RB_TOPK_CHUNK(4, T, P); else RB_TOPK_CHUNK(8, T, P); } while (0)
      const bool pipe = !(getenv("REIDB200_TOPK_PIPE") && atoi(getenv("REIDB200_TOPK_PIPE")) == 0);  // 0: one load round per 32 chunks (A/B)
      if (tri) {
        if (pipe) RB_TOPK_CHUNK_S(true, true); else RB_TOPK_CHUNK_S(true, false);
      } else {
        if (pipe) RB_TOPK_CHUNK_S(false, true); else RB_TOPK_CHUNK_S(false, false);
      }
#undef RB_TOPK_CHUNK_S
#undef RB_TOPK_CHUNK
    }
    RB_LAUNCH_CHECK("topk_chunk_warp_kernel");
    // rows the warp kernel flagged (candidate overflow: massive ties) are redone exactly
    ProfScope _ps2("topk_rows_redo_kernel", st);
    topk_rows_redo_kernel<512><<<min(ceil_div(rows, 512), 2 * num_sms()), 512, 0, st>>>(vals, rows, cols, ldv, k, descending, index_base, out_d,
                                                                       out_i, idx64, chunk_red, ld_chunk_red, tri ? 1 : 0);
  } else if (cols <= TOPK_SMEM_ROW_MAX) {
    ProfScope _ps("topk_rows_kernel", st);
    topk_rows_kernel<1, 512><<<rows, 512, align_up((size_t)cols * 4, 16), st>>>(vals, cols, ldv, cand_i, k, descending,
                                                                                index_base, out_d, out_i, idx64);
  } else {
    ProfScope _ps("topk_rows_kernel", st);
    topk_rows_kernel<0, 512><<<rows, 512, 0, st>>>(vals, cols, ldv, cand_i, k, descending, index_base, out_d, out_i, idx64);
  }
  RB_LAUNCH_CHECK("topk_rows_kernel");
  return REIDB200_OK;
}

// ------------------------------------------------------------------------------------ selection fused into the GEMM
// (SelArgs, gemm.cuh): the GEMM epilogue has left every row an append-only list of (key, column) entries that holds
// every element whose key was <= the row's threshold when it was produced, and a final threshold tau that is an upper
// bound of the row's k-th best key.  One warp per row keeps the entries <= tau (ballot compaction into shared memory),
// rank-sorts them by (key, column) and emits the k best -- ties -> lower column, like a stable argsort.  A row whose
// list overflowed in the epilogue, or whose candidates do not fit (thousands of exact ties), is marked and redone
// exactly by sel_redo_kernel.
constexpr int SL_WARPS = 4;

__global__ void __launch_bounds__(SL_WARPS * 32)
select_lists_kernel(SelArgs s, int rows, int cols, int k, int descending, int64_t index_base, float* __restrict__ out_d,
                    void* __restrict__ out_i, int idx64) {
  __shared__ uint32_t ckey[SL_WARPS][TW_CAP];
  __shared__ int32_t ccol[SL_WARPS][TW_CAP];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * SL_WARPS + warp;
  if (row >= rows) return;
  const uint32_t resv = s.resv[row];
  if (resv > (uint32_t)s.cap) return;  // overflowed in the epilogue: sel_redo_kernel
  const uint32_t tau = s.tau[row];
  const unsigned long long* e = s.list + (size_t)row * s.cap;
  const int kk = min(k, cols);
  uint32_t* mk = ckey[warp];
  int32_t* mc = ccol[warp];
  int count = 0;
  for (uint32_t i0 = 0; i0 < resv; i0 += 32) {
    const uint32_t i = i0 + lane;
    unsigned long long v = ~0ull;
    if (i < resv) v = __ldcg(e + i);
    const uint32_t key = (uint32_t)(v >> 32);
    const bool take = i < resv && key <= tau;
    const uint32_t tb = __ballot_sync(0xffffffffu, take);
    const int pos = count + __popc(tb & ((1u << lane) - 1));
    if (take && pos < TW_CAP) { mk[pos] = key; mc[pos] = (int32_t)(uint32_t)v; }
    count += __popc(tb);
  }
  if (count > TW_CAP) {
    // the threshold is looser than the buffer: find the exact kk-th smallest key of the list by bisection (re-reading the
    // list, which sits in L1 / L2) and collect again; only thousands of EXACT ties are left to the exact kernel
    uint32_t lo = 0, hi = tau;
    while (lo < hi) {
      const uint32_t mid = lo + ((hi - lo) >> 1);
      int cnt = 0;
      for (uint32_t i = lane; i < resv; i += 32) cnt += (uint32_t)(__ldcg(e + i) >> 32) <= mid;
      cnt = __reduce_add_sync(0xffffffffu, cnt);
      if (cnt >= kk) hi = mid; else lo = mid + 1;
    }
    count = 0;
    for (uint32_t i0 = 0; i0 < resv; i0 += 32) {
      const uint32_t i = i0 + lane;
      unsigned long long v = ~0ull;
      if (i < resv) v = __ldcg(e + i);
      const uint32_t key = (uint32_t)(v >> 32);
      const bool take = i < resv && key <= lo;
      const uint32_t tb = __ballot_sync(0xffffffffu, take);
      const int pos = count + __popc(tb & ((1u << lane) - 1));
      if (take && pos < TW_CAP) { mk[pos] = key; mc[pos] = (int32_t)(uint32_t)v; }
      count += __popc(tb);
    }
    if (count > TW_CAP) {  // (uniform) massive ties: hand the row to the exact kernel
      if (lane == 0) s.resv[row] = (uint32_t)s.cap + 1u;
      return;
    }
  }
  __syncwarp();
  float* od = out_d + (size_t)row * k;
  for (int x = lane; x < count; x += 32) {
    const uint32_t key = mk[x];
    const int col = mc[x];
    int rank = 0;
    for (int o = 0; o < count; ++o) {
      const uint32_t ok = mk[o];
      rank += (ok < key) || (ok == key && mc[o] < col);
    }
    if (rank < kk) {
      od[rank] = sel_val(key, descending);
      const int64_t gid = (int64_t)col + index_base;
      if (idx64) reinterpret_cast<int64_t*>(out_i)[(size_t)row * k + rank] = gid;
      else reinterpret_cast<int32_t*>(out_i)[(size_t)row * k + rank] = (int32_t)gid;
    }
  }
  for (int x = max(kk, count) + lane; x < k; x += 32) {  // fewer columns than k: pad like faiss (distance = +/-inf, id = -1)
    od[x] = descending ? -INFINITY : INFINITY;
    if (idx64) reinterpret_cast<int64_t*>(out_i)[(size_t)row * k + x] = -1;
    else reinterpret_cast<int32_t*>(out_i)[(size_t)row * k + x] = -1;
  }
}

// Exact redo of the (rare) rows marked resv > cap: a few persistent CTAs scan the marks; a marked row's whole distance row
// is recomputed from the packed operands -- x = (hi + lo) * 2^-e, fp32 FMA dot products, the GEMM's epilogue formula -- into
// this CTA's scratch row, and the exact CTA-per-row selection runs on it.
template <int NT>
__global__ void __launch_bounds__(NT)
sel_redo_kernel(SelArgs s, int rows, int cols, int k, int kind, const __half* __restrict__ pa, const float* __restrict__ xx,
                const float* __restrict__ isa, const __half* __restrict__ pb, const float* __restrict__ yy,
                const float* __restrict__ isb, int dp, int64_t index_base, float* __restrict__ out_d, void* __restrict__ out_i,
                int idx64, float* __restrict__ scratch) {
  __shared__ int todo[NT];
  __shared__ int ntodo;
  extern __shared__ __align__(16) float arow[];  // dp floats: the row's scaled features (hi + lo)
  float* my = scratch + (size_t)blockIdx.x * cols;
  const int desc = kind == REIDB200_DIST_DOT;
  for (int base = blockIdx.x * NT; base < rows; base += gridDim.x * NT) {
    if (threadIdx.x == 0) ntodo = 0;
    __syncthreads();
    const int r = base + threadIdx.x;
    if (r < rows && s.resv[r] > (uint32_t)s.cap) todo[atomicAdd(&ntodo, 1)] = r;
    __syncthreads();
    const int n = ntodo;
    for (int t = 0; t < n; ++t) {
      const int row = todo[t];
      const __half* a = pa + (size_t)row * 2 * dp;
      for (int x = threadIdx.x; x < dp; x += NT) arow[x] = __half2float(a[x]) + __half2float(a[dp + x]);
      __syncthreads();
      const float ia = isa[row], xr = xx ? xx[row] : 0.f;
      for (int c = threadIdx.x; c < cols; c += NT) {
        const __half2* b = reinterpret_cast<const __half2*>(pb + (size_t)c * 2 * dp);
        float acc = 0.f;
        for (int x = 0; x < dp / 2; ++x) {
          const float2 h = __half22float2(b[x]), l = __half22float2(b[dp / 2 + x]);
          acc = fmaf(arow[2 * x], h.x + l.x, acc);
          acc = fmaf(arow[2 * x + 1], h.y + l.y, acc);
        }
        my[c] = dist_epilogue(kind, __fmul_rn(__fmul_rn(acc, ia), isb[c]), xr, yy ? yy[c] : 0.f);
      }
      __syncthreads();
      topk_row_body<0, NT>(row, my, cols, 0, nullptr, k, desc, index_base, out_d, out_i, idx64);
      __syncthreads();
    }
  }
}

size_t sel_workspace_bytes(int m, int n, int k) {
  CarveSize c;
  c.take<uint32_t>(2 * (size_t)m + (size_t)m / 128 + 2);
  c.take<unsigned long long>((size_t)m * sel_sketch(k));
  c.take<unsigned long long>((size_t)m * sel_cap(k));
  c.take<float>((size_t)sel_redo_ctas(n) * n);
  return align_up(c.off, 256);
}

int sel_prepare(void* ws, size_t ws_bytes, int m, int n, int k, int descending, SelArgs* out, float** scratch_rows, cudaStream_t st) {
  RB_REQUIRE(sel_cap(k) > 0, "select: k=%d outside the fused selection's range", k);
  Carver c(ws, ws_bytes);
  const size_t n_gate = (size_t)m / 128 + 2;
  uint32_t* st2 = c.take<uint32_t>(2 * (size_t)m + n_gate);
  unsigned long long* sketch = c.take<unsigned long long>((size_t)m * sel_sketch(k));
  unsigned long long* list = c.take<unsigned long long>((size_t)m * sel_cap(k));
  float* scratch = c.take<float>((size_t)sel_redo_ctas(n) * n);
  RB_REQUIRE(c.ok(), "select: workspace too small");
  RB_CUDA(cudaMemsetAsync(st2, 0xff, (size_t)m * sizeof(uint32_t), st));                       // tau = +inf
  RB_CUDA(cudaMemsetAsync(st2 + m, 0, ((size_t)m + n_gate) * sizeof(uint32_t), st));           // resv = 0, gates closed
  RB_CUDA(cudaMemsetAsync(sketch, 0xff, (size_t)m * sel_sketch(k) * sizeof(unsigned long long), st));  // empty slots
  memset(out, 0, sizeof(*out));
  out->tau = st2; out->resv = st2 + m; out->gate = st2 + 2 * (size_t)m; out->list = list; out->sketch = sketch;
  out->cap = sel_cap(k); out->k = k; out->W = sel_sketch(k); out->window = 32; out->descending = descending;
  out->debug = getenv("REIDB200_SEL_DEBUG") ? atoi(getenv("REIDB200_SEL_DEBUG")) : 0;
  *scratch_rows = scratch;
  return REIDB200_OK;
}

int knn_select_finish(const SelArgs& sel, int m, int n, int k, int kind, const void* packedA, const float* xx, const float* isa,
                      const void* packedB, const float* yy, const float* isb, int d, int64_t index_base, float* out_d, void* out_i,
                      int idx64, float* scratch_rows, cudaStream_t st) {
  const int desc = kind == REIDB200_DIST_DOT;
  if (getenv("REIDB200_SEL_STATS")) {  // diagnostics only (synchronises): how long the candidate lists grew
    cudaStreamSynchronize(st);
    uint32_t* h = (uint32_t*)malloc((size_t)m * 4);
    cudaMemcpy(h, sel.resv, (size_t)m * 4, cudaMemcpyDeviceToHost);
    unsigned long long sum = 0; uint32_t mx = 0; int over = 0, big = 0;
    for (int i = 0; i < m; ++i) { sum += h[i]; mx = h[i] > mx ? h[i] : mx; over += h[i] > (uint32_t)sel.cap; big += h[i] > 512; }
    fprintf(stderr, "[sel] m=%d n=%d k=%d cap=%d: list length mean %.1f max %u, > 512: %d, overflow: %d\n", m, n, k, sel.cap,
            (double)sum / m, mx, big, over);
    free(h);
  }
  { ProfScope _ps("select_lists_kernel", st); select_lists_kernel<<<ceil_div(m, SL_WARPS), SL_WARPS * 32, 0, st>>>(sel, m, n, k, desc, index_base, out_d, out_i, idx64); }
  RB_LAUNCH_CHECK("select_lists_kernel");
  const int dp = padded_dim(d);
  static DeviceOnce attr;
  if (attr.pending()) RB_CUDA(cudaFuncSetAttribute(sel_redo_kernel<512>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
  RB_REQUIRE((size_t)dp * 4 <= 64 * 1024, "select: feature dimension %d too large for the redo kernel", d);
  { ProfScope _ps("sel_redo_kernel", st); sel_redo_kernel<512><<<sel_redo_ctas(n), 512, (size_t)dp * 4, st>>>(
      sel, m, n, k, kind, static_cast<const __half*>(packedA), xx, isa, static_cast<const __half*>(packedB), yy, isb, dp, index_base,
      out_d, out_i, idx64, scratch_rows); }
  RB_LAUNCH_CHECK("sel_redo_kernel");
  return REIDB200_OK;
}

// gallery columns per GEMM + select pass: distance block of at most 2^31 floats (8 GiB of the 180 GB)
// (REIDB200_KNN_BLOCK_FLOATS overrides the block size: lets the tests drive the multi-pass merge at small sizes)
// Measured at 10 k x 4 M x 2048, top-100 (profiles/r02_retrieval_blocks.log): the row select costs a fixed ~0.45 ms per
// pass over 10 k rows, the GEMM loses DRAM page locality as the rows of its output block move apart --
//   2 GiB: GEMM 373.8 + select 33.9 = 411.3 ms   8 GiB: 389.4 + 14.4 = 404.8   16 GiB: 395.0 + 11.9 = 407.5   32 GiB: 408.4
static int knn_chunk_cols(int nq, int nb) {
  long long block = 1LL << 31;
  if (const char* e = getenv("REIDB200_KNN_BLOCK_FLOATS")) { const long long v = atoll(e); if (v >= 256) block = v; }
  long long c = block / (nq > 0 ? nq : 1);
  c = c / 256 * 256;
  if (c < 256) c = 256;
  if (c > nb) c = nb;
  return (int)c;
}

}  // namespace reidb200

using namespace reidb200;

extern "C" int reidb200_topk_rows(const float* vals, int rows, int cols, int64_t ldv, const int64_t* cand_i, int k,
                                  int descending, int64_t index_base, float* out_d, void* out_i, int idx64,
                                  void* stream) {
  if (rows == 0) return REIDB200_OK;
  RB_REQUIRE(vals && out_d && out_i && rows > 0, "topk_rows: bad argument");
  return topk_rows_launch(vals, rows, cols, ldv, cand_i, k, descending, index_base, out_d, out_i, idx64,
                          (cudaStream_t)stream);
}

// Shared body of knn / knn_packed: operands already packed; `c` carves the distance block, the chunk reductions and the
// per-chunk candidate tables out of what is left of the workspace.
// the selection runs inside the GEMM epilogue (no distance block) for k <= 128; REIDB200_KNN_MATERIALISE keeps the
// distance-block path (tests / experiments)
// Measured (profiles/r02_selection_experiments.md): with the slow path switched off the selecting GEMM is as fast as the
// storing one (97 vs 94 ms per 10 k x 1 M x 2048 search) and the 9 ms row-select disappears, but every accepted element
// (~1450 per row at k = 100) costs ~3 us of a lone epilogue lane's time -- 141 ms in all -- so the distance-block path
// stays the default for k-NN; the selection is what the re-ranking step uses when N x N floats do not fit
// (rerank.cu rerank_select_mode).  REIDB200_KNN_SELECT=1 selects it here (tests run both).
static bool knn_fused(int k, int nb) {
  if (sel_cap(k) == 0) return false;
  if (const char* e = getenv("REIDB200_KNN_SELECT")) return atoi(e) != 0;
  (void)nb;
  return false;
}

static void knn_scratch_size(CarveSize& c, int nq, int nb, int k) {
  if (knn_fused(k, nb)) { c.take<char>(sel_workspace_bytes(nq, nb, k)); return; }
  const int chunk = knn_chunk_cols(nq, nb);
  const int nchunks = ceil_div(nb, chunk);
  c.take<float>((size_t)nq * align_up((size_t)chunk, 4));
  c.take<float>((size_t)nq * ceil_div(chunk, 32));
  if (nchunks > 1) {
    c.take<float>((size_t)nq * nchunks * k);
    c.take<int64_t>((size_t)nq * nchunks * k);
  }
}

static int knn_core(const char* pq, const float* qq, const float* isq, int nq, const char* pb, const float* bb,
                    const float* isb, int nb, int d, int k, int metric, int64_t index_base, float* out_d, void* out_i, int idx64, Carver& c, cudaStream_t st) {
  if (knn_fused(k, nb)) {
    // one GEMM over the whole gallery, candidates filtered against per-row thresholds in its epilogue, then the k best of
    // every row's list: no distance block, no per-block tables, no merge pass
    const size_t bytes = sel_workspace_bytes(nq, nb, k);
    char* sws = c.take<char>(bytes);
    RB_REQUIRE(c.ok(), "knn: workspace too small");
    SelArgs sel;
    float* redo_scratch = nullptr;
    int rc = sel_prepare(sws, bytes, nq, nb, k, metric == REIDB200_DIST_DOT, &sel, &redo_scratch, st);
    if (rc) return rc;
    rc = gemm_split_select_launch(pq, qq, isq, nq, pb, bb, isb, nb, d, metric, nullptr, 0, sel, st);
    if (rc) return rc;
    return knn_select_finish(sel, nq, nb, k, metric, pq, qq, isq, pb, bb, isb, d, index_base, out_d, out_i, idx64, redo_scratch, st);
  }
  const int chunk = knn_chunk_cols(nq, nb);
  const int nchunks = ceil_div(nb, chunk);
  const int dp = padded_dim(d);
  const int desc = metric == REIDB200_DIST_DOT;
  const int64_t ldd = (int64_t)align_up((size_t)chunk, 4);  // 16-byte aligned rows: vector epilogue stores
  float* dist = c.take<float>((size_t)nq * ldd);
  const int ldcr = ceil_div(chunk, 32);
  float* cred = c.take<float>((size_t)nq * ldcr);
  float* cd = nullptr;
  int64_t* cid = nullptr;
  if (nchunks > 1) {
    cd = c.take<float>((size_t)nq * nchunks * k);
    cid = c.take<int64_t>((size_t)nq * nchunks * k);
  }
  RB_REQUIRE(c.ok(), "knn: workspace too small");
  for (int ch = 0; ch < nchunks; ++ch) {
    const int c0 = ch * chunk;
    const int nc = min(chunk, nb - c0);
    int rc = gemm_split_launch(pq, qq, isq, nq, pb + (size_t)c0 * 2 * dp * sizeof(__half), bb + c0, isb + c0, nc, d, metric,
                               dist, ldd, st, cred, ldcr, desc);
    if (rc) return rc;
    if (nchunks == 1)
      return topk_rows_launch(dist, nq, nc, ldd, nullptr, k, desc, index_base, out_d, out_i, idx64, st, cred, ldcr);
    // per-chunk partial result -> table ch of [nchunks][nq][k]
    rc = topk_rows_launch(dist, nq, nc, ldd, nullptr, k, desc, index_base + c0, cd + (size_t)ch * nq * k,
                          cid + (size_t)ch * nq * k, 1, st, cred, ldcr);
    if (rc) return rc;
  }
  // merge: gather the per-chunk tables per query through transposing copies (nchunks*k candidates per query) and run
  // one more selection with explicit ids.
  float* md = dist;  // the distance block is free again: reuse as [nq, nchunks*k]
  int64_t* mi = reinterpret_cast<int64_t*>(dist + (size_t)nq * nchunks * k + 2);
  mi = reinterpret_cast<int64_t*>((reinterpret_cast<uintptr_t>(mi) + 7) & ~(uintptr_t)7);
  RB_REQUIRE((size_t)nq * nchunks * k * 12 + 16 <= (size_t)nq * ldd * 4, "knn: merge buffer does not fit (k too large)");
  for (int ch = 0; ch < nchunks; ++ch) {
    RB_CUDA(cudaMemcpy2DAsync(md + (size_t)ch * k, (size_t)nchunks * k * 4, cd + (size_t)ch * nq * k, (size_t)k * 4,
                              (size_t)k * 4, nq, cudaMemcpyDeviceToDevice, st));
    RB_CUDA(cudaMemcpy2DAsync(mi + (size_t)ch * k, (size_t)nchunks * k * 8, cid + (size_t)ch * nq * k, (size_t)k * 8,
                              (size_t)k * 8, nq, cudaMemcpyDeviceToDevice, st));
  }
  return topk_rows_launch(md, nq, nchunks * k, (int64_t)nchunks * k, mi, k, desc, 0, out_d, out_i, idx64, st);
}

extern "C" size_t reidb200_knn_workspace_bytes(int nq, int nb, int d, int k) {
  if (nq <= 0 || nb <= 0 || d <= 0 || k <= 0) return 0;
  CarveSize c;
  c.take<char>(packed_bytes(nq, d));
  c.take<float>(nq);
  c.take<float>(nq);
  c.take<char>(packed_bytes(nb, d));
  c.take<float>(nb);
  c.take<float>(nb);
  knn_scratch_size(c, nq, nb, k);
  return align_up(c.off, 256);
}

extern "C" int reidb200_knn(const float* xq, int nq, int64_t ldq, const float* xb, int nb, int64_t ldb, int d, int k,
                            int metric, int64_t index_base, float* out_d, void* out_i, int idx64, void* workspace,
                            size_t workspace_bytes, void* stream) {
  if (nq == 0) return REIDB200_OK;
  RB_REQUIRE(xq && xb && out_d && out_i && nq > 0 && nb > 0 && d > 0, "knn: bad argument nq=%d nb=%d d=%d", nq, nb, d);
  RB_REQUIRE(k > 0 && k <= TOPK_MAX_K, "knn: k=%d outside [1,%d]", k, TOPK_MAX_K);
  RB_REQUIRE(metric == REIDB200_DIST_SQL2 || metric == REIDB200_DIST_DOT, "knn: metric must be SQL2 or DOT");
  RB_REQUIRE(workspace && workspace_bytes >= reidb200_knn_workspace_bytes(nq, nb, d, k), "knn: workspace too small");
  RB_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, "knn: workspace must be 256-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  Carver c(workspace, workspace_bytes);
  char* pq = c.take<char>(packed_bytes(nq, d));
  float* qq = c.take<float>(nq);
  float* isq = c.take<float>(nq);
  char* pb = c.take<char>(packed_bytes(nb, d));
  float* bb = c.take<float>(nb);
  float* isb = c.take<float>(nb);
  int rc = pack_rows_launch(xq, nq, ldq, d, pq, qq, isq, st);
  if (rc) return rc;
  if (xq == xb && nq == nb && ldq == ldb) { pb = pq; bb = qq; isb = isq; }
  else { rc = pack_rows_launch(xb, nb, ldb, d, pb, bb, isb, st); if (rc) return rc; }
  return knn_core(pq, qq, isq, nq, pb, bb, isb, nb, d, k, metric, index_base, out_d, out_i, idx64, c, st);
}

extern "C" size_t reidb200_knn_packed_workspace_bytes(int nq, int nb, int d, int k) {
  if (nq <= 0 || nb <= 0 || d <= 0 || k <= 0) return 0;
  CarveSize c;
  c.take<char>(packed_bytes(nq, d));
  c.take<float>(nq);
  c.take<float>(nq);
  knn_scratch_size(c, nq, nb, k);
  return align_up(c.off, 256);
}

extern "C" int reidb200_knn_packed(const float* xq, int nq, int64_t ldq, const void* packed_b, const float* sqnorm_b,
                                   const float* iscale_b, int nb, int d, int k, int metric, int64_t index_base, float* out_d, void* out_i,
                                   int idx64, void* workspace, size_t workspace_bytes, void* stream) {
  if (nq == 0) return REIDB200_OK;
  RB_REQUIRE(xq && packed_b && sqnorm_b && iscale_b && out_d && out_i && nq > 0 && nb > 0 && d > 0,
             "knn_packed: bad argument nq=%d nb=%d d=%d", nq, nb, d);
  RB_REQUIRE(k > 0 && k <= TOPK_MAX_K, "knn_packed: k=%d outside [1,%d]", k, TOPK_MAX_K);
  RB_REQUIRE(metric == REIDB200_DIST_SQL2 || metric == REIDB200_DIST_DOT, "knn_packed: metric must be SQL2 or DOT");
  RB_REQUIRE(workspace && workspace_bytes >= reidb200_knn_packed_workspace_bytes(nq, nb, d, k),
             "knn_packed: workspace too small");
  RB_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255) == 0 && (reinterpret_cast<uintptr_t>(packed_b) & 127) == 0,
             "knn_packed: workspace must be 256-byte and the packed gallery 128-byte aligned");
  RB_REQUIRE(((reinterpret_cast<uintptr_t>(sqnorm_b) | reinterpret_cast<uintptr_t>(iscale_b)) & 15) == 0,
             "knn_packed: sqnorm_b / iscale_b must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  Carver c(workspace, workspace_bytes);
  char* pq = c.take<char>(packed_bytes(nq, d));
  float* qq = c.take<float>(nq);
  float* isq = c.take<float>(nq);
  int rc = pack_rows_launch(xq, nq, ldq, d, pq, qq, isq, st);
  if (rc) return rc;
  return knn_core(pq, qq, isq, nq, static_cast<const char*>(packed_b), sqnorm_b, iscale_b, nb, d, k, metric, index_base,
                  out_d, out_i, idx64, c, st);
}
