// (3) k-reciprocal re-ranking: reid/faiss_utils.py:142-244 (k_reciprocal_neigh,
// compute_jaccard_distance) restated with a SPARSE V (the reference's dense N x N V, V_qe and
// inverted-index scans become ELL rows of at most k1 + k1*(k1/2+1) and k2 times that many entries).
//
//   D2  initial_rank = top-k1 of the squared-L2 matrix (tensor-core GEMM + row select)
//   D3  R(i,k1), R(i,k1/2)                                    krecip_kernel       (warp / row)
//   D4  local expansion (2/3 rule), unique, softmax weights   expand_v_kernel     (CTA / row)
//   D5  query expansion: mean of the k2 nearest rows          qexpand_kernel      (CTA / row)
//   D6  inverted index (CSR by column)                        inv_* kernels
//   D7  Jaccard: min-sum over shared columns, 1 - t/(2-t)     jaccard_rows_kernel (CTA / output row,
//       dense accumulator staged in shared memory, coalesced dense row written once)
//   D8  optional lambda blend with the Euclidean distance (north_star; not in the reference)
#include <stdlib.h>

#include <algorithm>
#include <mutex>
#include <vector>
#include "evalmap.cuh"
#include "gemm.cuh"

namespace reidb200 {

// ------------------------------------------------------------------------------------ D3
// out_idx[i, 0..cnt) = forward neighbours c = rank[i,f], f < kc, with i in rank[c, 0..kc)
// (kc = min(k+1, k1_cols): the reference slices [:k+1] of a k1-column table).
__global__ void __launch_bounds__(256)
krecip_kernel(const int32_t* __restrict__ rank, int N, int k1_cols, int kc, int out_w, int32_t* __restrict__ out_idx,
              int32_t* __restrict__ out_cnt) {
  const int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (i >= N) return;
  int cnt = 0;
  for (int f0 = 0; f0 < kc; f0 += 32) {
    const int f = f0 + lane;
    bool keep = false;
    int c = -1;
    if (f < kc) {
      c = rank[(size_t)i * k1_cols + f];
      if (c >= 0 && c < N) {
        const int32_t* b = rank + (size_t)c * k1_cols;
        for (int t = 0; t < kc; ++t) keep |= (b[t] == i);
      }
    }
    const uint32_t ballot = __ballot_sync(0xffffffffu, keep);
    if (keep) out_idx[(size_t)i * out_w + cnt + __popc(ballot & ((1u << lane) - 1))] = c;
    cnt += __popc(ballot);
  }
  for (int t = cnt + lane; t < out_w; t += 32) out_idx[(size_t)i * out_w + t] = -1;
  if (lane == 0) out_cnt[i] = cnt;
}

// R(i,k1) and R(i,k1/2) in one pass over the neighbour rows (kh = min(k1/2 + 1, k1_cols) <= kc): the half list keeps
// the forward neighbours f < kh whose first kh backward neighbours contain i.
__global__ void __launch_bounds__(256)
krecip_both_kernel(const int32_t* __restrict__ rank, int N, int k1_cols, int kc, int kh, int32_t* __restrict__ full_idx,
                   int32_t* __restrict__ full_cnt, int32_t* __restrict__ half_idx, int32_t* __restrict__ half_cnt) {
  const int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (i >= N) return;
  const bool vec4 = (k1_cols & 3) == 0 && (kc & 3) == 0 && (reinterpret_cast<uintptr_t>(rank) & 15) == 0;
  int cnt = 0, cnth = 0;
  for (int f0 = 0; f0 < kc; f0 += 32) {
    const int f = f0 + lane;
    bool keep = false, keeph = false;
    int c = -1;
    if (f < kc) {
      c = rank[(size_t)i * k1_cols + f];
      if (c >= 0 && c < N) {
        const int32_t* b = rank + (size_t)c * k1_cols;
        if (vec4) {  // rows of whole 16-byte pieces: a quarter of the (uncoalesced: one row per lane) load instructions
          for (int t = 0; t < kc; t += 4) {
            const int4 v = __ldg(reinterpret_cast<const int4*>(b + t));
            const bool m0 = v.x == i, m1 = v.y == i, m2 = v.z == i, m3 = v.w == i;
            keep |= m0 | m1 | m2 | m3;
            keeph |= (m0 && t < kh) | (m1 && t + 1 < kh) | (m2 && t + 2 < kh) | (m3 && t + 3 < kh);
          }
        } else {
          for (int t = 0; t < kc; ++t) {
            const bool m = b[t] == i;
            keep |= m;
            keeph |= m && t < kh;
          }
        }
        keeph = keeph && f < kh;
      }
    }
    const uint32_t ballot = __ballot_sync(0xffffffffu, keep), balloth = __ballot_sync(0xffffffffu, keeph);
    const uint32_t below = (1u << lane) - 1;
    if (keep) full_idx[(size_t)i * kc + cnt + __popc(ballot & below)] = c;
    if (keeph) half_idx[(size_t)i * kh + cnth + __popc(balloth & below)] = c;
    cnt += __popc(ballot);
    cnth += __popc(balloth);
  }
  for (int t = cnt + lane; t < kc; t += 32) full_idx[(size_t)i * kc + t] = -1;
  for (int t = cnth + lane; t < kh; t += 32) half_idx[(size_t)i * kh + t] = -1;
  if (lane == 0) { full_cnt[i] = cnt; half_cnt[i] = cnth; }
}

// ------------------------------------------------------------------------------------ D4: where x_i . x_c comes from
// The reference computes dist = 2 - 2 x_i . X_e^T with one small torch.mm per row (faiss_utils.py:201-204).  Here the
// squared-L2 value of a pair comes from one of two places:
//   * a materialised squared-L2 matrix (reidb200_jaccard_from_rank: the caller owns one), or
//   * when the k-NN was selected inside the GEMM epilogue (no N x N matrix exists): the row's own top-k1 list if the
//     candidate is in it (the GEMM's value, bit for bit), else a direct fp32 dot product of the two feature rows
//     (warp-cooperative, rows are L2 resident) put through the GEMM's epilogue formula.  Both evaluation paths (fused /
//     dense output) run the same kernels, so their results stay bit-identical.
struct PairSrc {
  const float* sql2; int64_t lds;           // materialised matrix, or null
  int tri;                                  // the matrix is in triangle mode (gemm.cuh tri_stored): (i, c) may live at (c, i)
  const float* feats; int64_t ldf; int d;   // fp32 feature rows (used when sql2 is null)
  const int32_t* rank_i; const float* rank_d; int k1c;  // the rows' top-k1 lists
};

// one thread: squared L2 of (i, c) (the CTA fallback kernel for k1 > 32; the warp kernel has its own cooperative version)
__device__ float pair_sql2_thread(const PairSrc& P, const float* __restrict__ sqnorm, int i, int c) {
  if (P.sql2) return (P.tri && !tri_stored(i, c)) ? P.sql2[(size_t)c * P.lds + i] : P.sql2[(size_t)i * P.lds + c];
  for (int t = 0; t < P.k1c; ++t)
    if (P.rank_i[(size_t)i * P.k1c + t] == c) return P.rank_d[(size_t)i * P.k1c + t];
  const float* a = P.feats + (size_t)i * P.ldf;
  const float* b = P.feats + (size_t)c * P.ldf;
  float acc = 0.f;
  for (int k = 0; k < P.d; ++k) acc = fmaf(a[k], b[k], acc);
  return fmaf(-2.0f, acc, __fadd_rn(sqnorm[i], sqnorm[c]));
}

// one warp: x_i . x_c, every lane returns the sum (lane l owns elements 4l..4l+3, +128, ...)
__device__ __forceinline__ float pair_dot_warp(const float* __restrict__ a, const float* __restrict__ b, int d, int lane, bool vec) {
  float acc = 0.f;
  if (vec) {
    for (int k = lane * 4; k < d; k += 128) {
      const float4 x = __ldg(reinterpret_cast<const float4*>(a + k)), y = __ldg(reinterpret_cast<const float4*>(b + k));
      acc = fmaf(x.x, y.x, acc); acc = fmaf(x.y, y.y, acc); acc = fmaf(x.z, y.z, acc); acc = fmaf(x.w, y.w, acc);
    }
  } else {
    for (int k0 = lane * 4; k0 < d; k0 += 128)
      for (int e = 0; e < 4; ++e)
        if (k0 + e < d) acc = fmaf(__ldg(a + k0 + e), __ldg(b + k0 + e), acc);
  }
  return warp_sum(acc);
}

// ------------------------------------------------------------------------------------ D4
constexpr int EXP_THREADS = 128;

// exclusive scan of flags[0..n) in shared memory by one CTA; returns the total.  n <= 8192.
__device__ int block_exclusive_scan(int* data, int n, int* warp_tot) {
  const int tid = threadIdx.x, nthr = blockDim.x, lane = tid & 31, warp = tid >> 5;
  const int per = (n + nthr - 1) / nthr;
  const int b = tid * per, e = min(n, b + per);
  int s = 0;
  for (int t = b; t < e; ++t) s += data[t];
  int incl = s;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += v;
  }
  if (lane == 31) warp_tot[warp] = incl;
  __syncthreads();
  int before = 0, total = 0;
  for (int w = 0; w < nthr / 32; ++w) { if (w < warp) before += warp_tot[w]; total += warp_tot[w]; }
  int run = before + incl - s;
  for (int t = b; t < e; ++t) { const int v = data[t]; data[t] = run; run += v; }
  __syncthreads();
  return total;
}

// One CTA per row i.  Dynamic smem: K[P] flag[P] (ints) + val[P] (floats), P = cap rounded up to a power of two.
// The candidate lists are staged uncompacted in K -- slot t < k1 holds R(i,k1)[t], slot k1 + t*wh + a holds
// R(c_t,k1/2)[a] -- with INT_MAX in every unused slot; rejected candidates (2/3 rule) are overwritten with
// INT_MAX, one bitonic sort of the next power of two above the used extent puts the keys in np.unique's
// order, and neighbour comparison + a scan drops the duplicates.  (np.unique is a sort; the sorted unique
// set does not depend on how it is obtained.)
constexpr int EXP_MAX_K1 = 128;

__global__ void __launch_bounds__(EXP_THREADS)
expand_v_kernel(const int32_t* __restrict__ rfull, const int32_t* __restrict__ rfull_cnt, int wf,
                const int32_t* __restrict__ rhalf, const int32_t* __restrict__ rhalf_cnt, int wh,
                PairSrc src, const float* __restrict__ sqnorm, int N, int cap, int P,
                int use_f16, int32_t* __restrict__ vcols, float* __restrict__ vvals, int32_t* __restrict__ vcnt) {
  extern __shared__ int sm_i[];
  int* K = sm_i;
  int* flag = K + P;
  float* val = reinterpret_cast<float*>(flag + P);
  __shared__ int s_nc[EXP_MAX_K1], s_inter[EXP_MAX_K1];
  __shared__ int warp_tot[EXP_THREADS / 32];
  __shared__ float s_red[EXP_THREADS / 32];
  constexpr int SENT = 0x7fffffff;

  const int i = blockIdx.x;
  const int tid = threadIdx.x;
  const int nb = rfull_cnt[i];
  const int32_t* base = rfull + (size_t)i * wf;
  const int used = wf + nb * wh;  // slots [0, used) may hold keys
  int Pn = 32;
  while (Pn < used) Pn <<= 1;     // <= P
  for (int t = tid; t < Pn; t += EXP_THREADS) K[t] = SENT;
  if (tid < nb) { s_nc[tid] = rhalf_cnt[base[tid]]; s_inter[tid] = 0; }
  __syncthreads();
  for (int t = tid; t < nb; t += EXP_THREADS) K[t] = base[t];
  for (int e = tid; e < nb * wh; e += EXP_THREADS) {
    const int t = e / wh, a = e - t * wh;
    if (a < s_nc[t]) K[wf + e] = rhalf[(size_t)base[t] * wh + a];
  }
  __syncthreads();
  // 2/3 rule against the UNEXPANDED R(i,k1) (faiss_utils.py:192-197): inter_t = #(R(c_t,k1/2) & R(i,k1))
  for (int e = tid; e < nb * wh; e += EXP_THREADS) {
    const int t = e / wh;
    const int x = K[wf + e];
    if (x != SENT) {
      int hit = 0;
      for (int b2 = 0; b2 < nb; ++b2) hit += (K[b2] == x);
      if (hit) atomicAdd(s_inter + t, hit);
    }
  }
  __syncthreads();
  for (int e = tid; e < nb * wh; e += EXP_THREADS) {
    const int t = e / wh;
    if (!((double)s_inter[t] > (2.0 / 3.0) * (double)s_nc[t])) K[wf + e] = SENT;
  }
  __syncthreads();
  // bitonic sort of K[0, Pn), ascending (INT_MAX padding ends up last)
  for (int k = 2; k <= Pn; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int idx = tid; idx < (Pn >> 1); idx += EXP_THREADS) {
        const int i1 = ((idx & ~(j - 1)) << 1) | (idx & (j - 1));
        const int i2 = i1 + j;
        const int x = K[i1], y = K[i2];
        const bool up = (i1 & k) == 0;
        if ((x > y) == up) { K[i1] = y; K[i2] = x; }
      }
      __syncthreads();
    }
  }
  for (int r = tid; r < Pn; r += EXP_THREADS) flag[r] = (K[r] != SENT && (r == 0 || K[r - 1] != K[r])) ? 1 : 0;
  __syncthreads();
  const int nu = block_exclusive_scan(flag, Pn, warp_tot);  // flag[r] -> output slot of K[r]
  float lmax = -INFINITY;
  const float xi = sqnorm[i];
  for (int r = tid; r < Pn; r += EXP_THREADS) {
    const int c = K[r];
    if (c != SENT && (r == 0 || K[r - 1] != c)) {
      // dist = 2 - 2 x_i.x_c (faiss_utils.py:201-204), with x_i.x_c = (|x_i|^2 + |x_c|^2 - sql2)/2
      const float dot = 0.5f * ((xi + sqnorm[c]) - pair_sql2_thread(src, sqnorm, i, c));
      const float nd = -(2.0f - 2.0f * dot);
      val[flag[r]] = nd;
      vcols[(size_t)i * cap + flag[r]] = c;
      lmax = fmaxf(lmax, nd);
    }
  }
  // softmax over the nu entries
  lmax = fmaxf(lmax, __shfl_xor_sync(0xffffffffu, lmax, 16));
  lmax = fmaxf(lmax, __shfl_xor_sync(0xffffffffu, lmax, 8));
  lmax = fmaxf(lmax, __shfl_xor_sync(0xffffffffu, lmax, 4));
  lmax = fmaxf(lmax, __shfl_xor_sync(0xffffffffu, lmax, 2));
  lmax = fmaxf(lmax, __shfl_xor_sync(0xffffffffu, lmax, 1));
  if ((tid & 31) == 0) s_red[tid >> 5] = lmax;
  __syncthreads();
  float gmax = s_red[0];
  for (int w = 1; w < EXP_THREADS / 32; ++w) gmax = fmaxf(gmax, s_red[w]);
  __syncthreads();
  float lsum = 0.f;
  for (int u = tid; u < nu; u += EXP_THREADS) {
    const float e = expf(val[u] - gmax);
    val[u] = e;
    lsum += e;
  }
  lsum = warp_sum(lsum);
  if ((tid & 31) == 0) s_red[tid >> 5] = lsum;
  __syncthreads();
  float gsum = 0.f;
  for (int w = 0; w < EXP_THREADS / 32; ++w) gsum += s_red[w];
  for (int u = tid; u < nu; u += EXP_THREADS) {
    float v = __fdiv_rn(val[u], gsum);
    if (use_f16) v = __half2float(__float2half_rn(v));
    vvals[(size_t)i * cap + u] = v;
  }
  if (tid == 0) vcnt[i] = nu;
}

// Warp-per-row variant for the common small case (k1 <= 32 and k1 + k1 * wh <= 256 candidate slots, e.g. k1 = 20):
// same slots, same 2/3 rule, but the 256 keys live in registers (8 per lane), the sort is warp_bitonic_sort (no
// block barriers) and the duplicate removal is a neighbour comparison + a warp scan.
constexpr int EXW_WARPS = 4;

__global__ void __launch_bounds__(EXW_WARPS * 32)
expand_v_warp_kernel(const int32_t* __restrict__ rfull, const int32_t* __restrict__ rfull_cnt, int wf,
                     const int32_t* __restrict__ rhalf, const int32_t* __restrict__ rhalf_cnt, int wh,
                     PairSrc src, const float* __restrict__ sqnorm, int N, int cap,
                     int use_f16, int32_t* __restrict__ vcols, float* __restrict__ vvals, int32_t* __restrict__ vcnt) {
  __shared__ int s_base[EXW_WARPS][32], s_nc[EXW_WARPS][32], s_inter[EXW_WARPS][32];
  constexpr uint32_t SENT = 0x7fffffffu;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int i = blockIdx.x * EXW_WARPS + w;
  if (i >= N) return;  // whole warps leave; the kernel has no block-wide barrier
  const int nb = rfull_cnt[i];
  int b_mine = -1, nc_mine = 0;
  if (lane < nb) {
    b_mine = rfull[(size_t)i * wf + lane];
    nc_mine = rhalf_cnt[b_mine];
  }
  s_base[w][lane] = b_mine;
  s_nc[w][lane] = nc_mine;
  s_inter[w][lane] = 0;
  __syncwarp();
  const uint32_t inv_wh = (65536u + (uint32_t)wh - 1u) / (uint32_t)wh;  // q / wh == (q * inv_wh) >> 16 for q < 256
  uint32_t k[8], nov[8];
  int tl[8];
#pragma unroll
  for (int s = 0; s < 8; ++s) {
    const int e = lane * 8 + s;
    uint32_t key = SENT;
    int t = -1;
    if (e < wf) {
      if (e < nb) key = (uint32_t)s_base[w][e];
    } else {
      const uint32_t q = (uint32_t)(e - wf);
      const int tt = (int)((q * inv_wh) >> 16);
      const int a = (int)q - tt * wh;
      if (tt < nb && a < s_nc[w][tt]) {
        key = (uint32_t)rhalf[(size_t)s_base[w][tt] * wh + a];
        t = tt;
      }
    }
    k[s] = key;
    tl[s] = t;
  }
  // 2/3 rule against the UNEXPANDED R(i,k1) (faiss_utils.py:192-197): inter_t = #(R(c_t,k1/2) & R(i,k1))
  int hit[8];
#pragma unroll
  for (int s = 0; s < 8; ++s) hit[s] = 0;
  for (int b2 = 0; b2 < nb; ++b2) {
    const uint32_t bv = (uint32_t)s_base[w][b2];
#pragma unroll
    for (int s = 0; s < 8; ++s) hit[s] += (k[s] == bv);
  }
#pragma unroll
  for (int s = 0; s < 8; ++s)
    if (tl[s] >= 0 && hit[s]) atomicAdd(&s_inter[w][tl[s]], hit[s]);
  __syncwarp();
#pragma unroll
  for (int s = 0; s < 8; ++s)
    if (tl[s] >= 0 && !((double)s_inter[w][tl[s]] > (2.0 / 3.0) * (double)s_nc[w][tl[s]])) k[s] = SENT;
  warp_bitonic_sort<8, false>(k, nov, lane);
  // unique keys (np.unique order), their output slots
  const uint32_t last_prev = __shfl_up_sync(0xffffffffu, k[7], 1);
  bool head[8];
  int cnt = 0;
#pragma unroll
  for (int s = 0; s < 8; ++s) {
    const uint32_t prev = s == 0 ? last_prev : k[s - 1];
    head[s] = k[s] != SENT && ((lane == 0 && s == 0) || prev != k[s]);
    cnt += head[s] ? 1 : 0;
  }
  int nu;
  int slot = warp_exclusive_scan(cnt, lane, nu);
  const float xi = sqnorm[i];
  float nd[8];
  float lmax = -INFINITY;
  // squared L2 of (i, candidate) for this lane's unique candidates
  float d2[8];
  uint32_t need = 0;  // candidates whose value is neither in a matrix nor in the row's own top-k1 list
#pragma unroll
  for (int s = 0; s < 8; ++s) {
    d2[s] = 0.f;
    if (head[s]) {
      const int c = (int)k[s];
      if (src.sql2) {
        d2[s] = (src.tri && !tri_stored(i, c)) ? src.sql2[(size_t)c * src.lds + i] : src.sql2[(size_t)i * src.lds + c];
      } else {
        bool hit = false;
        for (int t = 0; t < src.k1c; ++t)
          if (src.rank_i[(size_t)i * src.k1c + t] == c) { d2[s] = src.rank_d[(size_t)i * src.k1c + t]; hit = true; }
        if (!hit) need |= 1u << s;
      }
    }
  }
  if (!src.sql2) {  // warp-uniform: the misses of all lanes, one cooperative dot product each
    const float* xa = src.feats + (size_t)i * src.ldf;
    const bool vec = (src.d & 3) == 0 && (src.ldf & 3) == 0 && (reinterpret_cast<uintptr_t>(src.feats) & 15) == 0;
#pragma unroll
    for (int s = 0; s < 8; ++s) {
      uint32_t todo = __ballot_sync(0xffffffffu, (need >> s) & 1u);
      while (todo) {
        const int owner = __ffs(todo) - 1;
        todo &= todo - 1;
        const int c = (int)__shfl_sync(0xffffffffu, k[s], owner);
        const float dot = pair_dot_warp(xa, src.feats + (size_t)c * src.ldf, src.d, lane, vec);
        if (lane == owner) d2[s] = fmaf(-2.0f, dot, __fadd_rn(xi, sqnorm[c]));
      }
    }
  }
#pragma unroll
  for (int s = 0; s < 8; ++s) {
    nd[s] = -INFINITY;
    if (head[s]) {
      // dist = 2 - 2 x_i.x_c (faiss_utils.py:201-204), with x_i.x_c = (|x_i|^2 + |x_c|^2 - sql2)/2
      const int c = (int)k[s];
      const float dot = 0.5f * ((xi + sqnorm[c]) - d2[s]);
      nd[s] = -(2.0f - 2.0f * dot);
      lmax = fmaxf(lmax, nd[s]);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) lmax = fmaxf(lmax, __shfl_xor_sync(0xffffffffu, lmax, o));
  float lsum = 0.f;
#pragma unroll
  for (int s = 0; s < 8; ++s)
    if (head[s]) { nd[s] = expf(nd[s] - lmax); lsum += nd[s]; }
  const float gsum = warp_sum(lsum);
#pragma unroll
  for (int s = 0; s < 8; ++s)
    if (head[s]) {
      float v = __fdiv_rn(nd[s], gsum);
      if (use_f16) v = __half2float(__float2half_rn(v));
      vcols[(size_t)i * cap + slot] = (int32_t)k[s];
      vvals[(size_t)i * cap + slot] = v;
      ++slot;
    }
  if (lane == 0) vcnt[i] = nu;
}

// ------------------------------------------------------------------------------------ D5
// V_qe[i] = mean(V[rank[i,:k2]], axis=0) (faiss_utils.py:209-214): merge of k2 sorted sparse rows.
// Sums run in neighbour order j = 0..k2-1 like the reference's row-by-row reduction (zeros are exact
// no-ops), then one fp32 division by k2.  Dynamic smem: mcol[cap2] head[cap2] (ints), mval[cap2].
constexpr int QE_THREADS = 128;
constexpr int QE_MAX_K2 = 32;

__global__ void __launch_bounds__(QE_THREADS)
qexpand_kernel(const int32_t* __restrict__ rank, int k1_cols, int k2, const int32_t* __restrict__ vcols,
               const float* __restrict__ vvals, const int32_t* __restrict__ vcnt, int cap1, int cap2, int use_f16,
               int32_t* __restrict__ qcols, float* __restrict__ qvals, int32_t* __restrict__ qcnt,
               const int32_t* __restrict__ row_list, const int32_t* __restrict__ row_list_len, int N,
               int32_t* __restrict__ colcnt, int32_t* __restrict__ qrank) {
  extern __shared__ int sm_i[];
  // rows: all of them, or (after qexpand_warp_kernel) the listed ones it left behind
  const int nrows = row_list ? *row_list_len : N;
  for (int idx = blockIdx.x; idx < nrows; idx += gridDim.x) {
  const int i = row_list ? row_list[idx] : idx;
  __syncthreads();  // shared memory of the previous row is free
  int* mcol = sm_i;
  int* head = mcol + cap2;
  float* mval = reinterpret_cast<float*>(head + cap2);
  __shared__ int s_row[QE_MAX_K2], s_len[QE_MAX_K2], s_off[QE_MAX_K2 + 1];
  __shared__ int warp_tot[QE_THREADS / 32];

  const int tid = threadIdx.x;
  if (tid < k2) {
    const int r = rank[(size_t)i * k1_cols + tid];
    s_row[tid] = r;
    s_len[tid] = vcnt[r];
  }
  __syncthreads();
  if (tid == 0) {
    int run = 0;
    for (int j = 0; j < k2; ++j) { s_off[j] = run; run += s_len[j]; }
    s_off[k2] = run;
  }
  __syncthreads();
  const int n = s_off[k2];
  // merged position of element (j,t): #elements with smaller column, or equal column and smaller j
  for (int e = tid; e < n; e += QE_THREADS) {
    int j = 0;
    while (e >= s_off[j + 1]) ++j;
    const int t = e - s_off[j];
    const int32_t* mine = vcols + (size_t)s_row[j] * cap1;
    const int c = mine[t];
    int pos = t;
    for (int o = 0; o < k2; ++o) {
      if (o == j) continue;
      const int32_t* oc = vcols + (size_t)s_row[o] * cap1;
      int lo = 0, hi = s_len[o];
      if (o < j) {  // upper bound: entries <= c
        while (lo < hi) { const int mid = (lo + hi) >> 1; if (oc[mid] <= c) lo = mid + 1; else hi = mid; }
      } else {      // lower bound: entries < c
        while (lo < hi) { const int mid = (lo + hi) >> 1; if (oc[mid] < c) lo = mid + 1; else hi = mid; }
      }
      pos += lo;
    }
    mcol[pos] = c;
    mval[pos] = vvals[(size_t)s_row[j] * cap1 + t];
  }
  __syncthreads();
  for (int p = tid; p < n; p += QE_THREADS) head[p] = (p == 0 || mcol[p - 1] != mcol[p]) ? 1 : 0;
  __syncthreads();
  // keep the head flags readable after the scan: stash them in the sign of mcol? no -- recompute below
  const int nu = block_exclusive_scan(head, n, warp_tot);
  for (int p = tid; p < n; p += QE_THREADS) {
    if (p == 0 || mcol[p - 1] != mcol[p]) {
      const int c = mcol[p];
      float s = mval[p];
      for (int t = p + 1; t < n && mcol[t] == c; ++t) s = __fadd_rn(s, mval[t]);
      float v = __fdiv_rn(s, (float)k2);
      if (use_f16) v = __half2float(__float2half_rn(v));
      const int slot = head[p];
      qcols[(size_t)i * cap2 + slot] = c;
      qvals[(size_t)i * cap2 + slot] = v;
      if (colcnt && v != 0.f) qrank[(size_t)i * cap2 + slot] = atomicAdd(colcnt + c, 1);  // D6 column counts, fused: the old value is the entry's place in its column
    }
  }
  if (tid == 0) qcnt[i] = nu;
  }
}

// Warp-per-row variant for rows whose k2 lists hold at most 256 entries in total (the common case): the entries are
// keyed (column, position in the concatenated lists), sorted by warp_bitonic_sort, and every run of equal columns is
// summed in neighbour order by its first element, the weights read from shared memory through the key's low bits.  Rows with more entries are
// appended to long_list and left to qexpand_kernel.
constexpr int QEW_WARPS = 4;

__global__ void __launch_bounds__(QEW_WARPS * 32)
qexpand_warp_kernel(const int32_t* __restrict__ rank, int k1_cols, int k2, const int32_t* __restrict__ vcols,
                    const float* __restrict__ vvals, const int32_t* __restrict__ vcnt, int N, int cap1, int cap2,
                    int use_f16, int32_t* __restrict__ qcols, float* __restrict__ qvals, int32_t* __restrict__ qcnt,
                    int32_t* __restrict__ colcnt, int32_t* __restrict__ qrank, int32_t* __restrict__ long_list,
                    int32_t* __restrict__ long_len) {
  __shared__ int s_row[QEW_WARPS][32], s_off[QEW_WARPS][33];
  __shared__ uint32_t s_key[QEW_WARPS][256];
  __shared__ float s_val[QEW_WARPS][256];
  constexpr uint32_t SENT = 0xffffffffu;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int i = blockIdx.x * QEW_WARPS + w;
  if (i >= N) return;  // whole warps leave; the kernel has no block-wide barrier
  int r = -1, len = 0;
  if (lane < k2) {
    r = rank[(size_t)i * k1_cols + lane];
    len = vcnt[r];
  }
  int n;
  const int off = warp_exclusive_scan(len, lane, n);
  if (n > 256) {
    if (lane == 0) long_list[atomicAdd(long_len, 1)] = i;
    return;
  }
  s_row[w][lane] = r;
  s_off[w][lane] = off;  // lanes >= k2 hold n
  if (lane == 31) s_off[w][32] = n;
  __syncwarp();
  // key = (column << 8) | element index e: e grows with the neighbour index j, so sorting the keys alone orders every
  // column's entries by neighbour; the weights stay in shared memory at their original positions (no payload)
  uint32_t k[8], nov[8];
#pragma unroll
  for (int s = 0; s < 8; ++s) {
    const int e = lane * 8 + s;
    k[s] = SENT;
    if (e < n) {
      int j = 0;
      while (e >= s_off[w][j + 1]) ++j;
      const size_t src = (size_t)s_row[w][j] * cap1 + (e - s_off[w][j]);
      k[s] = ((uint32_t)vcols[src] << 8) | (uint32_t)e;
      s_val[w][e] = vvals[src];
    }
  }
  warp_bitonic_sort<8, false>(k, nov, lane);
#pragma unroll
  for (int s = 0; s < 8; ++s) s_key[w][lane * 8 + s] = k[s];
  __syncwarp();
  const uint32_t last_prev = __shfl_up_sync(0xffffffffu, k[7], 1);
  bool head[8];
  int cnt = 0;
#pragma unroll
  for (int s = 0; s < 8; ++s) {
    const uint32_t prev = s == 0 ? last_prev : k[s - 1];
    head[s] = k[s] != SENT && ((lane == 0 && s == 0) || (prev >> 8) != (k[s] >> 8));
    cnt += head[s] ? 1 : 0;
  }
  int nu;
  int slot = warp_exclusive_scan(cnt, lane, nu);
#pragma unroll
  for (int s = 0; s < 8; ++s)
    if (head[s]) {
      const int e = lane * 8 + s;
      const uint32_t c = k[s] >> 8;
      float sum = s_val[w][k[s] & 255u];
      for (int t = e + 1; t < n && (s_key[w][t] >> 8) == c; ++t) sum = __fadd_rn(sum, s_val[w][s_key[w][t] & 255u]);
      float val = __fdiv_rn(sum, (float)k2);
      if (use_f16) val = __half2float(__float2half_rn(val));
      qcols[(size_t)i * cap2 + slot] = (int32_t)c;
      qvals[(size_t)i * cap2 + slot] = val;
      if (colcnt && val != 0.f) qrank[(size_t)i * cap2 + slot] = atomicAdd(colcnt + c, 1);  // D6 column counts, fused: the old value is the entry's place in its column
      ++slot;
    }
  if (lane == 0) qcnt[i] = nu;
}

// ------------------------------------------------------------------------------------ D6
__global__ void inv_count_kernel(const int32_t* __restrict__ cols, const float* __restrict__ vals,
                                 const int32_t* __restrict__ cnt, int N, int cap, int32_t* __restrict__ colcnt,
                                 int32_t* __restrict__ qrank) {
  const int i = blockIdx.x;
  const int n = cnt[i];
  for (int t = threadIdx.x; t < n; t += blockDim.x)
    if (vals[(size_t)i * cap + t] != 0.f) qrank[(size_t)i * cap + t] = atomicAdd(colcnt + cols[(size_t)i * cap + t], 1);
}

// colptr[0..N] = exclusive scan of colcnt; cursor = copy of colptr.  Single CTA; dynamic smem: N ints.
__global__ void __launch_bounds__(1024)
inv_scan_kernel(const int32_t* __restrict__ colcnt, int N, int32_t* __restrict__ colptr, int32_t* __restrict__ cursor) {
  extern __shared__ int sm_i[];
  __shared__ int wsum[32];
  const int total = cta_exclusive_scan_smem(colcnt, N, sm_i, wsum);
  for (int r = threadIdx.x; r < N; r += 1024) { const int v = sm_i[r]; colptr[r] = v; cursor[r] = v; }
  if (threadIdx.x == 0) colptr[N] = total;
}

// Warp per row, four entries per lane in flight.  An entry's place in its column was fixed when the column was counted
// (qrank = the return value of that atomic), so the fill is loads and stores only: with a cursor atomic per entry (CTA per
// row) it ran at 7 % issue activity behind ~100 stalled warps per issue and the atomics' throttle (ncu, hard data).
// The order inside an inverted list is whatever the counting atomics made it; its consumer sums order-independently.
constexpr int IF_WARPS = 8;
__global__ void __launch_bounds__(IF_WARPS * 32)
inv_fill_kernel(const int32_t* __restrict__ cols, const float* __restrict__ vals, const int32_t* __restrict__ cnt,
                const int32_t* __restrict__ qrank, int N, int cap, const int32_t* __restrict__ colptr,
                int32_t* __restrict__ inv_row, float* __restrict__ inv_val) {
  const int i = blockIdx.x * IF_WARPS + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (i >= N) return;
  const int n = cnt[i];
  constexpr int U = 4;
  for (int t0 = lane; t0 < n; t0 += 32 * U) {
    float v[U];
    int c[U], r[U], p[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int t = t0 + 32 * u;
      v[u] = t < n ? vals[(size_t)i * cap + t] : 0.f;
      c[u] = t < n ? cols[(size_t)i * cap + t] : 0;
      r[u] = t < n ? qrank[(size_t)i * cap + t] : 0;
    }
#pragma unroll
    for (int u = 0; u < U; ++u) p[u] = v[u] != 0.f ? __ldg(colptr + c[u]) + r[u] : -1;
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (p[u] >= 0) { inv_row[p[u]] = i; inv_val[p[u]] = v[u]; }
  }
}

// ------------------------------------------------------------------------------------ D7 (+ D8)
// One CTA per output row i: acc[r] = sum_j min(V[i,j], V[r,j]) over the inverted lists of the
// non-zeros of row i (faiss_utils.py:223-234), staged in shared memory (N floats), then the whole
// dense row 1 - t/(2-t) (clamped at 0, :236-240) is written once, coalesced.
// The scatter is flattened: the row's non-zeros and the extents of their inverted lists are staged in
// shared memory first (one dependent-load chain per CTA instead of one per entry), an exclusive scan of
// the list lengths turns (entry, position) into one index space, and every thread walks that space with
// a stride of the CTA size, several independent loads in flight.
// The accumulator is 2.30 fixed point (sums of minima of softmax weights lie in [0, 1]; every term is rounded to
// a multiple of 2^-30, the sum itself is exact): shared-memory integer atomics are native -- fp32 ones are
// compare-and-swap loops that serialise on the few hundred hot rows of a cluster -- and the result no longer
// depends on the order in which the atomics land.
// Dynamic smem: acc[N4] | e_val[cap] | e_beg[cap] (ints) | e_off[cap + 1] (ints), N4 = N rounded up to 4.
// Optional fused evaluation (D8 -> B4-B6): when the output block is the Q x G block of a re-ranked evaluation, the CTA
// that has just produced query row q still holds it in registers and reduces it to the query's AP / first hit right
// away (eval_row_in_registers, evalmap.cuh) -- the re-ranked matrix is then neither re-read nor, if `out` is null,
// written at all (2 x 214 MB of HBM traffic at Market size).
struct EvalFuse {
  const int32_t *ql, *qc, *gl, *gc;
  double* ap_q;
  int32_t *first, *hist, *status;
  int enabled;
};

constexpr int JR_THREADS = 512;
constexpr int JR_PF = 32;                        // squared-L2 values prefetched per thread (covers 16384 columns)
constexpr float JR_FIX = 1073741824.0f;          // 2^30
constexpr float JR_UNFIX = 1.0f / 1073741824.0f;

// 1 - t / (2 - t), clamped at 0 (faiss_utils.py:236-240).  Almost every entry of a row has t == 0 (no shared
// neighbour): that case is exactly 1 and stays inline, the division lives out of line (the output loop is unrolled).
static __device__ __noinline__ float jaccard_nonzero(float t, int use_f16) {
  if (use_f16) t = __half2float(__float2half_rn(t));
  const float jd = __fsub_rn(1.0f, __fdiv_rn(t, __fsub_rn(2.0f, t)));
  return jd < 0.f ? 0.f : jd;
}

// lambda blend of one entry (D8): jd * (1 - lambda) + sqrt(max(d2, 1e-12)) * lambda, every operation rounded on its own
static __device__ __noinline__ float jr_blend_value(float jd, float d2, float one_m_l, float lambda_value) {
  return __fadd_rn(__fmul_rn(jd, one_m_l), __fmul_rn(sqrtf(fmaxf(d2, 1e-12f)), lambda_value));
}

__device__ __forceinline__ float jaccard_value(float t, int use_f16) {
  return t == 0.f ? 1.0f : jaccard_nonzero(t, use_f16);
}

__global__ void __launch_bounds__(JR_THREADS, 2)
jaccard_rows_kernel(const int32_t* __restrict__ qcols, const float* __restrict__ qvals, const int32_t* __restrict__ qcnt,
                    int cap, const int32_t* __restrict__ colptr, const int32_t* __restrict__ inv_row,
                    const float* __restrict__ inv_val, int N, int row0, int col0, int ncols, float lambda_value,
                    const float* __restrict__ sql2, int64_t lds, int use_f16, void* __restrict__ out, int64_t ldo,
                    EvalFuse ev) {
  extern __shared__ __align__(16) float acc[];
  const int N4 = (N + 3) & ~3;
  unsigned int* acci = reinterpret_cast<unsigned int*>(acc);
  float* e_val = acc + N4;
  int* e_beg = reinterpret_cast<int*>(e_val + cap);
  int* e_off = e_beg + cap;
  __shared__ int warp_tot[JR_THREADS / 32];
  const int i = row0 + blockIdx.x;
  const int tid = threadIdx.x;
  // lambda blend: this row's slice of the squared-L2 matrix is requested NOW (JR_PF independent loads per thread,
  // the whole 64 KB slice of a Market-sized row in flight at once) and consumed after the scatter
  const bool blend = lambda_value > 0.f;
  const float* srow = sql2 + (size_t)i * lds + col0;
  float pre[JR_PF];
  // fused-evaluation path (nothing stored, fp32, blended): the thread holds groups of FOUR consecutive columns, so the
  // distance slice, the gallery labels and the accumulator are read with 128-bit accesses (ev_col<.., 4>)
  const bool fast = ev.enabled && blend && !out && !use_f16 && (col0 & 3) == 0 && (lds & 3) == 0 &&
                    ((reinterpret_cast<uintptr_t>(sql2) | reinterpret_cast<uintptr_t>(ev.gl)) & 15) == 0;
  if (fast) {
#pragma unroll
    for (int g = 0; g < JR_PF / 4; ++g) {
      const int c0 = 4 * (tid + g * JR_THREADS);
      // a group that straddles ncols still lies inside the row's allocation (lds >= N rounded up to 32 columns)
      const float4 v = c0 < ncols ? __ldcs(reinterpret_cast<const float4*>(srow + c0)) : make_float4(0.f, 0.f, 0.f, 0.f);
      pre[4 * g] = v.x; pre[4 * g + 1] = v.y; pre[4 * g + 2] = v.z; pre[4 * g + 3] = v.w;
    }
  } else if (blend) {
#pragma unroll
    for (int k = 0; k < JR_PF; ++k) {
      const int c = tid + k * JR_THREADS;
      pre[k] = c < ncols ? __ldcs(srow + c) : 0.f;
    }
  }
  for (int r = tid; r < N4 / 4; r += JR_THREADS) reinterpret_cast<float4*>(acc)[r] = make_float4(0.f, 0.f, 0.f, 0.f);
  const int n = qcnt[i];
  // fused evaluation: which of this thread's columns are good items of the query (labels only -- done here, under the
  // latency of the loads above and below, the row values arrive at the end)
  const uint32_t goodmask = !ev.enabled ? 0u
                            : fast ? eval_good_mask<JR_THREADS, JR_PF, 4>(ncols, (int)blockIdx.x, ev.ql, ev.qc, ev.gl, ev.gc)
                                   : eval_good_mask<JR_THREADS, JR_PF, 1>(ncols, (int)blockIdx.x, ev.ql, ev.qc, ev.gl, ev.gc);
  for (int t = tid; t < n; t += JR_THREADS) {
    const float v = qvals[(size_t)i * cap + t];
    const int j = qcols[(size_t)i * cap + t];
    const int b = colptr[j], e = colptr[j + 1];
    e_val[t] = v;
    e_beg[t] = b;
    e_off[t] = v == 0.f ? 0 : e - b;
  }
  __syncthreads();
  const int total = block_exclusive_scan(e_off, n, warp_tot);  // e_off[t] = first flat index of entry t
  if (tid == 0) e_off[n] = total;
  __syncthreads();
  {
    int t = 0;
    constexpr int U = 4;
    for (int w0 = tid; w0 < total; w0 += U * JR_THREADS) {
      int rr[U];
      float vv[U], mv[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int w = w0 + u * JR_THREADS;
        rr[u] = -1;
        if (w < total) {
          while (w >= e_off[t + 1]) ++t;  // entries with an empty list are skipped (equal offsets)
          const int p = e_beg[t] + (w - e_off[t]);
          rr[u] = __ldg(inv_row + p);
          vv[u] = __ldg(inv_val + p);
          mv[u] = e_val[t];
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (rr[u] >= 0) atomicAdd(acci + rr[u], __float2uint_rn(__fmul_rn(fminf(mv[u], vv[u]), JR_FIX)));
    }
  }
  __syncthreads();
  if (!fast)  // the fused path converts the few non-zero sums where it reads them
  for (int r = tid; r < N4 / 4; r += JR_THREADS) {  // fixed point -> fp32, in place
    const uint4 a4 = reinterpret_cast<const uint4*>(acci)[r];
    if ((a4.x | a4.y | a4.z | a4.w) != 0u)
      reinterpret_cast<float4*>(acc)[r] = make_float4(__fmul_rn(__uint2float_rn(a4.x), JR_UNFIX), __fmul_rn(__uint2float_rn(a4.y), JR_UNFIX),
                                                      __fmul_rn(__uint2float_rn(a4.z), JR_UNFIX), __fmul_rn(__uint2float_rn(a4.w), JR_UNFIX));
  }
  if (!fast) __syncthreads();  // `fast` is uniform over the CTA
  const float one_m_l = 1.0f - lambda_value;
  float* orow = reinterpret_cast<float*>(out) + (size_t)blockIdx.x * ldo;
  __half* hrow = reinterpret_cast<__half*>(out) + (size_t)blockIdx.x * ldo;
  if (fast) {
    // Almost every entry of a row has no shared neighbour (t == 0, Jaccard distance exactly 1): its final value
    // (1 - lambda) + lambda * sqrt(d2) is monotone in the squared distance the thread already holds, so the
    // evaluation finishes it LAZILY -- only for the good items and for the entries whose d2 is below a bound derived
    // from the worst good item (eval_row_in_registers).  Entries with shared neighbours are finished here, before the
    // accumulator's shared memory becomes the scratch.
    const unsigned int* arow = acci + col0;
    uint32_t donemask = 0;
#pragma unroll
    for (int g = 0; g < JR_PF / 4; ++g) {
      const int c0 = 4 * (tid + g * JR_THREADS);
      if (c0 < ncols) {  // acc holds N rounded up to 4 entries: the straddling group stays in bounds
        const uint4 a4 = *reinterpret_cast<const uint4*>(arow + c0);  // still 2.30 fixed point (no conversion pass here)
        const unsigned int aa[4] = {a4.x, a4.y, a4.z, a4.w};
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          if (aa[u] != 0u && c0 + u < ncols) {
            const float t = __fmul_rn(__uint2float_rn(aa[u]), JR_UNFIX);
            pre[4 * g + u] = jr_blend_value(jaccard_nonzero(t, 0), pre[4 * g + u], one_m_l, lambda_value);
            donemask |= 1u << (4 * g + u);
          }
        }
      }
    }
    __syncthreads();  // the accumulator is dead: its shared memory becomes the evaluation scratch
    auto finish = [&](float d2) { return jr_blend_value(1.0f, d2, one_m_l, lambda_value); };
    // finish(d2) <= w  =>  lambda * sqrt(d2) <= w - (1 - lambda) + a few ulp of w; the slack below is 5 x that
    auto raw_bound = [&](float w) {
      const float sb = __fdiv_rn(__fadd_rn(__fsub_rn(w, one_m_l), __fmul_rn(fabsf(w), 1e-6f)), lambda_value);
      if (!(sb >= 0.f)) return __int_as_float(0xff800000);  // -inf: no entry without shared neighbours can be that close
      const float s1 = __fmul_rn(sb, 1.00001f);
      return fmaxf(__fmul_rn(__fmul_rn(s1, s1), 1.0001f), 1.1e-12f);  // >= the 1e-12 clamp of negative d2
    };
    eval_row_in_registers<JR_THREADS, JR_PF, 4>(pre, donemask, goodmask, finish, raw_bound, ncols, (int)blockIdx.x, 0, ev.ql, ev.qc,
                                             ev.gl, ev.gc, ev.ap_q, ev.first, ev.hist, ev.status,
                                             reinterpret_cast<EvalScratch*>(acc));
    return;
  } else {
#pragma unroll
  for (int k = 0; k < JR_PF; ++k) {
    const int c = tid + k * JR_THREADS;
    if (c < ncols) {
      float jd = jaccard_value(acc[col0 + c], use_f16);
      if (blend) jd = __fadd_rn(__fmul_rn(jd, one_m_l), __fmul_rn(sqrtf(fmaxf(pre[k], 1e-12f)), lambda_value));
      pre[k] = jd;  // the row stays in registers for the fused evaluation
      if (out) {
        if (use_f16) hrow[c] = __float2half_rn(jd);
        else __stcs(orow + c, jd);
      }
    }
  }
  }
  for (int c = tid + JR_PF * JR_THREADS; c < ncols; c += JR_THREADS) {  // rows wider than JR_PF * JR_THREADS columns
    float jd = jaccard_value(acc[col0 + c], use_f16);
    if (blend) jd = __fadd_rn(__fmul_rn(jd, one_m_l), __fmul_rn(sqrtf(fmaxf(__ldcs(srow + c), 1e-12f)), lambda_value));
    if (use_f16) hrow[c] = __float2half_rn(jd);
    else __stcs(orow + c, jd);
  }
  if (ev.enabled) {  // host side guarantees ncols <= JR_PF * JR_THREADS, fp32 output, shared memory >= EvalScratch
    __syncthreads();  // the accumulator is dead: its shared memory becomes the evaluation scratch
    eval_row_in_registers<JR_THREADS, JR_PF, 1>(pre, ~0u, goodmask, [](float v) { return v; }, [](float w) { return w; }, ncols,
                                             (int)blockIdx.x, 0, ev.ql, ev.qc, ev.gl, ev.gc, ev.ap_q, ev.first, ev.hist, ev.status,
                                             reinterpret_cast<EvalScratch*>(acc));
  }
}

// ------------------------------------------------------------------------------------ D7 + D8 + B4-B6, evaluation only
// The fused re-ranked EVALUATION of one query row (fp32, lambda > 0, nothing stored) as a kernel of its own: same
// arithmetic as jaccard_rows_kernel's fused path, bit for bit, organised around what a row of a HARD evaluation set costs
// (75 non-zeros in the V_qe row, 6300 scatter terms, a fifth of the columns with shared neighbours, 75 -- up to 6000 --
// columns ranked before the worst good item; profiles/r02_jaccard_eval_rows.md):
//   * the accumulator covers only the output columns [col0, col0 + ncols) (partners outside it are dropped at the scatter)
//     and is read exactly once more, by the thread that holds the column's squared distance in a register;
//   * the query's good items and same-camera junk come from per-query lists built once per evaluation
//     (query_lists_kernel), the label -1 junk from a bit mask in the row kernel's register layout (junk_layout_kernel):
//     no label is read per row;
//   * the scatter walks (entry, position) space in warp-contiguous chunks, lanes interleaved: coalesced list reads, each
//     lane's entry index advances monotonically, every load of a batch issued before the first loaded value is used;
//   * per column ONE branch-free superset test (a quadratic lower bound of the Jaccard distance, no division, no square
//     root), 32 independent chains per thread; the columns that pass are staged and their exact values computed by the
//     whole CTA afterwards, evenly spread.
// Dynamic smem: acc[nc4] (2.30 fixed point) | max(entries: e_val[cap] e_beg[cap] e_off[cap + 1], EvalScratch) |
//               glist[JE_GLIST] gd2[JE_GLIST] jlist[QL_STRIDE].
// REIDB200_JE_DEBUG (timing experiments): 2 = no scatter, 4 = no candidates (both give wrong results), 16 = cycles per
// phase and row statistics into status[8..] (tools/je_phases.py).
constexpr int JE_GLIST = EV_GOOD_CAP;

constexpr int QL_STRIDE = 128;  // gallery items with the query's id kept per query by query_lists_kernel (more: label scan in the row kernel)
constexpr int JE_CAND = EV_GOOD_CAP;  // candidate columns staged per row (more: finished in place)

// Per query: the gallery columns that carry the query's id (evaluate.py:66-72), bit 31 set when the camera is the same
// (junk) and clear when it differs (a good item).  Eight queries per CTA share one pass over the label vector.
// qlcnt[q] may exceed QL_STRIDE: the row kernel then scans the labels itself.
constexpr int QLK_Q = 8;
// rare path, out of line on purpose: inlined at its 17 call sites the kernel did not fit the instruction cache
// (ncu: 28 stalled warps per issue waiting for instructions, 55 us for 0.2 MB of labels)
static __device__ __noinline__ void ql_hit(int j, int32_t l, const int32_t* s_l, const int32_t* s_c, int* s_n, int q0, int Q,
                                           const int32_t* __restrict__ gc, int32_t* __restrict__ qlist) {
  const int32_t c = __ldg(gc + j);
  for (int x = 0; x < QLK_Q; ++x)
    if (l == s_l[x] && q0 + x < Q) {
      const int pos = atomicAdd(&s_n[x], 1);
      if (pos < QL_STRIDE) qlist[(size_t)(q0 + x) * QL_STRIDE + pos] = j | (c == s_c[x] ? (int32_t)0x80000000 : 0);
    }
}

__global__ void __launch_bounds__(256)
query_lists_kernel(const int32_t* __restrict__ ql, const int32_t* __restrict__ qc, const int32_t* __restrict__ gl,
                   const int32_t* __restrict__ gc, int Q, int G, int32_t* __restrict__ qlist, int32_t* __restrict__ qlcnt) {
  __shared__ int32_t s_l[QLK_Q], s_c[QLK_Q];
  __shared__ int s_n[QLK_Q];
  const int q0 = blockIdx.x * QLK_Q, tid = threadIdx.x;
  if (tid < QLK_Q) {
    const int q = min(q0 + tid, Q - 1);
    s_l[tid] = ql[q]; s_c[tid] = qc[q]; s_n[tid] = 0;
  }
  __syncthreads();
  int32_t ll[QLK_Q];
#pragma unroll
  for (int x = 0; x < QLK_Q; ++x) ll[x] = s_l[x];
  auto hit = [&](int j, int32_t l) { ql_hit(j, l, s_l, s_c, s_n, q0, Q, gc, qlist); };
  auto any_of = [&](int32_t l) {
    bool any = false;
#pragma unroll
    for (int x = 0; x < QLK_Q; ++x) any |= l == ll[x];
    return any;
  };
  // four labels per load, four loads in flight per thread (the pass is a chain of L2 round trips otherwise)
  const int G4 = (reinterpret_cast<uintptr_t>(gl) & 15) == 0 ? G >> 2 : 0;
  constexpr int UL = 4;
  for (int j0 = tid; j0 < G4; j0 += 256 * UL) {
    int4 lab[UL];
#pragma unroll
    for (int u = 0; u < UL; ++u) {
      const int j4 = j0 + u * 256;
      lab[u] = j4 < G4 ? __ldg(reinterpret_cast<const int4*>(gl) + j4) : make_int4(ll[0], ll[0], ll[0], ll[0]);
    }
#pragma unroll
    for (int u = 0; u < UL; ++u) {
      const int j4 = j0 + u * 256;
      if (j4 >= G4) continue;
      if (any_of(lab[u].x)) hit(4 * j4, lab[u].x);
      if (any_of(lab[u].y)) hit(4 * j4 + 1, lab[u].y);
      if (any_of(lab[u].z)) hit(4 * j4 + 2, lab[u].z);
      if (any_of(lab[u].w)) hit(4 * j4 + 3, lab[u].w);
    }
  }
  for (int j = 4 * G4 + tid; j < G; j += 256) {
    const int32_t l = __ldg(gl + j);
    if (any_of(l)) hit(j, l);
  }
  __syncthreads();
  if (tid < QLK_Q && q0 + tid < Q) qlcnt[q0 + tid] = s_n[tid];
}

struct JeRow {  // what the exact per-column path needs, kept in one place so that the call passes one pointer
  EvalScratch* sc;
  const int* jlist;  // same id, same camera columns of this query (query_lists_kernel), n_junk of them
  int n_junk;
  int ngood;
  float worst_v, one_m_l, lambda_value;
};

// a column that may precede the worst good item: finish its value, exact compare, junk test, count
static __device__ __noinline__ void je_column(const JeRow& r, unsigned int a, float d2, int j) {
  const float jd = a ? jaccard_nonzero(__fmul_rn(__uint2float_rn(a), JR_UNFIX), 0) : 1.0f;
  const float v = jr_blend_value(jd, d2, r.one_m_l, r.lambda_value);
  if (!(v <= r.worst_v)) return;
  const uint32_t kj = ev_key(v, 0);
  EvalScratch* sc = r.sc;
  int lo = 0, hi = r.ngood - 1;
  if (!ev_before(kj, j, sc->skey[hi], sc->scol[hi], 0)) return;
  for (int e = 0; e < r.n_junk; ++e)
    if (r.jlist[e] == j) return;  // same id, same camera: junk (evaluate.py:71-72)
  while (lo < hi) {  // smallest s with j before good s (exists: j is before the worst good)
    const int mid = (lo + hi) >> 1;
    if (ev_before(kj, j, sc->skey[mid], sc->scol[mid], 0)) hi = mid; else lo = mid + 1;
  }
  atomicAdd(&sc->gpos[lo], 1);
}

// exclusive scan of data[0..n] (n + 1 values, data[n] = 0 on entry => data[n] = total on exit) by ONE warp
__device__ __forceinline__ void warp_exclusive_scan(int* data, int n1, int lane) {
  const int per = (n1 + 31) >> 5;
  const int b = lane * per, e = min(n1, b + per);
  int s = 0;
  for (int t = b; t < e; ++t) s += data[t];
  int incl = s;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += v;
  }
  int run = incl - s;
  for (int t = b; t < e; ++t) { const int v = data[t]; data[t] = run; run += v; }
}

// bit 4 g + u of word tid: column 4 (tid + g JR_THREADS) + u of the output block is junk for EVERY query (label -1,
// evaluate.py:71) or lies beyond the block -- jaccard_eval_rows_kernel's thread tid holds exactly these columns
__global__ void __launch_bounds__(JR_THREADS)
junk_layout_kernel(const int32_t* __restrict__ gl, int ncols, uint32_t* __restrict__ junk32) {
  const int tid = threadIdx.x;
  uint32_t m = 0;
  int32_t lab[JR_PF];
#pragma unroll
  for (int k = 0; k < JR_PF; ++k) {  // all 32 loads in flight before the first compare
    const int j = 4 * (tid + (k >> 2) * JR_THREADS) + (k & 3);
    lab[k] = j < ncols ? __ldg(gl + j) : -1;
  }
#pragma unroll
  for (int k = 0; k < JR_PF; ++k)
    if (lab[k] == -1) m |= 1u << k;
  junk32[tid] = m;
}

struct JeLists {  // per-query label lists (query_lists_kernel) + the label -1 layout (junk_layout_kernel)
  const int32_t* qlist;
  const int32_t* qlcnt;
  const uint32_t* junk32;
};

__global__ void __launch_bounds__(JR_THREADS, 2)
jaccard_eval_rows_kernel(const int32_t* __restrict__ qcols, const float* __restrict__ qvals, const int32_t* __restrict__ qcnt,
                         int cap, const int32_t* __restrict__ colptr, const int32_t* __restrict__ inv_row,
                         const float* __restrict__ inv_val, int row0, int col0, int ncols, float lambda_value,
                         const float* __restrict__ sql2, int64_t lds, int region_bytes, JeLists lists, EvalFuse ev) {
  extern __shared__ __align__(16) unsigned int je_smem[];
  const int nc4 = (ncols + 3) & ~3;
  unsigned int* acci = je_smem;
  unsigned char* region = reinterpret_cast<unsigned char*>(acci + nc4);
  float* e_val = reinterpret_cast<float*>(region);
  int* e_beg = reinterpret_cast<int*>(e_val + cap);
  int* e_off = e_beg + cap;
  EvalScratch* sc = reinterpret_cast<EvalScratch*>(region);  // alive once the entries are dead (after the scatter)
  int* glist = reinterpret_cast<int*>(region + region_bytes);  // good columns [JE_GLIST]
  float* gd2 = reinterpret_cast<float*>(glist + JE_GLIST);     // their squared distances [JE_GLIST]
  int* jlist = reinterpret_cast<int*>(gd2 + JE_GLIST);         // same id, same camera columns [QL_STRIDE]
  __shared__ int s_ngood, s_njunk, s_ncand;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int q = blockIdx.x;
  const int i = row0 + q;
  // timing experiments (REIDB200_JE_DEBUG bit 16): thread 0 adds the cycles of each phase, /16, to status[8 + phase]
  long long tick = (ev.enabled & 16) ? clock64() : 0;
  auto phase_done = [&](int ph) {
    if ((ev.enabled & 16) && tid == 0) {
      const long long now = clock64();
      atomicAdd(ev.status + 8 + ph, (int)((now - tick) >> 4));
      tick = now;
    }
  };

  // ---- the row's squared distances: requested now (the whole slice in flight), first used after the scatter
  const float* srow = sql2 + (size_t)i * lds + col0;
  float pre[JR_PF];
#pragma unroll
  for (int g = 0; g < JR_PF / 4; ++g) {
    const int c0 = 4 * (tid + g * JR_THREADS);
    // a group that straddles ncols still lies inside the row's allocation (lds >= the row length rounded up to 4)
    // (streaming: the slice is read once; the inverted lists and label lists it would evict from L2 are re-read by every row)
    const float4 v = c0 < ncols ? __ldcs(reinterpret_cast<const float4*>(srow + c0)) : make_float4(0.f, 0.f, 0.f, 0.f);
    pre[4 * g] = v.x; pre[4 * g + 1] = v.y; pre[4 * g + 2] = v.z; pre[4 * g + 3] = v.w;
  }
  const int nl = lists.qlcnt[q];
  if (tid == 0) { s_ncand = 0; s_ngood = 0; s_njunk = 0; }
  // bit 4 g + u: this thread's column 4 (tid + g JR_THREADS) + u has label -1 or lies beyond the block
  uint32_t junkmask = __ldg(lists.junk32 + tid);
  // ---- the non-zeros of V_qe row i and the extents of their inverted lists.  The first JR_THREADS slots are read
  // without waiting for the row's count (what the slots behind it hold is never used): one dependent round trip less
  // on the kernel's critical path.
  float v0 = 0.f;
  int j0 = 0;
  if (tid < cap) { v0 = __ldg(qvals + (size_t)i * cap + tid); j0 = __ldg(qcols + (size_t)i * cap + tid); }
  const int n = qcnt[i];
  if (tid < n) {
    const int b = colptr[j0], e = colptr[j0 + 1];
    e_val[tid] = v0;
    e_beg[tid] = b;
    e_off[tid] = v0 == 0.f ? 0 : e - b;
  }
  for (int t = tid + JR_THREADS; t < n; t += JR_THREADS) {
    const float v = qvals[(size_t)i * cap + t];
    const int j = qcols[(size_t)i * cap + t];
    const int b = colptr[j], e = colptr[j + 1];
    e_val[t] = v;
    e_beg[t] = b;
    e_off[t] = v == 0.f ? 0 : e - b;
  }
  if (tid == 0) e_off[n] = 0;  // the scan turns it into the total
  for (int r = tid; r < nc4 / 4; r += JR_THREADS) reinterpret_cast<uint4*>(acci)[r] = make_uint4(0u, 0u, 0u, 0u);
  __syncthreads();  // (also: the counters are set for everybody)
  phase_done(0);
  // ---- warp 0 scans the list lengths; the other warps fetch the query's label lists meanwhile: good items (same id,
  // other camera, evaluate.py:66-70) with their squared distances, and the same-camera junk (evaluate.py:71-72)
  if (warp == 0) warp_exclusive_scan(e_off, n + 1, lane);
  if (nl <= QL_STRIDE) {
    for (int e = tid - 32; e < nl && warp != 0; e += JR_THREADS - 32) {
      const int32_t item = __ldg(lists.qlist + (size_t)q * QL_STRIDE + e);
      const int c = item & 0x7fffffff;
      if (item < 0) {
        jlist[atomicAdd(&s_njunk, 1)] = c;
      } else {
        const int pos = atomicAdd(&s_ngood, 1);
        glist[pos] = c;
        gd2[pos] = __ldg(srow + c);
      }
    }
  } else {
    // more items with the query's id than query_lists_kernel keeps: scan the labels here; the same-camera junk goes
    // straight into the junk bits of the threads that own those columns (n_junk stays 0)
    const int32_t my_l = ev.ql[q], my_c = ev.qc[q];
    for (int k = 0; k < JR_PF; ++k) {
      const int j = 4 * (tid + (k >> 2) * JR_THREADS) + (k & 3);
      if (j < ncols && __ldg(ev.gl + j) == my_l) {
        if (__ldg(ev.gc + j) == my_c) {
          junkmask |= 1u << k;
        } else {
          const int pos = atomicAdd(&s_ngood, 1);
          if (pos < JE_GLIST) { glist[pos] = j; gd2[pos] = __ldg(srow + j); }
        }
      }
    }
  }
  __syncthreads();
  phase_done(1);
  const int total = e_off[n];
  // ---- scatter: flat (entry, position) space in warp-contiguous chunks, lanes interleaved
  {
    constexpr int NW = JR_THREADS / 32, U = 4;
    const int per_warp = ((total + NW - 1) / NW + 31) & ~31;
    const int wb = warp * per_warp, we = min(total, wb + per_warp);
    if (wb < we && !(ev.enabled & 2)) {
      int t;
      {  // largest t with e_off[t] <= first index of this lane (entries with an empty list share their successor's offset)
        const int w = min(wb + lane, we - 1);
        int lo = 0, hi = n - 1;
        while (lo < hi) {
          const int mid = (lo + hi + 1) >> 1;
          if (e_off[mid] <= w) lo = mid; else hi = mid - 1;
        }
        t = lo;
      }
      int lo_t = e_off[t], hi_t = e_off[t + 1], beg_t = e_beg[t];  // the lane's current entry, reloaded when it changes
      float val_t = e_val[t];
      for (int w0 = wb + lane; w0 < we; w0 += U * 32) {
        // NOTHING in the first loop may read a loaded value (the warp would wait there, in order, before it has issued
        // the loads of the next u: one L2 round trip per term instead of one per batch -- measured 3x on this phase);
        // indices behind the chunk's end are clamped to its last element and dropped in the second loop
        int rr[U];
        float vv[U], mv[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int w = min(w0 + u * 32, we - 1);
          if (w >= hi_t) {
            do { ++t; hi_t = e_off[t + 1]; } while (w >= hi_t);
            lo_t = e_off[t]; beg_t = e_beg[t]; val_t = e_val[t];
          }
          const int p = beg_t + (w - lo_t);
          rr[u] = __ldg(inv_row + p);
          vv[u] = __ldg(inv_val + p);
          mv[u] = val_t;
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int c = rr[u] - col0;
          if ((unsigned)c < (unsigned)ncols && w0 + u * 32 < we) atomicAdd(acci + c, __float2uint_rn(__fmul_rn(fminf(mv[u], vv[u]), JR_FIX)));
        }
      }
    }
  }
  __syncthreads();  // the accumulator is complete; the entries are dead: their shared memory is the evaluation scratch now
  phase_done(2);
  const int ngood = s_ngood;
  if (ngood == 0) {  // the cmc[0] == -1 sentinel: query skipped (evaluate.py:43-45,80-82)
    if (tid == 0) { ev.ap_q[q] = 0.0; ev.first[q] = -1; }
    return;
  }
  if (ngood > EV_GOOD_CAP) {  // poison the result so that the caller fails loudly
    if (tid == 0) { atomicExch(ev.status, 1); ev.ap_q[q] = __longlong_as_double(0x7ff8000000000000LL); ev.first[q] = -1; }
    return;
  }
  const float one_m_l = 1.0f - lambda_value;
  if (ngood <= 32) {
    // the usual case: final values and the rank-sort inside warp 0 (shuffles, no block barrier in between)
    if (warp == 0) {
      uint32_t k = 0xffffffffu;
      int c = 0x7fffffff;
      if (lane < ngood) {
        c = glist[lane];
        const unsigned int a = acci[c];
        const float jd = a ? jaccard_nonzero(__fmul_rn(__uint2float_rn(a), JR_UNFIX), 0) : 1.0f;
        k = ev_key(jr_blend_value(jd, gd2[lane], one_m_l, lambda_value), 0);
      }
      int r = 0;
#pragma unroll
      for (int f = 0; f < 32; ++f) {  // (lanes behind ngood hold the largest key and column: never "before" anything)
        const uint32_t kf = __shfl_sync(0xffffffffu, k, f);
        const int cf = __shfl_sync(0xffffffffu, c, f);
        r += ev_before(kf, cf, k, c, 0);
      }
      if (lane < ngood) { sc->skey[r] = k; sc->scol[r] = c; }
      sc->gpos[lane] = 0;
      if (lane == 0) sc->gpos[32] = 0;
    }
  } else {
  for (int e = tid; e < ngood; e += JR_THREADS) {  // final values of the good items
    const int c = glist[e];
    const unsigned int a = acci[c];
    const float jd = a ? jaccard_nonzero(__fmul_rn(__uint2float_rn(a), JR_UNFIX), 0) : 1.0f;
    sc->gcol[e] = c;
    sc->gkey[e] = ev_key(jr_blend_value(jd, gd2[e], one_m_l, lambda_value), 0);
  }
  __syncthreads();
  for (int e = tid; e < ngood; e += JR_THREADS) {  // rank-sort the goods into ranking order
    const uint32_t k = sc->gkey[e];
    const int c = sc->gcol[e];
    int r = 0;
    for (int f = 0; f < ngood; ++f) r += ev_before(sc->gkey[f], sc->gcol[f], k, c, 0);
    sc->skey[r] = k;
    sc->scol[r] = c;
  }
  for (int e = tid; e <= ngood; e += JR_THREADS) sc->gpos[e] = 0;
  }
  __syncthreads();
  phase_done(3);
  JeRow jr;
  jr.sc = sc; jr.ngood = ngood; jr.jlist = jlist; jr.n_junk = s_njunk;
  jr.one_m_l = one_m_l; jr.lambda_value = lambda_value;
  {
    const uint32_t worst_o = sc->skey[ngood - 1];
    jr.worst_v = __uint_as_float((worst_o & 0x80000000u) ? (worst_o & 0x7fffffffu) : ~worst_o);
  }
  // One branch-free superset test for every column, with or without shared neighbours (on hard data a fifth of a row has
  // some): with t = the accumulated sum, jd = 1 - t / (2 - t) >= 1 - (t + t^2) / 2 on [0, 1], so
  //   final <= w  =>  lambda * sqrt(d2) <= w - (1 - lambda) * (1 - (t + t^2) / 2)  =>  d2 <= (c0 + c1 * (t + t^2))^2
  // with c0 = (w - (1 - lambda)) / lambda and c1 = (1 - lambda) / (2 lambda), both rounded up generously (relative 1e-5,
  // absolute 4e-6 (|w| + 1): far above the few ulp the exact value can differ by; the 1.2e-12 covers the clamp of d2 at
  // 1e-12).  The columns that pass are only STAGED here (column + squared distance): their exact values -- an IEEE
  // division, a square root and a binary search each -- are computed afterwards, spread evenly over the CTA, instead
  // of by whichever warp happens to own the query's neighbourhood.
  const float w_s = __fadd_rn(jr.worst_v, __fmul_rn(__fadd_rn(fabsf(jr.worst_v), 1.0f), 4e-6f));
  const float c0 = __fmul_rn(__fdiv_rn(__fsub_rn(w_s, one_m_l), lambda_value), w_s >= one_m_l ? 1.00001f : 0.99999f);
  const float c1 = __fmul_rn(__fdiv_rn(one_m_l, __fadd_rn(lambda_value, lambda_value)), 1.00002f);
  int* cand_col = sc->gcol;                                  // (the unsorted good lists are dead after the rank-sort)
  float* cand_d2 = reinterpret_cast<float*>(sc->gkey);
  // pass 1: the 32 tests of the thread, branch-free (32 independent chains for the scheduler: with a branch per column
  // the pass ran at the latency of one chain after the other); pass 2: the few set bits
  uint32_t pass = 0;
#pragma unroll
  for (int g = 0; g < JR_PF / 4; ++g) {
    const int c0g = 4 * (tid + g * JR_THREADS);
    // (a group beyond the block reads the accumulator's first entries instead: its columns carry the junk bit)
    const uint4 a4 = *reinterpret_cast<const uint4*>(acci + (c0g < ncols ? c0g : 0));
    const unsigned int aa[4] = {a4.x, a4.y, a4.z, a4.w};
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const float t = __fmul_rn(__uint2float_rn(aa[u]), JR_UNFIX);
      const float b = fmaxf(__fmaf_rn(c1, __fmaf_rn(t, t, t), c0), 0.f);
      pass |= (pre[4 * g + u] <= __fmaf_rn(__fmul_rn(b, b), 1.0001f, 1.2e-12f) ? 1u : 0u) << (4 * g + u);
    }
  }
  pass &= ~junkmask;
  if (ev.enabled & 4) pass = 0;
  while (pass) {
    const int k = __ffs(pass) - 1;
    pass &= pass - 1;
    float d2 = 0.f;
#pragma unroll
    for (int x = 0; x < JR_PF; ++x) d2 = x == k ? pre[x] : d2;  // (a register file has no dynamic index)
    const int c = 4 * (tid + (k >> 2) * JR_THREADS) + (k & 3);
    const int pos = atomicAdd(&s_ncand, 1);
    if (pos < JE_CAND) { cand_col[pos] = c; cand_d2[pos] = d2; }
    else je_column(jr, acci[c], d2, c);
  }
  __syncthreads();
  phase_done(4);
  if ((ev.enabled & 16) && tid == 0) {  // row statistics of the timing experiments
    atomicAdd(ev.status + 16, s_ncand);
    atomicMax(ev.status + 17, s_ncand);
    if (s_ncand > JE_CAND) atomicAdd(ev.status + 18, 1);
    atomicAdd(ev.status + 19, total >> 4);
    atomicMax(ev.status + 20, total);
    atomicAdd(ev.status + 21, n);
  }
  {
    const int nc = min(s_ncand, JE_CAND);
    for (int e = tid; e < nc; e += JR_THREADS) {
      const int c = cand_col[e];
      je_column(jr, acci[c], cand_d2[e], c);
    }
  }
  __syncthreads();
  phase_done(5);
  eval_ap_tail<JR_THREADS>(sc, ngood, q, ev.ap_q, ev.first, ev.hist);
  phase_done(6);
}
// ------------------------------------------------------------------------------------ host side
struct RerankDims {
  int half, wh, cap1, cap2;
  int64_t lds;
};

static RerankDims rerank_dims(int N, int k1, int k2) {
  RerankDims d;
  d.half = (int)rint(k1 / 2.0);  // np.around: round half to even (faiss_utils.py:187)
  d.wh = d.half + 1 < k1 ? d.half + 1 : k1;
  d.cap1 = k1 + k1 * d.wh;
  d.cap2 = k2 > 1 ? k2 * d.cap1 : d.cap1;
  d.lds = (int64_t)align_up((size_t)N, 32);  // 128-byte aligned rows: every 32 x 32 output box of the GEMM covers whole cache lines
  return d;
}

struct SparseWs {
  int32_t *rfull, *rfull_cnt, *rhalf, *rhalf_cnt, *v1cols, *v1cnt, *v2cols, *v2cnt, *colcnt, *colptr, *cursor, *inv_row;
  uint32_t* junk32;  // label == -1 bits of the output columns in jaccard_eval_rows_kernel's register layout
  int32_t *qlist, *qlcnt;  // per output row (query): gallery columns with the query's id (query_lists_kernel)
  int32_t* qrank;          // place of every V_qe entry inside its column of the inverted index
  float *v1vals, *v2vals, *inv_val;
};

template <class C>
static void carve_sparse(C& c, int N, int k1, int k2, const RerankDims& d, SparseWs* w) {
  SparseWs t;
  t.rfull = c.template take<int32_t>((size_t)N * k1);
  t.rfull_cnt = c.template take<int32_t>(N);
  t.rhalf = c.template take<int32_t>((size_t)N * d.wh);
  t.rhalf_cnt = c.template take<int32_t>(N);
  t.v1cols = c.template take<int32_t>((size_t)N * d.cap1);
  t.v1vals = c.template take<float>((size_t)N * d.cap1);
  t.v1cnt = c.template take<int32_t>(N);
  if (k2 > 1) {
    t.v2cols = c.template take<int32_t>((size_t)N * d.cap2);
    t.v2vals = c.template take<float>((size_t)N * d.cap2);
    t.v2cnt = c.template take<int32_t>(N);
  } else {
    t.v2cols = t.v1cols; t.v2vals = t.v1vals; t.v2cnt = t.v1cnt;
  }
  t.colcnt = c.template take<int32_t>(N + 2);  // [N + 1] = number of rows left to the CTA-wide qexpand kernel
  t.colptr = c.template take<int32_t>(N + 1);
  t.cursor = c.template take<int32_t>(N + 1);
  t.inv_row = c.template take<int32_t>((size_t)N * d.cap2);
  t.inv_val = c.template take<float>((size_t)N * d.cap2);
  t.junk32 = c.template take<uint32_t>(JR_THREADS);
  t.qlist = c.template take<int32_t>((size_t)N * QL_STRIDE);
  t.qlcnt = c.template take<int32_t>(N);
  t.qrank = c.template take<int32_t>((size_t)N * d.cap2);
  if (w) *w = t;
}

struct NullCarver {  // size-only twin of Carver
  size_t off = 0;
  template <typename T>
  T* take(size_t count) { off = align_up(off, 256); off += count * sizeof(T); return nullptr; }
};

static int rerank_check(int N, int k1, int k2, const RerankDims& d) {
  RB_REQUIRE(k1 >= 1 && k1 <= 128 && k1 < N, "jaccard: need 1 <= k1 <= 128 and k1 < N (k1=%d, N=%d)", k1, N);
  RB_REQUIRE(k2 >= 1 && k2 <= QE_MAX_K2 && k2 <= k1, "jaccard: need 1 <= k2 <= min(k1,%d) (k2=%d)", QE_MAX_K2, k2);
  if (d.cap1 > 4096 || d.cap2 > 8192 || (size_t)((N + 3) & ~3) * 4 + (size_t)d.cap2 * 12 + 16 > 232448 - 256) {
    set_error("jaccard: size outside this build's limits (4 N + 12 cap2 <= 227 KB of shared memory: N=%d, cap1=%d <= 4096, cap2=%d <= 8192)", N, d.cap1, d.cap2);
    return REIDB200_ELIMIT;
  }
  return REIDB200_OK;
}

static size_t jaccard_rows_smem(int N, int cap) { return (size_t)((N + 3) & ~3) * 4 + (size_t)cap * 12 + 16; }

// A second stream per caller stream for work that depends on the caller's INPUTS only (the label lists of the fused
// evaluation): forked from `st` by an event, joined back by another, so it runs beside the sparse stages -- also inside a
// CUDA graph capture of `st`, where the two events become the fork and join edges of the graph.
struct SideLane {
  cudaStream_t owner, side;
  cudaEvent_t fork, join;
};

static SideLane* side_lane(cudaStream_t st) {
  static std::mutex mu;
  static std::vector<SideLane*> lanes;
  std::lock_guard<std::mutex> lock(mu);
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return nullptr;
  // (stream handles are per device, and the default stream exists on every device: lanes are keyed by (device, stream))
  static std::vector<int> lane_dev;
  for (size_t i = 0; i < lanes.size(); ++i)
    if (lanes[i]->owner == st && lane_dev[i] == dev) return lanes[i];
  SideLane* l = new SideLane();
  l->owner = st;
  if (cudaStreamCreateWithFlags(&l->side, cudaStreamNonBlocking) != cudaSuccess ||
      cudaEventCreateWithFlags(&l->fork, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&l->join, cudaEventDisableTiming) != cudaSuccess) {
    delete l;
    return nullptr;
  }
  lanes.push_back(l);
  lane_dev.push_back(dev);
  return l;
}

// `sql2` (row stride lds) is read by the lambda blend only, for the output region; `src` says where the expansion's pair
// values come from (PairSrc)
static int jaccard_sparse_stages(const int32_t* rank, const float* sql2, int64_t lds, const float* sqnorm, int N, int k1,
                                 int k2, int use_f16, float lambda_value, int row0, int nrows, int col0, int ncols,
                                 void* out, int64_t ldo, const RerankDims& d, const SparseWs& w, cudaStream_t st,
                                 const PairSrc& src, const EvalFuse* ev = nullptr) {
  static DeviceOnce attr;
  if (attr.pending()) {
    RB_CUDA(cudaFuncSetAttribute(expand_v_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 4096 * 12));
    RB_CUDA(cudaFuncSetAttribute(qexpand_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 8192 * 12));
    RB_CUDA(cudaFuncSetAttribute(inv_scan_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448 - 256));
    RB_CUDA(cudaFuncSetAttribute(jaccard_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448 - 256));
    RB_CUDA(cudaFuncSetAttribute(jaccard_eval_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448 - 256));
  }
  // evaluation only (nothing stored, fp32, blended, 128-bit aligned row slices): the kernel built for it, and its label
  // lists -- which need the labels only -- on the side lane, beside the sparse stages
  const int jr_cap = k2 > 1 ? d.cap2 : d.cap1;
  const size_t je_region = align_up(std::max((size_t)jr_cap * 12 + 16, sizeof(EvalScratch)), 16);
  const size_t je_smem = (size_t)((ncols + 3) & ~3) * 4 + je_region + (size_t)JE_GLIST * 8 + (size_t)QL_STRIDE * 4;
  const bool eval_kernel = ev && !out && !use_f16 && lambda_value > 0.f && ncols <= JR_PF * JR_THREADS && (col0 & 3) == 0 &&
                           (lds & 3) == 0 && ((reinterpret_cast<uintptr_t>(sql2) | reinterpret_cast<uintptr_t>(ev->gl)) & 15) == 0 &&
                           je_smem <= 232448 - 256 && !getenv("REIDB200_JR_OLD");
  SideLane* lane = nullptr;
  if (eval_kernel) {
    lane = getenv("REIDB200_NO_SIDE_LANE") ? nullptr : side_lane(st);
    cudaStream_t ls = st;
    if (lane) {
      RB_CUDA(cudaEventRecord(lane->fork, st));
      RB_CUDA(cudaStreamWaitEvent(lane->side, lane->fork, 0));
      ls = lane->side;
    }
    { ProfScope _ps("junk_layout_kernel", ls); junk_layout_kernel<<<1, JR_THREADS, 0, ls>>>(ev->gl, ncols, w.junk32); }
    RB_LAUNCH_CHECK("junk_layout_kernel");
    { ProfScope _ps("query_lists_kernel", ls); query_lists_kernel<<<ceil_div(nrows, QLK_Q), 256, 0, ls>>>(
        ev->ql, ev->qc, ev->gl, ev->gc, nrows, ncols, w.qlist, w.qlcnt); }
    RB_LAUNCH_CHECK("query_lists_kernel");
    if (lane) RB_CUDA(cudaEventRecord(lane->join, ls));
  }
  const int wpb = 8;
  { ProfScope _ps("krecip_kernel", st); krecip_both_kernel<<<ceil_div(N, wpb), wpb * 32, 0, st>>>(rank, N, k1, k1, d.wh, w.rfull, w.rfull_cnt, w.rhalf, w.rhalf_cnt); }
  RB_LAUNCH_CHECK("krecip_both_kernel");
  if (k1 <= 32 && k1 + k1 * d.wh <= 256) {
    { ProfScope _ps("expand_v_kernel", st); expand_v_warp_kernel<<<ceil_div(N, EXW_WARPS), EXW_WARPS * 32, 0, st>>>(
        w.rfull, w.rfull_cnt, k1, w.rhalf, w.rhalf_cnt, d.wh, src, sqnorm, N, d.cap1, use_f16, w.v1cols, w.v1vals, w.v1cnt); }
  } else {
    int expP = 32;
    while (expP < d.cap1) expP <<= 1;
    { ProfScope _ps("expand_v_kernel", st); expand_v_kernel<<<N, EXP_THREADS, (size_t)expP * 12, st>>>(w.rfull, w.rfull_cnt, k1, w.rhalf, w.rhalf_cnt, d.wh, src,
                                                                  sqnorm, N, d.cap1, expP, use_f16, w.v1cols, w.v1vals,
                                                                  w.v1cnt); }
  }
  RB_LAUNCH_CHECK("expand_v_kernel");
  RB_CUDA(cudaMemsetAsync(w.colcnt, 0, (size_t)(N + 2) * sizeof(int32_t), st));
  if (k2 > 1) {
    // D5 + the column counts of D6: warp kernel first, the CTA kernel picks up the rows it marked as too long
    { ProfScope _ps("qexpand_kernel", st); qexpand_warp_kernel<<<ceil_div(N, QEW_WARPS), QEW_WARPS * 32, 0, st>>>(
        rank, k1, k2, w.v1cols, w.v1vals, w.v1cnt, N, d.cap1, d.cap2, use_f16, w.v2cols, w.v2vals, w.v2cnt, w.colcnt,
        w.qrank, w.cursor, w.colcnt + N + 1); }  // the cursor array is free until inv_scan_kernel: it holds the long-row list
    RB_LAUNCH_CHECK("qexpand_warp_kernel");
    { ProfScope _ps("qexpand_kernel(long rows)", st); qexpand_kernel<<<min(N, 2 * num_sms()), QE_THREADS, (size_t)d.cap2 * 12, st>>>(
        rank, k1, k2, w.v1cols, w.v1vals, w.v1cnt, d.cap1, d.cap2, use_f16, w.v2cols, w.v2vals, w.v2cnt, w.cursor, w.colcnt + N + 1, N,
        w.colcnt, w.qrank); }
    RB_LAUNCH_CHECK("qexpand_kernel");
  } else {
    { ProfScope _ps("inv_count_kernel", st); inv_count_kernel<<<N, 128, 0, st>>>(w.v2cols, w.v2vals, w.v2cnt, N, d.cap2, w.colcnt, w.qrank); }
    RB_LAUNCH_CHECK("inv_count_kernel");
  }
  { ProfScope _ps("inv_scan_kernel", st); inv_scan_kernel<<<1, 1024, (size_t)N * sizeof(int), st>>>(w.colcnt, N, w.colptr, w.cursor); }
  RB_LAUNCH_CHECK("inv_scan_kernel");
  { ProfScope _ps("inv_fill_kernel", st); inv_fill_kernel<<<ceil_div(N, IF_WARPS), IF_WARPS * 32, 0, st>>>(w.v2cols, w.v2vals, w.v2cnt, w.qrank, N, d.cap2, w.colptr, w.inv_row, w.inv_val); }
  RB_LAUNCH_CHECK("inv_fill_kernel");
  EvalFuse evf;
  memset(&evf, 0, sizeof(evf));
  size_t jr_smem = jaccard_rows_smem(N, jr_cap);
  if (ev) {
    evf = *ev;
    evf.enabled = 1;
    if (jr_smem < sizeof(EvalScratch) + 16) jr_smem = sizeof(EvalScratch) + 16;
  }
  if (eval_kernel) {
    if (const char* dbg = getenv("REIDB200_JE_DEBUG")) evf.enabled |= atoi(dbg) & ~1;  // timing experiments only (wrong results)
    if (lane) RB_CUDA(cudaStreamWaitEvent(st, lane->join, 0));
    JeLists jl;
    jl.qlist = w.qlist; jl.qlcnt = w.qlcnt; jl.junk32 = w.junk32;
    { ProfScope _ps("jaccard_rows_kernel", st); jaccard_eval_rows_kernel<<<nrows, JR_THREADS, je_smem, st>>>(
        w.v2cols, w.v2vals, w.v2cnt, d.cap2, w.colptr, w.inv_row, w.inv_val, row0, col0, ncols, lambda_value, sql2, lds,
        (int)je_region, jl, evf); }
    RB_LAUNCH_CHECK("jaccard_eval_rows_kernel");
    return REIDB200_OK;
  }
  { ProfScope _ps("jaccard_rows_kernel", st); jaccard_rows_kernel<<<nrows, JR_THREADS, jr_smem, st>>>(
      w.v2cols, w.v2vals, w.v2cnt, d.cap2, w.colptr, w.inv_row, w.inv_val, N, row0, col0, ncols, lambda_value, sql2, lds,
      use_f16, out, ldo, evf); }
  RB_LAUNCH_CHECK("jaccard_rows_kernel");
  return REIDB200_OK;
}

}  // namespace reidb200

using namespace reidb200;

// Workspace of the k-NN + sparse stages of one merged set.  The squared-L2 matrix is NOT materialised: the initial ranking
// is selected inside the GEMM epilogue (gemm.cuh SelArgs) and only the rows the lambda blend reads (`blend_rows` of them,
// 0 when lambda == 0) are stored, in a block of whole 32-row groups.
struct JacWs {
  char* packed; float* sqn; float* isc;
  float* blk;            // selection mode: [blk_rows, lds] rows of the squared-L2 matrix kept for the blend (or null)
  float* mat; float* cred;  // matrix mode: the whole [N, lds] squared-L2 matrix + its per-(row, 32-column chunk) minima
  uint32_t* cred2;          // matrix mode: second word per chunk of the triangle-mode GEMM (gemm.cuh gemm_split_tri_launch)
  float* rank_d; int32_t* rank_i;
  char* selws; size_t selbytes;
  SparseWs w;
};

// Two ways to get the initial ranking (D2):
//   matrix mode     the N x N squared-L2 matrix is written by the GEMM (with per-chunk minima) and the chunk-guided row
//                   select reads ~k chunks per row.  Fastest up to a few 10^4 rows: at N = 19281 the GEMM with its stores
//                   takes 0.47 ms + 0.095 ms of selection.
//   selection mode  the k-NN is selected INSIDE the GEMM epilogue (gemm.cuh SelArgs): nothing but the rows the lambda blend
//                   reads is stored.  Measured at N = 19281: 0.40 ms of GEMM + 0.2 ms for the bootstrap wave (every row's
//                   first tile fills its list and sketch while the other CTAs wait at the gates) + the appends of ~0.45 %
//                   of all elements through L2 atomics -- slower than matrix mode there, but the only way once N x N
//                   floats stop fitting (N = 150 k: 90 GB) and cheaper per element as N grows (the acceptance rate falls
//                   like k ln N / N).  Chosen from N; REIDB200_RERANK_SELECT=0/1 forces it (tests run both).
static bool rerank_select_mode(int N) {
  if (const char* e = getenv("REIDB200_RERANK_SELECT")) return atoi(e) != 0;
  return N > 32768;
}

static int blend_block_rows(int blend_rows, int N) {  // rows of whole 32-row groups that cover any `blend_rows` consecutive rows
  if (blend_rows <= 0) return 0;
  const int r = ((blend_rows + 31) & ~31) + 32;
  return r < ((N + 31) & ~31) ? r : ((N + 31) & ~31);
}

template <class C>
static void carve_jaccard(C& c, int N, int d, int k1, int k2, const RerankDims& rd, int blend_rows, JacWs* out) {
  JacWs t;
  t.packed = c.template take<char>(packed_bytes(N, d));
  t.sqn = c.template take<float>(N);
  t.isc = c.template take<float>(N);
  t.blk = t.mat = t.cred = nullptr;
  t.cred2 = nullptr;
  t.selws = nullptr; t.selbytes = 0;
  if (rerank_select_mode(N)) {
    const int br = blend_block_rows(blend_rows, N);
    t.blk = br ? c.template take<float>((size_t)br * rd.lds) : nullptr;
    t.selbytes = sel_workspace_bytes(N, N, k1);
    t.selws = c.template take<char>(t.selbytes);
  } else {
    t.mat = c.template take<float>((size_t)N * rd.lds);
    t.cred = c.template take<float>((size_t)N * ceil_div((int)rd.lds, 32));
    t.cred2 = c.template take<uint32_t>((size_t)N * ceil_div((int)rd.lds, 32));
  }
  t.rank_d = c.template take<float>((size_t)N * k1);
  t.rank_i = c.template take<int32_t>((size_t)N * k1);
  carve_sparse(c, N, k1, k2, rd, &t.w);
  if (out) *out = t;
}

extern "C" size_t reidb200_jaccard_region_workspace_bytes(int N, int d, int k1, int k2, int blend_rows) {
  if (N <= 0 || d <= 0 || k1 <= 0 || k2 <= 0) return 0;
  const RerankDims rd = rerank_dims(N, k1, k2);
  NullCarver c;
  carve_jaccard(c, N, d, k1, k2, rd, blend_rows, nullptr);
  return align_up(c.off, 256);
}

extern "C" size_t reidb200_jaccard_workspace_bytes(int N, int d, int k1, int k2) {
  return reidb200_jaccard_region_workspace_bytes(N, d, k1, k2, N);  // any region, any lambda
}

static int region_check(int N, int row0, int nrows, int col0, int ncols, int64_t ldo) {
  RB_REQUIRE(row0 >= 0 && nrows >= 0 && row0 + nrows <= N && col0 >= 0 && ncols >= 0 && col0 + ncols <= N,
             "jaccard: output region [%d,+%d) x [%d,+%d) outside N=%d", row0, nrows, col0, ncols, N);
  RB_REQUIRE(ldo >= ncols, "jaccard: ldo too small");
  return REIDB200_OK;
}

extern "C" int reidb200_jaccard_from_rank(const int32_t* rank, const float* sql2, int64_t lds, const float* sqnorm, int N,
                                          int k1, int k2, int use_float16, float lambda_value, int row0, int nrows,
                                          int col0, int ncols, void* out, int64_t ldo, void* workspace,
                                          size_t workspace_bytes, void* stream) {
  RB_REQUIRE(rank && sql2 && sqnorm && out && N > 0 && lds >= N, "jaccard_from_rank: bad argument");
  const RerankDims rd = rerank_dims(N, k1, k2);
  int rc = rerank_check(N, k1, k2, rd);
  if (rc) return rc;
  rc = region_check(N, row0, nrows, col0, ncols, ldo);
  if (rc) return rc;
  if (nrows == 0 || ncols == 0) return REIDB200_OK;
  NullCarver nc;
  carve_sparse(nc, N, k1, k2, rd, nullptr);
  RB_REQUIRE(workspace && workspace_bytes >= nc.off, "jaccard_from_rank: workspace too small");
  Carver c(workspace, workspace_bytes);
  SparseWs w;
  carve_sparse(c, N, k1, k2, rd, &w);
  PairSrc src;
  memset(&src, 0, sizeof(src));
  src.sql2 = sql2; src.lds = lds;
  return jaccard_sparse_stages(rank, sql2, lds, sqnorm, N, k1, k2, use_float16, lambda_value, row0, nrows, col0, ncols,
                               out, ldo, rd, w, (cudaStream_t)stream, src);
}

// D2 for one merged set without the N x N matrix: pack, GEMM with the selection in its epilogue (the rows
// [blend_row0, blend_row0 + blend_rows) x all columns of the squared-L2 matrix are still stored for the lambda blend, into
// ws.blk), k best entries of every list.  Returns the VIRTUAL base of the stored rows (row i at base + i * lds) or null.
static int knn_stage(const float* feats, int N, int64_t ldf, int d, int k1, const RerankDims& rd, int blend_row0, int blend_rows,
                     int blend_col0, int blend_cols, const JacWs& ws, const float** sql2_virtual, bool* have_matrix, cudaStream_t st,
                     bool* tri_out = nullptr) {
  int rc = pack_rows_launch(feats, N, ldf, d, ws.packed, ws.sqn, ws.isc, st);
  if (rc) return rc;
  if (ws.mat) {  // matrix mode
    const int ldcr = ceil_div((int)rd.lds, 32);
    // Triangle mode: every element stored once (half of the GEMM's output bytes).  Only for a caller that can live with
    // it (tri_out != null: its expansion reads single elements through PairSrc::tri) and whose blend block is stored in
    // place: rows whose tile block is not below the block of any of its columns.
    // REIDB200_TRI=0 keeps the full matrix (A/B timing, tests); REIDB200_TRI=2 makes a call that cannot use it fail (tests).
    const int tri_env = getenv("REIDB200_TRI") ? atoi(getenv("REIDB200_TRI")) : 1;
    const bool tri = tri_out && tri_env != 0 && tri_mode_ok(N, k1) && !getenv("REIDB200_GEMM_NOSYM") && !getenv("REIDB200_GEMM_1CTA") &&
                     (blend_rows <= 0 || blend_cols <= 0 || tri_stored(blend_row0, blend_col0 + blend_cols - 1));
    RB_REQUIRE(tri || !tri_out || tri_env != 2, "rerank: REIDB200_TRI=2 but N=%d, k1=%d cannot run in triangle mode", N, k1);
    if (tri_out) *tri_out = tri;
    if (tri) {
      rc = gemm_split_tri_launch(ws.packed, ws.sqn, ws.isc, N, d, ws.mat, rd.lds, st, ws.cred, ws.cred2, ldcr);
      if (rc) return rc;
      *sql2_virtual = ws.mat;
      *have_matrix = true;
      return topk_rows_launch(ws.mat, N, N, rd.lds, nullptr, k1, 0, 0, ws.rank_d, ws.rank_i, 0, st, ws.cred, ldcr, ws.cred2);
    }
    rc = gemm_split_launch(ws.packed, ws.sqn, ws.isc, N, ws.packed, ws.sqn, ws.isc, N, d, REIDB200_DIST_SQL2, ws.mat, rd.lds, st, ws.cred, ldcr, 0, 0);
    if (rc) return rc;
    *sql2_virtual = ws.mat;
    *have_matrix = true;
    return topk_rows_launch(ws.mat, N, N, rd.lds, nullptr, k1, 0, 0, ws.rank_d, ws.rank_i, 0, st, ws.cred, ldcr);
  }
  SelArgs sel;
  float* redo_scratch = nullptr;
  rc = sel_prepare(ws.selws, ws.selbytes, N, N, k1, 0, &sel, &redo_scratch, st);
  if (rc) return rc;
  float* vbase = nullptr;
  if (blend_rows > 0) {
    const int r0a = blend_row0 & ~31;  // the epilogue stores whole 32 x 32 boxes
    RB_REQUIRE(ws.blk && ((blend_row0 + blend_rows + 31) & ~31) - r0a <= blend_block_rows(blend_rows, N), "jaccard: blend block too small");
    vbase = ws.blk - (ptrdiff_t)r0a * rd.lds;
    sel.st_r0 = blend_row0; sel.st_r1 = blend_row0 + blend_rows; sel.st_c0 = blend_col0; sel.st_c1 = blend_col0 + blend_cols;
  }
  rc = gemm_split_select_launch(ws.packed, ws.sqn, ws.isc, N, ws.packed, ws.sqn, ws.isc, N, d, REIDB200_DIST_SQL2, vbase, rd.lds, sel, st);
  if (rc) return rc;
  *sql2_virtual = vbase;
  *have_matrix = false;
  return knn_select_finish(sel, N, N, k1, REIDB200_DIST_SQL2, ws.packed, ws.sqn, ws.isc, ws.packed, ws.sqn, ws.isc, d, 0, ws.rank_d,
                           ws.rank_i, 0, redo_scratch, st);
}

extern "C" int reidb200_jaccard(const float* feats, int N, int64_t ldf, int d, int k1, int k2, int use_float16,
                                float lambda_value, int row0, int nrows, int col0, int ncols, void* out, int64_t ldo,
                                int32_t* rank_out, void* workspace, size_t workspace_bytes, void* stream) {
  RB_REQUIRE(feats && out && N > 0 && d > 0 && ldf >= d, "jaccard: bad argument");
  const RerankDims rd = rerank_dims(N, k1, k2);
  int rc = rerank_check(N, k1, k2, rd);
  if (rc) return rc;
  rc = region_check(N, row0, nrows, col0, ncols, ldo);
  if (rc) return rc;
  const int blend_rows = lambda_value > 0.f ? nrows : 0;
  RB_REQUIRE(workspace && workspace_bytes >= reidb200_jaccard_region_workspace_bytes(N, d, k1, k2, blend_rows), "jaccard: workspace too small");
  RB_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, "jaccard: workspace must be 256-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  Carver c(workspace, workspace_bytes);
  JacWs ws;
  carve_jaccard(c, N, d, k1, k2, rd, blend_rows, &ws);
  const float* sql2v = nullptr;
  bool have_matrix = false;
  rc = knn_stage(feats, N, ldf, d, k1, rd, row0, blend_rows, col0, ncols, ws, &sql2v, &have_matrix, st);
  if (rc) return rc;
  if (rank_out) RB_CUDA(cudaMemcpyAsync(rank_out, ws.rank_i, (size_t)N * k1 * sizeof(int32_t), cudaMemcpyDeviceToDevice, st));
  if (nrows == 0 || ncols == 0) return REIDB200_OK;
  PairSrc src;
  memset(&src, 0, sizeof(src));
  if (have_matrix) { src.sql2 = sql2v; src.lds = rd.lds; }
  else { src.feats = feats; src.ldf = ldf; src.d = d; src.rank_i = ws.rank_i; src.rank_d = ws.rank_d; src.k1c = k1; }
  return jaccard_sparse_stages(ws.rank_i, sql2v, rd.lds, ws.sqn, N, k1, k2, use_float16, lambda_value, row0, nrows, col0, ncols,
                               out, ldo, rd, ws.w, st, src);
}

// ------------------------------------------------------------------------------------ row-sharded re-ranking
// Stages D3-D8 for a block of OUTPUT ROWS from a given global initial ranking: what one rank of a multi-GPU re-ranking
// runs after the k-NN tables of all ranks have been all-gathered.  No squared-L2 matrix of the merged set is needed: pair
// values of the expansion come from the rows' top-k1 lists (rank_d) or a direct dot product of the feature rows; the
// lambda blend reads `blend_sql2`, the squared L2 of just this block (rows [row0, +nrows) x columns [col0, +ncols)).
__global__ void __launch_bounds__(256)
row_sqnorm_kernel(const float* __restrict__ x, int rows, int64_t ldx, int d, float* __restrict__ sqnorm) {
  const int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (r >= rows) return;
  float s = 0.f;
  for (int k = lane; k < d; k += 32) { const float v = __ldg(x + (size_t)r * ldx + k); s = fmaf(v, v, s); }  // pack_rows_kernel's order
  s = warp_sum(s);
  if (lane == 0) sqnorm[r] = s;
}

extern "C" size_t reidb200_rerank_rows_workspace_bytes(int N, int d, int k1, int k2) {
  if (N <= 0 || d <= 0 || k1 <= 0 || k2 <= 0) return 0;
  const RerankDims rd = rerank_dims(N, k1, k2);
  NullCarver c;
  c.take<float>(N);
  carve_sparse(c, N, k1, k2, rd, nullptr);
  return align_up(c.off, 256);
}

extern "C" int reidb200_rerank_rows_from_rank(const float* feats, int N, int64_t ldf, int d, const int32_t* rank_i,
                                              const float* rank_d, int k1, int k2, int use_float16, float lambda_value,
                                              int row0, int nrows, int col0, int ncols, const float* blend_sql2, int64_t ldb,
                                              void* out, int64_t ldo, const int32_t* ql, const int32_t* qc, const int32_t* gl,
                                              const int32_t* gc, int32_t* hist, double* ap_q, int32_t* first, int32_t* status,
                                              void* workspace, size_t workspace_bytes, void* stream) {
  RB_REQUIRE(feats && rank_i && rank_d && N > 0 && d > 0 && ldf >= d, "rerank_rows_from_rank: bad argument");
  const RerankDims rd = rerank_dims(N, k1, k2);
  int rc = rerank_check(N, k1, k2, rd);
  if (rc) return rc;
  rc = region_check(N, row0, nrows, col0, ncols, out ? ldo : ncols);
  if (rc) return rc;
  const bool eval = hist != nullptr;
  RB_REQUIRE(out || eval, "rerank_rows_from_rank: neither an output block nor an evaluation requested");
  RB_REQUIRE(!eval || (ql && qc && gl && gc && ap_q && first && status && !use_float16 && ncols <= JR_PF * JR_THREADS),
             "rerank_rows_from_rank: the fused evaluation needs labels, fp32 and at most %d columns", JR_PF * JR_THREADS);
  RB_REQUIRE(lambda_value <= 0.f || (blend_sql2 && ldb >= ncols), "rerank_rows_from_rank: the lambda blend needs the block's squared L2");
  RB_REQUIRE(workspace && workspace_bytes >= reidb200_rerank_rows_workspace_bytes(N, d, k1, k2), "rerank_rows_from_rank: workspace too small");
  if (nrows == 0 || ncols == 0) return REIDB200_OK;
  cudaStream_t st = (cudaStream_t)stream;
  Carver c(workspace, workspace_bytes);
  float* sqn = c.take<float>(N);
  SparseWs w;
  carve_sparse(c, N, k1, k2, rd, &w);
  { ProfScope _ps("row_sqnorm_kernel", st); row_sqnorm_kernel<<<ceil_div(N, 8), 256, 0, st>>>(feats, N, ldf, d, sqn); }
  RB_LAUNCH_CHECK("row_sqnorm_kernel");
  PairSrc src;
  memset(&src, 0, sizeof(src));
  src.feats = feats; src.ldf = ldf; src.d = d; src.rank_i = rank_i; src.rank_d = rank_d; src.k1c = k1;
  // jaccard_rows reads blend[i * lds + col0 + c] with the ABSOLUTE row i: hand it the virtual base of the block
  const float* vbase = blend_sql2 ? blend_sql2 - (ptrdiff_t)row0 * ldb - col0 : nullptr;
  EvalFuse ev;
  memset(&ev, 0, sizeof(ev));
  if (eval) { ev.ql = ql; ev.qc = qc; ev.gl = gl; ev.gc = gc; ev.hist = hist; ev.ap_q = ap_q; ev.first = first; ev.status = status; }
  return jaccard_sparse_stages(rank_i, vbase, ldb, sqn, N, k1, k2, use_float16, lambda_value, row0, nrows, col0, ncols, out, ldo, rd,
                               w, st, src, eval ? &ev : nullptr);
}

extern "C" int reidb200_k_reciprocal(const int32_t* rank, int N, int k1_cols, int k, int32_t* out_idx, int32_t* out_cnt,
                                     void* stream) {
  RB_REQUIRE(rank && out_idx && out_cnt && N > 0 && k1_cols > 0 && k > 0, "k_reciprocal: bad argument");
  const int kc = k + 1 < k1_cols ? k + 1 : k1_cols;
  { ProfScope _ps("krecip_kernel", (cudaStream_t)stream); krecip_kernel<<<ceil_div(N, 8), 256, 0, (cudaStream_t)stream>>>(rank, N, k1_cols, kc, k + 1, out_idx, out_cnt); }
  RB_LAUNCH_CHECK("krecip_kernel");
  return REIDB200_OK;
}

// Re-ranked evaluation in one call: merged = [gallery (G rows); query (Q rows)], Jaccard re-ranking blended with the
// Euclidean distance on the Q x G block, CMC / mAP reduced by the CTA that produces each query row (EvalFuse).
// out_dist may be null: the re-ranked matrix is then never written.
extern "C" size_t reidb200_rerank_eval_workspace_bytes(int N, int d, int k1, int k2, int Q, int G) {
  if (N <= 0 || d <= 0 || k1 <= 0 || k2 <= 0 || Q <= 0 || G <= 0) return 0;
  CarveSize c;
  c.take<char>(reidb200_jaccard_region_workspace_bytes(N, d, k1, k2, Q));
  c.take<int32_t>(G);   // first-hit histogram
  c.take<double>(Q);    // per-query AP
  c.take<int32_t>(Q);   // first hit
  c.take<int32_t>(64);  // status
  return align_up(c.off, 256);
}

static int rerank_eval_impl(const float* feats, int G, int Q, int64_t ldf, int d, int k1, int k2, float lambda_value,
                            const int32_t* ql, const int32_t* qc, const int32_t* gl, const int32_t* gc,
                            float* out_dist, int64_t ldo, float* out_cmc, double* out_ap, double* out_ap_q,
                            int32_t* out_first, void* workspace, size_t workspace_bytes, void* stream) {
  RB_REQUIRE(feats && ql && qc && gl && gc && out_cmc && out_ap && G > 0 && Q > 0 && d > 0 && ldf >= d,
             "rerank_eval: bad argument");
  RB_REQUIRE(G <= JR_PF * JR_THREADS, "rerank_eval: the fused evaluation holds at most %d gallery columns per query row (G=%d); "
             "use reidb200_jaccard + reidb200_eval_cmc_map", JR_PF * JR_THREADS, G);
  RB_REQUIRE(out_dist == nullptr || ldo >= G, "rerank_eval: ldo too small");
  const int N = G + Q;
  RB_REQUIRE(workspace && workspace_bytes >= reidb200_rerank_eval_workspace_bytes(N, d, k1, k2, Q, G), "rerank_eval: workspace too small");
  RB_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, "rerank_eval: workspace must be 256-byte aligned");
  const RerankDims rd = rerank_dims(N, k1, k2);
  int rc = rerank_check(N, k1, k2, rd);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  const int blend_rows = lambda_value > 0.f ? Q : 0;
  Carver c0(workspace, workspace_bytes);
  const size_t jbytes = reidb200_jaccard_region_workspace_bytes(N, d, k1, k2, Q);
  char* jws = c0.take<char>(jbytes);
  EvalFuse ev;
  ev.ql = ql; ev.qc = qc; ev.gl = gl; ev.gc = gc;
  ev.hist = c0.take<int32_t>(G);
  ev.ap_q = c0.take<double>(Q);
  ev.first = c0.take<int32_t>(Q);
  ev.status = c0.take<int32_t>(64);
  ev.enabled = 1;
  if (out_ap_q) ev.ap_q = out_ap_q;
  if (out_first) ev.first = out_first;
  RB_CUDA(cudaMemsetAsync(ev.hist, 0, (size_t)G * sizeof(int32_t), st));
  RB_CUDA(cudaMemsetAsync(ev.status, 0, 64 * sizeof(int32_t), st));
  Carver c(jws, jbytes);
  JacWs ws;
  carve_jaccard(c, N, d, k1, k2, rd, Q, &ws);
  const float* sql2v = nullptr;
  bool have_matrix = false;
  bool tri = false;  // the query rows come LAST in the merged set, so their gallery columns are stored in place in triangle mode
  rc = knn_stage(feats, N, ldf, d, k1, rd, G, blend_rows, 0, G, ws, &sql2v, &have_matrix, st, &tri);  // selection mode stores the query x gallery block only
  if (rc) return rc;
  PairSrc src;
  memset(&src, 0, sizeof(src));
  if (have_matrix) { src.sql2 = sql2v; src.lds = rd.lds; src.tri = tri ? 1 : 0; }
  else { src.feats = feats; src.ldf = ldf; src.d = d; src.rank_i = ws.rank_i; src.rank_d = ws.rank_d; src.k1c = k1; }
  rc = jaccard_sparse_stages(ws.rank_i, sql2v, rd.lds, ws.sqn, N, k1, k2, 0, lambda_value, G, Q, 0, G, out_dist, ldo, rd, ws.w, st, src, &ev);
  if (rc) return rc;
  return eval_finalize_launch(ev.hist, ev.ap_q, Q, G, out_cmc, out_ap, st);
}

extern "C" int reidb200_rerank_eval(const float* feats, int G, int Q, int64_t ldf, int d, int k1, int k2, float lambda_value,
                                    const int32_t* ql, const int32_t* qc, const int32_t* gl, const int32_t* gc,
                                    float* out_dist, int64_t ldo, float* out_cmc, double* out_ap, void* workspace,
                                    size_t workspace_bytes, void* stream) {
  return rerank_eval_impl(feats, G, Q, ldf, d, k1, k2, lambda_value, ql, qc, gl, gc, out_dist, ldo, out_cmc, out_ap, nullptr,
                          nullptr, workspace, workspace_bytes, stream);
}

extern "C" int reidb200_rerank_eval_pq(const float* feats, int G, int Q, int64_t ldf, int d, int k1, int k2, float lambda_value,
                                       const int32_t* ql, const int32_t* qc, const int32_t* gl, const int32_t* gc,
                                       float* out_dist, int64_t ldo, float* out_cmc, double* out_ap, double* out_ap_q,
                                       int32_t* out_first, void* workspace, size_t workspace_bytes, void* stream) {
  return rerank_eval_impl(feats, G, Q, ldf, d, k1, k2, lambda_value, ql, qc, gl, gc, out_dist, ldo, out_cmc, out_ap, out_ap_q,
                          out_first, workspace, workspace_bytes, stream);
}
