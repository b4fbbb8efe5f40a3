// Shared pieces of the CMC / mAP reduction (reid/evaluate.py:33-105): ranking keys, the tie order of the reference, and
// the per-query evaluation of a score row that a CTA already holds in registers (used by the fused re-rank + evaluate
// kernel in rerank.cu; evalmap.cu's eval_query_kernel streams the row from memory with the same logic).
#pragma once
#include "common.cuh"

namespace reidb200 {

constexpr int EV_GOOD_CAP = 1024;

__device__ __forceinline__ uint32_t ev_key(float v, int descending) {
  const uint32_t u = float_to_ordered(v);
  return descending ? ~u : u;  // smaller key = ranked earlier
}

// Tie order of the reference: ascending argsort then [::-1] for scores (higher index first), plain ascending
// (lower index first) for distances.
__device__ __forceinline__ bool ev_before(uint32_t ka, int ja, uint32_t kb, int jb, int descending) {
  return ka < kb || (ka == kb && (descending ? ja > jb : ja < jb));
}

// shared-memory scratch of one evaluation (28.3 KB); may alias memory the caller no longer needs
struct EvalScratch {
  int32_t gcol[EV_GOOD_CAP], scol[EV_GOOD_CAP];
  uint32_t gkey[EV_GOOD_CAP], skey[EV_GOOD_CAP];
  int32_t gpos[EV_GOOD_CAP + 1];
  int32_t ngood;
  double terms[EV_GOOD_CAP];
};

// The two rare paths of the register-resident evaluation, kept out of line: the loops over the PF register values are
// fully unrolled (static register indices) and would otherwise carry PF copies of these bodies (instruction-cache
// misses were the top stall of the first fused version).
static __device__ __noinline__ void ev_add_good(EvalScratch* sc, int j, uint32_t key) {
  const int pos = atomicAdd(&sc->ngood, 1);
  if (pos < EV_GOOD_CAP) { sc->gcol[pos] = j; sc->gkey[pos] = key; }
}

static __device__ __noinline__ void ev_count_before(EvalScratch* sc, int ngood, uint32_t kj, int j, int descending, int32_t my_l,
                                             int32_t my_c, const int32_t* __restrict__ gl, const int32_t* __restrict__ gc) {
  if (!ev_before(kj, j, sc->skey[ngood - 1], sc->scol[ngood - 1], descending)) return;
  const int32_t l = __ldg(gl + j);
  if (l == -1 || (l == my_l && __ldg(gc + j) == my_c)) return;  // junk (evaluate.py:71-72)
  int lo = 0, hi = ngood - 1;  // smallest s with j before good s (exists: j is before the worst good)
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (ev_before(kj, j, sc->skey[mid], sc->scol[mid], descending)) hi = mid; else lo = mid + 1;
  }
  atomicAdd(&sc->gpos[lo], 1);
}

// Which column of the row sits in register k of thread tid.  VEC == 1: column tid + k * NT (consecutive threads hold
// consecutive columns).  VEC == 4: the thread holds groups of four consecutive columns, group g = k / 4 starting at
// 4 * (tid + g * NT) -- the row, the labels and the accumulator are then read with 128-bit accesses.
template <int NT, int VEC>
__device__ __forceinline__ int ev_col(int tid, int k) {
  return VEC == 1 ? tid + k * NT : VEC * (tid + (k / VEC) * NT) + (k % VEC);
}

// Label pass of the register-resident evaluation, separated so that the caller can run it EARLY (it needs the labels
// only, not the row values: in the fused re-rank kernel it sits under the latency of the kernel's first global loads
// instead of adding four dependent L2 round trips to the tail).  Bit k of the result = column tid + k * NT is a good
// item of query q (same id, different camera: evaluate.py:66-70).
template <int NT, int PF, int VEC = 1>
__device__ __forceinline__ uint32_t eval_good_mask(int G, int q, const int32_t* __restrict__ ql, const int32_t* __restrict__ qc,
                                                   const int32_t* __restrict__ gl, const int32_t* __restrict__ gc) {
  static_assert(PF <= 32, "one mask bit per register value");
  static_assert(VEC == 1 || (VEC == 4 && PF % 4 == 0), "VEC is 1 or 4");
  const int tid = threadIdx.x;
  const int32_t my_l = ql[q], my_c = qc[q];
  uint32_t mask = 0;
  if (VEC == 4) {  // gl must be 16-byte aligned (the caller checks)
    constexpr int GB = 2;  // groups per batch: the caller's row values already fill half of the register file
#pragma unroll
    for (int g0 = 0; g0 < PF / 4; g0 += GB) {
      int4 lab[GB];
#pragma unroll
      for (int u = 0; u < GB; ++u) {
        const int j0 = 4 * (tid + (g0 + u) * NT);
        if (j0 + 3 < G) {
          lab[u] = __ldg(reinterpret_cast<const int4*>(gl + j0));
        } else {  // the group that straddles the end of the label vector
          lab[u].x = j0 < G ? __ldg(gl + j0) : -2;
          lab[u].y = j0 + 1 < G ? __ldg(gl + j0 + 1) : -2;
          lab[u].z = j0 + 2 < G ? __ldg(gl + j0 + 2) : -2;
          lab[u].w = -2;
        }
      }
#pragma unroll
      for (int u = 0; u < GB; ++u) {
        const int g = g0 + u;
        const int j0 = 4 * (tid + g * NT);
        if (lab[u].x == my_l && j0 < G && __ldg(gc + j0) != my_c) mask |= 1u << (4 * g);
        if (lab[u].y == my_l && j0 + 1 < G && __ldg(gc + j0 + 1) != my_c) mask |= 1u << (4 * g + 1);
        if (lab[u].z == my_l && j0 + 2 < G && __ldg(gc + j0 + 2) != my_c) mask |= 1u << (4 * g + 2);
        if (lab[u].w == my_l && j0 + 3 < G && __ldg(gc + j0 + 3) != my_c) mask |= 1u << (4 * g + 3);
      }
    }
    return mask;
  }
#pragma unroll
  for (int k0 = 0; k0 < PF; k0 += 8) {  // labels in batches of eight loads (the label vectors are L2 resident)
    int32_t lab[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int j = tid + (k0 + u) * NT;
      lab[u] = (k0 + u < PF && j < G) ? __ldg(gl + j) : -2;
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int j = tid + (k0 + u) * NT;
      if (k0 + u < PF && j < G && lab[u] == my_l && __ldg(gc + j) != my_c) mask |= 1u << (k0 + u);
    }
  }
  return mask;
}

// Tail of one query's evaluation, after the counting pass: sc->gpos[s] = number of non-junk items that precede good item
// s but not good item s - 1 (all NT threads call it, right after the barrier that ends the counting pass).
template <int NT>
__device__ __forceinline__ void eval_ap_tail(EvalScratch* sc, int ngood, int q, double* __restrict__ ap_q,
                                             int32_t* __restrict__ first, int32_t* __restrict__ hist) {
  const int tid = threadIdx.x;
  if (ngood <= 32) {
    // the usual case (a Market query has ~17 good items): prefix, AP terms and the ordered sum stay inside warp 0 --
    // shuffles instead of two more block barriers and two serial loops through shared memory; same operations in
    // the same order as the general path below
    if (tid < 32) {
      int pe = tid < ngood ? sc->gpos[tid] : 0;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {  // inclusive prefix: position of good s among the non-junk items
        const int up = __shfl_up_sync(0xffffffffu, pe, o);
        if (tid >= o) pe += up;
      }
      const double d_recall = 1.0 / (double)ngood;
      double term = 0.0;
      if (tid < ngood) {
        const double precision = (double)(tid + 1) * 1.0 / (double)(pe + 1);
        const double old_precision = pe != 0 ? (double)tid * 1.0 / (double)pe : 1.0;
        term = d_recall * (old_precision + precision) / 2;
      }
      double ap = 0.0;
      for (int t = 0; t < ngood; ++t) ap = ap + __shfl_sync(0xffffffffu, term, t);
      if (tid == 0) {
        ap_q[q] = ap;
        first[q] = pe;
        atomicAdd(hist + pe, 1);
      }
    }
    return;
  }
  if (tid == 0) {  // inclusive prefix: position of good s among the non-junk items
    int run = 0;
    for (int e = 0; e < ngood; ++e) { run += sc->gpos[e]; sc->gpos[e] = run; }
  }
  __syncthreads();
  // AP: sum over goods in ascending position of d_recall * (old_precision + precision) / 2 (evaluate.py:94-103)
  const double d_recall = 1.0 / (double)ngood;
  for (int i = tid; i < ngood; i += NT) {
    const int pe = sc->gpos[i];  // rows_good[i]
    const double precision = (double)(i + 1) * 1.0 / (double)(pe + 1);
    const double old_precision = pe != 0 ? (double)i * 1.0 / (double)pe : 1.0;
    sc->terms[i] = d_recall * (old_precision + precision) / 2;
  }
  __syncthreads();
  if (tid == 0) {
    double ap = 0.0;
    for (int t = 0; t < ngood; ++t) ap = ap + sc->terms[t];
    ap_q[q] = ap;
    first[q] = sc->gpos[0];
    atomicAdd(hist + sc->gpos[0], 1);
  }
}

// One query whose G row values live in registers: thread tid holds column ev_col<NT, VEC>(tid, k) in val[k], k < PF
// (G <= NT * PF);
// goodmask = eval_good_mask<NT, PF>() of the same query.
// LAZY values: bit k of donemask says val[k] is the final ranking value; a clear bit says val[k] is still a RAW value
// r with final = finish(r), finish monotone non-decreasing (distances, descending == 0 only).  Raw values are finished
// only where the ranking needs them: the good items, and the columns with r <= raw_bound(worst good value), where
// raw_bound(w) >= every r with finish(r) <= w (a superset test -- the exact comparison follows on the finished value,
// so the result is identical to finishing everything first).  Pass donemask = ~0u when all values are final.
// Same two passes as eval_query_kernel: (1) good items = same id & other camera, rank-sorted; (2) every non-junk column
// ranked before the worst good item bumps the histogram slot of the first good item it precedes.  All NT threads of
// the CTA must call it (block barriers inside).  Writes ap_q[q], first[q], hist[first hit]; sets *status on overflow.
template <int NT, int PF, int VEC, class Finish, class RawBound>
__device__ __forceinline__ void eval_row_in_registers(float (&val)[PF], uint32_t donemask, uint32_t goodmask, Finish finish,
                                                      RawBound raw_bound, int G, int q, int descending,
                                                      const int32_t* __restrict__ ql, const int32_t* __restrict__ qc,
                                                      const int32_t* __restrict__ gl, const int32_t* __restrict__ gc,
                                                      double* __restrict__ ap_q, int32_t* __restrict__ first,
                                                      int32_t* __restrict__ hist, int32_t* __restrict__ status,
                                                      EvalScratch* sc) {
  const int tid = threadIdx.x;
  const int32_t my_l = ql[q], my_c = qc[q];
  if (tid == 0) sc->ngood = 0;
  __syncthreads();
  if (goodmask) {  // a handful of threads per CTA
#pragma unroll
    for (int k = 0; k < PF; ++k)
      if ((goodmask >> k) & 1u) {
        if (!((donemask >> k) & 1u)) { val[k] = finish(val[k]); donemask |= 1u << k; }
        ev_add_good(sc, ev_col<NT, VEC>(tid, k), ev_key(val[k], descending));
      }
  }
  __syncthreads();
  const int ngood = sc->ngood;
  if (ngood == 0) {  // the cmc[0] == -1 sentinel: query skipped (evaluate.py:43-45,80-82)
    if (tid == 0) { ap_q[q] = 0.0; first[q] = -1; }
    return;
  }
  if (ngood > EV_GOOD_CAP) {  // poison the result so that the caller fails loudly
    if (tid == 0) { atomicExch(status, 1); ap_q[q] = __longlong_as_double(0x7ff8000000000000LL); first[q] = -1; }
    return;
  }
  for (int e = tid; e < ngood; e += NT) {  // rank-sort the goods into ranking order
    const uint32_t k = sc->gkey[e];
    const int c = sc->gcol[e];
    int r = 0;
    for (int f = 0; f < ngood; ++f) r += ev_before(sc->gkey[f], sc->gcol[f], k, c, descending);
    sc->skey[r] = k;
    sc->scol[r] = c;
  }
  for (int e = tid; e <= ngood; e += NT) sc->gpos[e] = 0;
  __syncthreads();
  const uint32_t worst_k = sc->skey[ngood - 1];
  const uint32_t worst_o = descending ? ~worst_k : worst_k;
  const float worst_v = __uint_as_float((worst_o & 0x80000000u) ? (worst_o & 0x7fffffffu) : ~worst_o);
  const float raw_t = raw_bound(worst_v);
  const int kfull = (G / (NT * VEC)) * VEC;  // leading registers that hold a valid column in every thread
#pragma unroll
  for (int k = 0; k < PF; ++k) {
    const int j = ev_col<NT, VEC>(tid, k);
    if (k >= kfull && j >= G) continue;
    float v = val[k];
    if (!((donemask >> k) & 1u)) {
      if (!(v <= raw_t)) continue;  // the common case: a raw value that cannot be ranked before the worst good item
      v = finish(v);
    }
    if (descending ? !(v >= worst_v) : !(v <= worst_v)) continue;  // cheap superset test
    ev_count_before(sc, ngood, ev_key(v, descending), j, descending, my_l, my_c, gl, gc);
  }
  __syncthreads();
  eval_ap_tail<NT>(sc, ngood, q, ap_q, first, hist);
}

// CMC[r] = float(#queries with first hit <= r) / Q ; mAP = (sum_q ap_q in query order) / Q (evalmap.cu, single CTA).
int eval_finalize_launch(const int32_t* hist, const double* ap_q, int Q, int G, float* out_cmc, double* out_ap,
                         cudaStream_t st);

}  // namespace reidb200
