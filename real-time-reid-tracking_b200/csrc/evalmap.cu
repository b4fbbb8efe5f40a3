// (4) CMC / mAP reduction (reid/evaluate.py:33-105: evaluate_all, evaluate, compute_mAP).
//
// No per-query sort: for the i-th best "good" gallery item g of a query,
//     rows_good[i] = #{non-junk j ranked before g}
// so each query needs only |good| counting passes over its score row.  AP is accumulated in the
// reference's order (ascending rows_good, fp64) and the per-query APs are summed in query order, so
// the mAP is bit-identical to the reference whenever the rankings are.
#include "common.cuh"

namespace reidb200 {

constexpr int EV_THREADS = 256;
constexpr int EV_GOOD_CAP = 1024;

__device__ __forceinline__ uint32_t ev_key(float v, int descending) {
  const uint32_t u = float_to_ordered(v);
  return descending ? ~u : u;  // smaller key = ranked earlier
}

// Key of gallery item j as ranked for this query: junk items (label -1, or same id & same camera,
// evaluate.py:71-72) get the largest key so that they are never counted.
__device__ __forceinline__ uint32_t ev_masked_key(const float* row, const int32_t* gl, const int32_t* gc, int j,
                                                  int32_t my_l, int32_t my_c, int descending) {
  const int32_t l = __ldg(gl + j);
  const bool junk = (l == -1) || (l == my_l && __ldg(gc + j) == my_c);
  return junk ? 0xffffffffu : ev_key(__ldg(row + j), descending);
}

// One CTA per query.  Pass 1 stages the junk-masked keys of the whole score row in shared memory (when it
// fits: G * 4 bytes of dynamic smem) and collects the good items; pass 2 counts, for four good items per warp at
// a time, the non-junk items ranked before them.
template <bool SMEM_ROW>
__global__ void __launch_bounds__(EV_THREADS)
eval_query_kernel(const float* __restrict__ mat, int G, int64_t ldm, int descending, const int32_t* __restrict__ ql,
                  const int32_t* __restrict__ qc, const int32_t* __restrict__ gl, const int32_t* __restrict__ gc,
                  double* __restrict__ ap_q, int32_t* __restrict__ first, int32_t* __restrict__ hist,
                  int32_t* __restrict__ status) {
  extern __shared__ uint32_t skey[];
  __shared__ int32_t gcol[EV_GOOD_CAP];
  __shared__ uint32_t gkey[EV_GOOD_CAP];
  __shared__ int32_t gpos[EV_GOOD_CAP];
  __shared__ int s_ngood;
  __shared__ double s_terms[EV_GOOD_CAP];
  __shared__ int s_best;

  const int q = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const float* row = mat + (size_t)q * ldm;
  const int32_t my_l = ql[q], my_c = qc[q];
  if (tid == 0) s_ngood = 0;
  __syncthreads();

  // good = same id & different camera (evaluate.py:66-70).  Four independent elements per thread per trip so
  // that several loads are in flight; good items are rare, so a plain shared-memory counter collects them.
  for (int j0 = 0; j0 < G; j0 += 4 * EV_THREADS) {
    int32_t l[4];
    float v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int j = j0 + u * EV_THREADS + tid;
      l[u] = j < G ? __ldg(gl + j) : -2;
      v[u] = j < G ? __ldg(row + j) : 0.f;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int j = j0 + u * EV_THREADS + tid;
      if (j >= G) continue;
      const uint32_t key = ev_key(v[u], descending);
      bool junk = l[u] == -1;
      if (l[u] == my_l) {
        if (__ldg(gc + j) == my_c) junk = true;
        else {
          const int pos = atomicAdd(&s_ngood, 1);
          if (pos < EV_GOOD_CAP) { gcol[pos] = j; gkey[pos] = key; }
        }
      }
      if (SMEM_ROW) skey[j] = junk ? 0xffffffffu : key;
    }
  }
  __syncthreads();
  const int ngood = s_ngood;
  if (ngood == 0) {  // the cmc[0] == -1 sentinel: query skipped (evaluate.py:43-45,80-82)
    if (tid == 0) { ap_q[q] = 0.0; first[q] = -1; }
    return;
  }
  if (ngood > EV_GOOD_CAP) {
    // more good matches than this build holds: poison the result so that the caller fails loudly
    if (tid == 0) { atomicExch(status, 1); ap_q[q] = __longlong_as_double(0x7ff8000000000000LL); first[q] = -1; }
    return;
  }

  // position of every good item among the non-junk items: four good items per warp per sweep of the row.
  // Tie order of the reference: ascending argsort then [::-1] for scores (higher index first), plain ascending
  // (lower index first) for distances.
  for (int e0 = warp * 4; e0 < ngood; e0 += (EV_THREADS / 32) * 4) {
    uint32_t k[4];
    int col[4], cnt[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const bool valid = e0 + u < ngood;
      k[u] = valid ? gkey[e0 + u] : 0u;
      col[u] = valid ? gcol[e0 + u] : -1;
      cnt[u] = 0;
    }
#pragma unroll 4
    for (int j = lane; j < G; j += 32) {
      const uint32_t kj = SMEM_ROW ? skey[j] : ev_masked_key(row, gl, gc, j, my_l, my_c, descending);
#pragma unroll
      for (int u = 0; u < 4; ++u)
        cnt[u] += (kj < k[u]) || (kj == k[u] && kj != 0xffffffffu && (descending ? j > col[u] : j < col[u]));
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int c = warp_sum_i(cnt[u]);
      if (lane == 0 && e0 + u < ngood) gpos[e0 + u] = c;
    }
  }
  __syncthreads();

  // AP: sum over goods in ascending position of d_recall * (old_precision + precision) / 2 (evaluate.py:94-103)
  const double d_recall = 1.0 / (double)ngood;
  for (int e = tid; e < ngood; e += EV_THREADS) {
    const int pe = gpos[e];
    int i = 0;  // number of goods ranked before this one (positions are distinct)
    for (int f = 0; f < ngood; ++f) i += gpos[f] < pe;
    const double precision = (double)(i + 1) * 1.0 / (double)(pe + 1);
    const double old_precision = pe != 0 ? (double)i * 1.0 / (double)pe : 1.0;
    s_terms[i] = d_recall * (old_precision + precision) / 2;
    if (i == 0) s_best = pe;
  }
  __syncthreads();
  double ap = 0.0;
  int best = 0;
  if (tid == 0) {
    for (int t = 0; t < ngood; ++t) ap = ap + s_terms[t];
    best = s_best;
  }
  if (tid == 0) {
    ap_q[q] = ap;
    first[q] = best;
    atomicAdd(hist + best, 1);
  }
}

// CMC[r] = float(#queries with first hit <= r) / Q ; mAP = (sum_q ap_q in query order) / Q
__global__ void __launch_bounds__(1024)
eval_finalize_kernel(const int32_t* __restrict__ hist, const double* __restrict__ ap_q, int Q, int G,
                     float* __restrict__ out_cmc, double* __restrict__ out_ap) {
  __shared__ int wsum[32];
  __shared__ int carry_s;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) carry_s = 0;
  __syncthreads();
  for (int r0 = 0; r0 < G; r0 += 1024) {
    const int r = r0 + tid;
    int v = r < G ? hist[r] : 0;
    // inclusive scan inside the warp
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int n = __shfl_up_sync(0xffffffffu, v, o);
      if (lane >= o) v += n;
    }
    if (lane == 31) wsum[warp] = v;
    __syncthreads();
    int before = carry_s;
    for (int w = 0; w < warp; ++w) before += wsum[w];
    const int total = v + before;
    if (r < G) out_cmc[r] = __fdiv_rn((float)total, (float)Q);
    __syncthreads();
    if (tid == 1023) carry_s = total;
    __syncthreads();
  }
  // mAP: the reference adds the per-query APs in query order (fp64); here 1024 contiguous chunks are summed in
  // order and the chunk sums are added in order (same value up to ~1e-16 relative).
  __shared__ double part[1024];
  const int per = (Q + 1023) / 1024;
  double acc = 0.0;
  for (int q = tid * per; q < min(Q, (tid + 1) * per); ++q) acc += ap_q[q];
  part[tid] = acc;
  __syncthreads();
  if (tid == 0) {
    double ap = 0.0;
    for (int t = 0; t < 1024; ++t) ap += part[t];
    *out_ap = ap / (double)Q;
  }
}

}  // namespace reidb200

using namespace reidb200;

extern "C" size_t reidb200_eval_workspace_bytes(int Q, int G) {
  if (Q <= 0 || G <= 0) return 0;
  CarveSize c;
  c.take<int32_t>(G);       // first-hit histogram
  c.take<double>(Q);        // per-query AP
  c.take<int32_t>(Q);       // first hit
  c.take<int32_t>(64);      // status
  return align_up(c.off, 256);
}

extern "C" int reidb200_eval_cmc_map(const float* mat, int Q, int G, int64_t ldm, int descending, const int32_t* ql,
                                     const int32_t* qc, const int32_t* gl, const int32_t* gc, float* out_cmc,
                                     double* out_ap, double* out_ap_q, int32_t* out_first, void* workspace,
                                     size_t workspace_bytes, void* stream) {
  RB_REQUIRE(Q > 0 && G > 0, "eval_cmc_map: empty query or gallery set (Q=%d, G=%d)", Q, G);
  RB_REQUIRE(mat && ql && qc && gl && gc && out_cmc && out_ap && ldm >= G, "eval_cmc_map: bad argument");
  RB_REQUIRE(workspace && workspace_bytes >= reidb200_eval_workspace_bytes(Q, G), "eval_cmc_map: workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  Carver c(workspace, workspace_bytes);
  int32_t* hist = c.take<int32_t>(G);
  double* apq = c.take<double>(Q);
  int32_t* first = c.take<int32_t>(Q);
  int32_t* status = c.take<int32_t>(64);
  if (out_ap_q) apq = out_ap_q;
  if (out_first) first = out_first;
  RB_CUDA(cudaMemsetAsync(hist, 0, (size_t)G * sizeof(int32_t), st));
  RB_CUDA(cudaMemsetAsync(status, 0, 64 * sizeof(int32_t), st));
  const size_t row_bytes = (size_t)G * sizeof(uint32_t);
  if (row_bytes <= 160 * 1024) {
    static bool attr = false;
    if (!attr) {
      RB_CUDA(cudaFuncSetAttribute(eval_query_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
      attr = true;
    }
    ProfScope _ps("eval_query_kernel", st);
    eval_query_kernel<true><<<Q, EV_THREADS, row_bytes, st>>>(mat, G, ldm, descending, ql, qc, gl, gc, apq, first, hist, status);
  } else {
    ProfScope _ps("eval_query_kernel", st);
    eval_query_kernel<false><<<Q, EV_THREADS, 0, st>>>(mat, G, ldm, descending, ql, qc, gl, gc, apq, first, hist, status);
  }
  RB_LAUNCH_CHECK("eval_query_kernel");
  { ProfScope _ps("eval_finalize_kernel", st); eval_finalize_kernel<<<1, 1024, 0, st>>>(hist, apq, Q, G, out_cmc, out_ap); }
  RB_LAUNCH_CHECK("eval_finalize_kernel");
  return REIDB200_OK;
}
