// (4) CMC / mAP reduction (reid/evaluate.py:33-105: evaluate_all, evaluate, compute_mAP).
//
// No per-query sort: for the i-th best "good" gallery item g of a query,
//     rows_good[i] = #{non-junk j ranked before g}
// so each query needs only |good| counting passes over its score row.  AP is accumulated in the
// reference's order (ascending rows_good, fp64) and the per-query APs are summed in query order, so
// the mAP is bit-identical to the reference whenever the rankings are.
#include "evalmap.cuh"

namespace reidb200 {

constexpr int EV_THREADS = 256;
constexpr int EV_UNROLL = 16;  // row values in flight per thread in the streaming pass (4 CTAs x 16 KB per SM)

// One CTA per query, one streaming pass over the score row.
//   pass 1  collect the good items (same id, different camera: evaluate.py:66-70), rank-sort them;
//   pass 2  every non-junk item j (junk = label -1 or same id & same camera, evaluate.py:71-72) that is ranked
//           before the WORST good item finds, by binary search in the sorted goods, the first good item s it
//           precedes and bumps hist[s]; position(good s) = sum_{b <= s} hist[b].
// Tie order of the reference: ascending argsort then [::-1] for scores (higher index first), plain ascending
// (lower index first) for distances.
__global__ void __launch_bounds__(EV_THREADS, 4)
eval_query_kernel(const float* __restrict__ mat, int G, int64_t ldm, int descending, const int32_t* __restrict__ ql,
                  const int32_t* __restrict__ qc, const int32_t* __restrict__ gl, const int32_t* __restrict__ gc,
                  double* __restrict__ ap_q, int32_t* __restrict__ first, int32_t* __restrict__ hist,
                  int32_t* __restrict__ status) {
  __shared__ int32_t gcol[EV_GOOD_CAP], scol[EV_GOOD_CAP];
  __shared__ uint32_t gkey[EV_GOOD_CAP], skey[EV_GOOD_CAP];
  __shared__ int32_t gpos[EV_GOOD_CAP + 1];
  __shared__ int s_ngood;
  __shared__ double s_terms[EV_GOOD_CAP];

  const int q = blockIdx.x;
  const int tid = threadIdx.x;
  const float* row = mat + (size_t)q * ldm;
  const int32_t my_l = ql[q], my_c = qc[q];
  if (tid == 0) s_ngood = 0;
  __syncthreads();

  // ---- pass 1: good items (labels only; the gallery label vectors are L2 resident across queries)
  auto good_candidate = [&](int j) {  // label matches: good unless same camera
    if (__ldg(gc + j) != my_c) {
      const int pos = atomicAdd(&s_ngood, 1);
      if (pos < EV_GOOD_CAP) { gcol[pos] = j; gkey[pos] = ev_key(__ldg(row + j), descending); }
    }
  };
  const int G4 = (reinterpret_cast<uintptr_t>(gl) & 15) == 0 ? G >> 2 : 0;  // 128-bit label loads when aligned
  for (int j0 = 0; j0 < G4; j0 += 4 * EV_THREADS) {
    int4 l[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int j4 = j0 + u * EV_THREADS + tid;
      l[u] = j4 < G4 ? __ldg(reinterpret_cast<const int4*>(gl) + j4) : make_int4(-2, -2, -2, -2);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int j = 4 * (j0 + u * EV_THREADS + tid);
      if (l[u].x == my_l) good_candidate(j);
      if (l[u].y == my_l) good_candidate(j + 1);
      if (l[u].z == my_l) good_candidate(j + 2);
      if (l[u].w == my_l) good_candidate(j + 3);
    }
  }
  for (int j = 4 * G4 + tid; j < G; j += EV_THREADS)
    if (__ldg(gl + j) == my_l) good_candidate(j);
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
  // rank-sort the goods into ranking order; gpos doubles as the histogram
  for (int e = tid; e < ngood; e += EV_THREADS) {
    const uint32_t k = gkey[e];
    const int c = gcol[e];
    int r = 0;
    for (int f = 0; f < ngood; ++f) r += ev_before(gkey[f], gcol[f], k, c, descending);
    skey[r] = k;
    scol[r] = c;
  }
  for (int e = tid; e <= ngood; e += EV_THREADS) gpos[e] = 0;
  __syncthreads();
  const uint32_t worst_k = skey[ngood - 1];
  const int worst_c = scol[ngood - 1];
  // float value of the worst good item: a plain float comparison against it is a cheap superset test of "ranked
  // before the worst good" (the key order refines the float order: -0 < +0, index tie-breaks)
  const uint32_t worst_o = descending ? ~worst_k : worst_k;
  const float worst_v = __uint_as_float((worst_o & 0x80000000u) ? (worst_o & 0x7fffffffu) : ~worst_o);

  // ---- pass 2: stream the row once
  for (int j0 = 0; j0 < G; j0 += EV_UNROLL * EV_THREADS) {
    float v[EV_UNROLL];
#pragma unroll
    for (int u = 0; u < EV_UNROLL; ++u) {
      const int j = j0 + u * EV_THREADS + tid;
      v[u] = j < G ? __ldcs(row + j) : 0.f;
    }
#pragma unroll
    for (int u = 0; u < EV_UNROLL; ++u) {
      const int j = j0 + u * EV_THREADS + tid;
      if (j >= G) continue;
      if (descending ? !(v[u] >= worst_v) : !(v[u] <= worst_v)) continue;  // the common case
      const uint32_t kj = ev_key(v[u], descending);
      if (!ev_before(kj, j, worst_k, worst_c, descending)) continue;
      const int32_t l = __ldg(gl + j);
      if (l == -1 || (l == my_l && __ldg(gc + j) == my_c)) continue;  // junk
      int lo = 0, hi = ngood - 1;  // smallest s with j before good s (exists: j is before the worst good)
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (ev_before(kj, j, skey[mid], scol[mid], descending)) hi = mid; else lo = mid + 1;
      }
      atomicAdd(&gpos[lo], 1);
    }
  }
  __syncthreads();
  if (tid == 0) {  // inclusive prefix: position of good s among the non-junk items
    int run = 0;
    for (int e = 0; e < ngood; ++e) { run += gpos[e]; gpos[e] = run; }
  }
  __syncthreads();

  // AP: sum over goods in ascending position of d_recall * (old_precision + precision) / 2 (evaluate.py:94-103)
  const double d_recall = 1.0 / (double)ngood;
  for (int i = tid; i < ngood; i += EV_THREADS) {
    const int pe = gpos[i];  // rows_good[i]
    const double precision = (double)(i + 1) * 1.0 / (double)(pe + 1);
    const double old_precision = pe != 0 ? (double)i * 1.0 / (double)pe : 1.0;
    s_terms[i] = d_recall * (old_precision + precision) / 2;
  }
  __syncthreads();
  if (tid == 0) {
    double ap = 0.0;
    for (int t = 0; t < ngood; ++t) ap = ap + s_terms[t];
    ap_q[q] = ap;
    first[q] = gpos[0];
    atomicAdd(hist + gpos[0], 1);
  }
}

// CMC[r] = float(#queries with first hit <= r) / Q ; mAP = (sum_q ap_q in query order) / Q
// Single CTA; dynamic smem: G + 1 ints (cta_exclusive_scan_smem).
__global__ void __launch_bounds__(1024)
eval_finalize_kernel(const int32_t* __restrict__ hist, const double* __restrict__ ap_q, int Q, int G,
                     float* __restrict__ out_cmc, double* __restrict__ out_ap) {
  extern __shared__ int sbuf[];
  __shared__ int wsum[32];
  __shared__ double part[1024];
  const int tid = threadIdx.x;
  // mAP partial sums: the reference adds the per-query APs in query order (fp64); here 1024 contiguous chunks are
  // summed in order and the chunk sums are added in order (same value up to ~1e-16 relative).
  const int perq = (Q + 1023) / 1024;
  double acc = 0.0;
  for (int q = tid * perq; q < min(Q, (tid + 1) * perq); ++q) acc += ap_q[q];
  part[tid] = acc;
  const int total = cta_exclusive_scan_smem(hist, G, sbuf, wsum);
  if (tid == 0) sbuf[G] = total;
  __syncthreads();
  for (int r = tid; r < G; r += 1024) out_cmc[r] = __fdiv_rn((float)sbuf[r + 1], (float)Q);  // inclusive prefix
  if (tid < 32) {  // 32 chunk sums per lane, in order, then the 32 lane sums in order
    double a2 = 0.0;
    for (int t = 0; t < 32; ++t) a2 += part[tid * 32 + t];
    __syncwarp();
    part[tid] = a2;
    __syncwarp();
    if (tid == 0) {
      double ap = 0.0;
      for (int t = 0; t < 32; ++t) ap += part[t];
      *out_ap = ap / (double)Q;
    }
  }
}

int eval_finalize_launch(const int32_t* hist, const double* ap_q, int Q, int G, float* out_cmc, double* out_ap,
                         cudaStream_t st) {
  static DeviceOnce attr;
  if (attr.pending()) {
    RB_CUDA(cudaFuncSetAttribute(eval_finalize_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  }
  RB_REQUIRE((size_t)(G + 1) * sizeof(int) <= 200 * 1024, "eval_cmc_map: gallery larger than this build's limit (G=%d <= 51199)", G);
  { ProfScope _ps("eval_finalize_kernel", st); eval_finalize_kernel<<<1, 1024, (size_t)(G + 1) * sizeof(int), st>>>(hist, ap_q, Q, G, out_cmc, out_ap); }
  RB_LAUNCH_CHECK("eval_finalize_kernel");
  return REIDB200_OK;
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
  { ProfScope _ps("eval_query_kernel", st); eval_query_kernel<<<Q, EV_THREADS, 0, st>>>(mat, G, ldm, descending, ql, qc, gl, gc, apq, first, hist, status); }
  RB_LAUNCH_CHECK("eval_query_kernel");
  return eval_finalize_launch(hist, apq, Q, G, out_cmc, out_ap, st);
}

// ------------------------------------------------------------------------------------ allocation-free variant
// No counterpart in the reference: added for callers that cannot allocate
// (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  Same kernels as reidb200_eval_cmc_map;
// the scratch (first-hit histogram, per-query AP / first hit when the caller does not want them, status words) lives in
// ONE file-scope device array that every GPU loading the library holds until exit.  All calls share it and it keeps the
// last call's contents: the histogram and status words are cleared on the stream at the start of every call, the
// per-query arrays are fully overwritten; two calls that overlap on the GPU (different streams) race.
constexpr int EVAL_STATIC_MAX_Q = REIDB200_EVAL_STATIC_MAX_Q;
constexpr int EVAL_STATIC_MAX_G = REIDB200_EVAL_STATIC_MAX_G;
constexpr size_t EVAL_STATIC_O_HIST = 0;                                                              // i32[MAX_G]
constexpr size_t EVAL_STATIC_O_APQ = EVAL_STATIC_O_HIST + (size_t)EVAL_STATIC_MAX_G * sizeof(int32_t);  // f64[MAX_Q]
constexpr size_t EVAL_STATIC_O_FIRST = EVAL_STATIC_O_APQ + (size_t)EVAL_STATIC_MAX_Q * sizeof(double);  // i32[MAX_Q]
constexpr size_t EVAL_STATIC_O_STATUS = EVAL_STATIC_O_FIRST + (size_t)EVAL_STATIC_MAX_Q * sizeof(int32_t);  // i32[64]
constexpr size_t EVAL_STATIC_BYTES = EVAL_STATIC_O_STATUS + 64 * sizeof(int32_t);                     // 229 632 B
static_assert(EVAL_STATIC_O_APQ % 256 == 0 && EVAL_STATIC_O_FIRST % 256 == 0 && EVAL_STATIC_O_STATUS % 256 == 0, "scratch pieces stay 256-byte aligned");

static __device__ __align__(256) unsigned char g_eval_static_scratch[EVAL_STATIC_BYTES];

// The array's name is only a stand-in on the host: its device address is looked up once per device, with that device
// current, and that pointer is what the kernels get.
static int eval_static_scratch(unsigned char** out) {
  static unsigned char* addr[64] = {};
  int dev = 0;
  RB_CUDA(cudaGetDevice(&dev));
  RB_REQUIRE(dev >= 0 && dev < 64, "eval_cmc_map_static: device ordinal %d out of range", dev);
  if (!addr[dev]) {
    void* p = nullptr;
    RB_CUDA(cudaGetSymbolAddress(&p, g_eval_static_scratch));
    RB_REQUIRE(p != nullptr, "eval_cmc_map_static: no device address for the static scratch");
    addr[dev] = static_cast<unsigned char*>(p);
  }
  *out = addr[dev];
  return REIDB200_OK;
}

extern "C" int reidb200_eval_cmc_map_static(const float* mat, int Q, int G, int64_t ldm, int descending, const int32_t* ql,
                                            const int32_t* qc, const int32_t* gl, const int32_t* gc, float* out_cmc,
                                            double* out_ap, double* out_ap_q, int32_t* out_first, void* stream) {
  RB_REQUIRE(Q > 0 && G > 0, "eval_cmc_map_static: empty query or gallery set (Q=%d, G=%d)", Q, G);
  RB_REQUIRE(Q <= EVAL_STATIC_MAX_Q && G <= EVAL_STATIC_MAX_G,
             "eval_cmc_map_static: Q=%d / G=%d exceed the static scratch (Q <= %d, G <= %d): use reidb200_eval_cmc_map with a workspace",
             Q, G, EVAL_STATIC_MAX_Q, EVAL_STATIC_MAX_G);
  RB_REQUIRE(mat && ql && qc && gl && gc && out_cmc && out_ap && ldm >= G, "eval_cmc_map_static: bad argument");
  cudaStream_t st = (cudaStream_t)stream;
  unsigned char* base = nullptr;
  if (int rc = eval_static_scratch(&base)) return rc;
  int32_t* hist = reinterpret_cast<int32_t*>(base + EVAL_STATIC_O_HIST);
  double* apq = out_ap_q ? out_ap_q : reinterpret_cast<double*>(base + EVAL_STATIC_O_APQ);
  int32_t* first = out_first ? out_first : reinterpret_cast<int32_t*>(base + EVAL_STATIC_O_FIRST);
  int32_t* status = reinterpret_cast<int32_t*>(base + EVAL_STATIC_O_STATUS);
  RB_CUDA(cudaMemsetAsync(hist, 0, (size_t)G * sizeof(int32_t), st));       // (the previous call's counts are still there)
  RB_CUDA(cudaMemsetAsync(status, 0, 64 * sizeof(int32_t), st));
  { ProfScope _ps("eval_query_kernel", st); eval_query_kernel<<<Q, EV_THREADS, 0, st>>>(mat, G, ldm, descending, ql, qc, gl, gc, apq, first, hist, status); }
  RB_LAUNCH_CHECK("eval_query_kernel");
  return eval_finalize_launch(hist, apq, Q, G, out_cmc, out_ap, st);
}
