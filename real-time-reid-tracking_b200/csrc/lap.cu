// N1: the linear assignment of min_cost_matching on the device-resident cost matrix
//   cost[cost > max_distance] = max_distance + 1e-5 ; linear_sum_assignment ; pairs above max_distance dropped
//   boxmot/trackers/strongsort/sort/linear_assignment.py:58-77 (scipy.optimize.linear_sum_assignment, _lsap.cpp)
// After the clamp every gated entry costs the same G > any admissible entry, so the optimum splits over the connected
// components of the bipartite graph of ADMISSIBLE (cost <= max_distance) entries: an assignment pays
// sum_S c + (k - |S|) * G for its admissible pairs S, i.e. it minimises sum_S (c - G) over matchings S, component by
// component, and whatever scipy pairs through gated entries is dropped again by the reference's post-filter.  A
// tracker frame has many small components (a track sees one to three detections inside its gate), so
//   lap_components_kernel  one CTA: active rows / columns, union-find over the admissible entries in shared memory,
//                          rows and columns sorted by component
//   lap_solve_kernel       one WARP per component: the shortest-augmenting-path algorithm scipy uses (Crouse 2016), in
//                          float64 with the same reduced costs and dual updates, on the clamped sub-matrix
// Matches are identical to scipy's whenever the optimum is unique (generic float64 costs); between assignments of exactly
// equal total cost either may be returned.
#include "common.cuh"

namespace reidb200 {

constexpr int LAP_MAX_SIDE = 2048;

__device__ __forceinline__ int sm_find(volatile int* parent, int x) {
  int p = parent[x];
  while (p != x) { x = p; p = parent[x]; }
  return x;
}

__device__ __forceinline__ void sm_unite(int* parent, int a, int b) {
  while (true) {
    a = sm_find(parent, a);
    b = sm_find(parent, b);
    if (a == b) return;
    if (a < b) { const int t = a; a = b; b = t; }
    if (atomicCAS(parent + a, a, b) == a) return;
  }
}

// info[0] = number of components, info[1] = active rows, info[2] = active columns
__global__ void __launch_bounds__(1024)
lap_components_kernel(const double* __restrict__ cost, size_t ld, const int32_t* __restrict__ row_cand, int nrc, int M,
                      double max_distance, const int32_t* __restrict__ match_col_of_row, const int32_t* __restrict__ det_matched,
                      int32_t* __restrict__ rows_sorted, int32_t* __restrict__ cols_sorted, int32_t* __restrict__ comps,
                      int32_t* __restrict__ info, int debug_stop) {
  extern __shared__ int sm[];
  __shared__ int wsum[32];
  __shared__ int s_ncomp;
  int* ar = sm;                    // active rows (original indices), ascending position in row_cand
  int* ac = ar + nrc;              // active columns
  int* parent = ac + M;            // nodes: active row a -> a, active column b -> nr + b
  int* root = parent + nrc + M;
  int* buf = root + nrc + M;       // max(nrc, M) ints
  const int tid = threadIdx.x;
  for (int k = tid; k < nrc; k += 1024) root[k] = match_col_of_row[row_cand[k]] < 0 ? 1 : 0;
  __syncthreads();
  const int nr = cta_exclusive_scan_smem(root, nrc, buf, wsum);
  for (int k = tid; k < nrc; k += 1024)
    if (root[k]) ar[buf[k]] = row_cand[k];
  __syncthreads();
  for (int j = tid; j < M; j += 1024) root[j] = det_matched[j] == 0 ? 1 : 0;
  __syncthreads();
  const int nc = cta_exclusive_scan_smem(root, M, buf, wsum);
  for (int j = tid; j < M; j += 1024)
    if (root[j]) ac[buf[j]] = j;
  if (tid == 0) s_ncomp = 0;
  __syncthreads();
  const int n = nr + nc;
  for (int k = tid; k < n; k += 1024) parent[k] = k;
  __syncthreads();
  if (debug_stop == 1) { if (tid == 0) info[0] = 0; return; }  // (REIDB200_LAP_DEBUG: timing only)
  if (nr > 0 && nc > 0) {
    // a warp takes two rows at a time, its lanes stride the columns: sixteen loads in flight per lane, no index division
    const int wp = tid >> 5, lane = tid & 31;
    for (int a0 = wp; a0 < nr; a0 += 64) {
      const int a1 = a0 + 32;
      const double* r0 = cost + (size_t)ar[a0] * ld;
      const double* r1 = a1 < nr ? cost + (size_t)ar[a1] * ld : r0;
      for (int b0 = 0; b0 < nc; b0 += 256) {
        double c0[8], c1[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const int b = b0 + k * 32 + lane;
          const int col = b < nc ? ac[b] : 0;
          c0[k] = b < nc ? r0[col] : INFINITY;
          c1[k] = (b < nc && a1 < nr) ? r1[col] : INFINITY;
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const int b = b0 + k * 32 + lane;
          if (c0[k] <= max_distance) sm_unite(parent, a0, nr + b);
          if (c1[k] <= max_distance) sm_unite(parent, a1, nr + b);
        }
      }
    }
  }
  __syncthreads();
  if (debug_stop == 2) { if (tid == 0) info[0] = 0; return; }
  for (int k = tid; k < n; k += 1024) root[k] = sm_find(parent, k);
  __syncthreads();
  // rows and columns in (component root, index) order through one unique key per node; a component with a row is led
  // by its lowest row (rows carry the lower node numbers)
  int* key = parent;  // (the union-find forest is no longer needed)
  for (int k = tid; k < n; k += 1024) key[k] = root[k] * 4096 + (k < nr ? k : k - nr);
  __syncthreads();
  for (int k = tid; k < n; k += 1024) {
    const int mine = key[k];
    int rank = 0;
    if (k < nr) {
      for (int o = 0; o < nr; ++o) rank += key[o] < mine ? 1 : 0;
      rows_sorted[rank] = ar[k];
      if (root[k] == k) {  // leader: sizes of its component and the columns sorted before it
        int cnt = 0, before = 0, ncols = 0;
        for (int o = 0; o < nr; ++o) cnt += root[o] == k ? 1 : 0;
        for (int o = nr; o < n; ++o) {
          before += root[o] < k ? 1 : 0;
          ncols += root[o] == k ? 1 : 0;
        }
        if (ncols > 0) {
          const int c = atomicAdd(&s_ncomp, 1);
          comps[4 * c] = rank; comps[4 * c + 1] = cnt; comps[4 * c + 2] = before; comps[4 * c + 3] = ncols;
        }
      }
    } else {
      for (int o = nr; o < n; ++o) rank += key[o] < mine ? 1 : 0;
      cols_sorted[rank] = ac[k - nr];
    }
  }
  __syncthreads();
  if (tid == 0) { info[0] = s_ncomp; info[1] = nr; info[2] = nc; }
}

struct LapWork {
  double *u, *v, *sp;                   // duals of the A / B side, shortest path costs (each T + M)
  int32_t *path, *a4b, *b4a, *SR, *SC;  // each T + M
};

// Warp arg-min over (value, free column preferred, lower column index): the order scipy's scan resolves equal costs in.
__device__ __forceinline__ void lap_argmin(double& best, int& best_free, int& best_b) {
#pragma unroll
  for (int o = 16; o; o >>= 1) {
    const double ob = __shfl_xor_sync(0xffffffffu, best, o);
    const int obb = __shfl_xor_sync(0xffffffffu, best_b, o), of = __shfl_xor_sync(0xffffffffu, best_free, o);
    if (ob < best || (ob == best && (of > best_free || (of == best_free && obb < best_b)))) { best = ob; best_b = obb; best_free = of; }
  }
}

constexpr int LAP_SOLVE_WARPS = 4;

// A component of at most 32 x 32: the clamped sub-matrix sits in shared memory, lane b owns column b (dual v, shortest
// path cost, path, owner) and lane a owns row a (dual u, assigned column) in registers; the same arithmetic in the same
// order as the general path below.  cs = this warp's 32 x 33 doubles.
__device__ __forceinline__ void lap_solve_small(double* cs, int na, int nb, int lane, double max_distance, int& b4a_out,
                                                bool& admissible_out) {
  double u_a = 0.0, v_b = 0.0;
  int b4a_a = -1, a4b_b = -1;
  for (int cur = 0; cur < na; ++cur) {
    double sp_b = INFINITY;
    int path_b = -1;
    bool SC_b = false, SR_a = false;
    int sink = -1, i = cur;
    double min_val = 0.0;
    while (sink < 0) {
      if (lane == i) SR_a = true;
      const double ui = __shfl_sync(0xffffffffu, u_a, i);
      double best = INFINITY;
      int best_b = INT_MAX, best_free = 0;
      if (lane < nb && !SC_b) {
        const double r = min_val + cs[i * 33 + lane] - ui - v_b;
        if (r < sp_b) { sp_b = r; path_b = i; }
        best = sp_b; best_b = lane; best_free = a4b_b < 0 ? 1 : 0;
      }
      lap_argmin(best, best_free, best_b);
      if (best_b == INT_MAX) {  // cannot happen with the clamped (finite) costs; never spin
        b4a_out = -1;
        admissible_out = false;
        return;
      }
      min_val = best;
      if (lane == best_b) SC_b = true;
      const int owner = __shfl_sync(0xffffffffu, a4b_b, best_b);
      if (owner < 0) sink = best_b; else i = owner;
    }
    const double sp_of_mine = __shfl_sync(0xffffffffu, sp_b, b4a_a < 0 ? 0 : b4a_a);
    if (lane == cur) u_a = u_a + min_val;
    else if (SR_a) u_a = u_a + (min_val - sp_of_mine);
    if (SC_b) v_b = v_b - (min_val - sp_b);
    int j = sink;
    for (int hop = 0; hop <= na; ++hop) {  // augment along the path back to cur
      const int a = __shfl_sync(0xffffffffu, path_b, j);
      if (lane == j) a4b_b = a;
      const int t = __shfl_sync(0xffffffffu, b4a_a, a);
      if (lane == a) b4a_a = j;
      j = t;
      if (a == cur) break;
    }
  }
  b4a_out = b4a_a;
  admissible_out = lane < na && cs[lane * 33 + (b4a_a < 0 ? 0 : b4a_a)] <= max_distance;
}

__global__ void __launch_bounds__(32 * LAP_SOLVE_WARPS)
lap_solve_kernel(const double* __restrict__ cost, size_t ld, double max_distance, const int32_t* __restrict__ rows_sorted,
                 const int32_t* __restrict__ cols_sorted, const int32_t* __restrict__ comps, const int32_t* __restrict__ info,
                 LapWork w, int32_t* __restrict__ match_col_of_row, int32_t* __restrict__ det_matched, int32_t* __restrict__ status) {
  __shared__ double s_cost[LAP_SOLVE_WARPS][32 * 33];
  const int comp = blockIdx.x * LAP_SOLVE_WARPS + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (comp >= info[0]) return;
  const int row_off = comps[4 * comp], nrows = comps[4 * comp + 1], col_off = comps[4 * comp + 2], ncols = comps[4 * comp + 3];
  const int32_t* R = rows_sorted + row_off;
  const int32_t* C = cols_sorted + col_off;
  if (nrows == 1 && ncols == 1) {  // the usual case: one track, one detection inside its gate
    if (lane == 0) { match_col_of_row[R[0]] = C[0]; det_matched[C[0]] = 1; }
    return;
  }
  const bool tr = nrows > ncols;  // scipy transposes so that the augmented side is the smaller one
  const int na = tr ? ncols : nrows, nb = tr ? nrows : ncols;
  const int32_t* A = tr ? C : R;
  const int32_t* B = tr ? R : C;
  const double gated = max_distance + 1e-5;
  auto cst = [&](int a, int b) {
    const double x = tr ? cost[(size_t)B[b] * ld + A[a]] : cost[(size_t)A[a] * ld + B[b]];
    return x <= max_distance ? x : gated;
  };
  if (nb <= 32) {  // the usual size: registers + shared memory
    double* cs = s_cost[threadIdx.x >> 5];
    for (int a = 0; a < na; ++a)
      if (lane < nb) cs[a * 33 + lane] = cst(a, lane);
    __syncwarp();
    int b4a_a;
    bool ok;
    lap_solve_small(cs, na, nb, lane, max_distance, b4a_a, ok);
    if (ok) {
      const int row = tr ? B[b4a_a] : A[lane], col = tr ? A[lane] : B[b4a_a];
      match_col_of_row[row] = col;
      det_matched[col] = 1;
    }
    return;
  }
  const int base = row_off + col_off;
  double* u = w.u + base;
  double* v = w.v + base;
  double* sp = w.sp + base;
  int32_t *path = w.path + base, *a4b = w.a4b + base, *b4a = w.b4a + base, *SR = w.SR + base, *SC = w.SC + base;
  for (int a = lane; a < na; a += 32) { u[a] = 0.0; b4a[a] = -1; }
  for (int b = lane; b < nb; b += 32) { v[b] = 0.0; a4b[b] = -1; }
  __syncwarp();
  for (int cur = 0; cur < na; ++cur) {
    for (int b = lane; b < nb; b += 32) { sp[b] = INFINITY; SC[b] = 0; }
    for (int a = lane; a < na; a += 32) SR[a] = 0;
    __syncwarp();
    int sink = -1, i = cur, guard = 0;
    double min_val = 0.0;
    while (sink < 0) {
      if (lane == 0) SR[i] = 1;
      const double ui = u[i];
      double best = INFINITY;
      int best_b = INT_MAX, best_free = 0;
      for (int b = lane; b < nb; b += 32) {
        if (SC[b]) continue;
        const double r = min_val + cst(i, b) - ui - v[b];
        double s = sp[b];
        if (r < s) { s = r; sp[b] = r; path[b] = i; }
        const int fr = a4b[b] < 0 ? 1 : 0;  // scipy prefers an unassigned column among equal costs
        if (s < best || (s == best && (fr > best_free || (fr == best_free && b < best_b)))) { best = s; best_b = b; best_free = fr; }
      }
      lap_argmin(best, best_free, best_b);
      if (best_b == INT_MAX || ++guard > nb) {  // cannot happen with finite costs; never spin
        if (lane == 0) atomicExch(status, 1);
        return;
      }
      min_val = best;
      if (lane == 0) SC[best_b] = 1;
      __syncwarp();
      const int owner = a4b[best_b];
      if (owner < 0) sink = best_b; else i = owner;
    }
    if (lane == 0) u[cur] = u[cur] + min_val;
    for (int a = lane; a < na; a += 32)
      if (SR[a] && a != cur) u[a] = u[a] + (min_val - sp[b4a[a]]);
    for (int b = lane; b < nb; b += 32)
      if (SC[b]) v[b] = v[b] - (min_val - sp[b]);
    __syncwarp();
    if (lane == 0) {  // augment along the path back to cur
      int j = sink;
      for (int hop = 0; hop <= na; ++hop) {  // (a path visits every A node at most once)
        const int a = path[j];
        a4b[j] = a;
        const int t = b4a[a];
        b4a[a] = j;
        j = t;
        if (a == cur) break;
      }
    }
    __syncwarp();
  }
  for (int a = lane; a < na; a += 32) {
    const int b = b4a[a];
    const int row = tr ? B[b] : A[a], col = tr ? A[a] : B[b];
    if (cost[(size_t)row * ld + col] <= max_distance) {  // the reference's post-filter (linear_assignment.py:72-76)
      match_col_of_row[row] = col;
      det_matched[col] = 1;
    }
  }
}

}  // namespace reidb200

using namespace reidb200;

static size_t lap_align(size_t x) { return (x + 255) & ~(size_t)255; }

extern "C" size_t reidb200_lap_workspace_bytes(int T, int M) {
  const size_t n = (size_t)(T > 0 ? T : 0) + (size_t)(M > 0 ? M : 0);
  return lap_align(n * 4) * 2 + lap_align(n * 16 + 64) + lap_align(n * 8) * 3 + lap_align(n * 4) * 5 + 256;
}

extern "C" int reidb200_lap_gated(const double* cost, int T, int M, size_t ld, const int32_t* row_cand, int nrc, double max_distance,
                                  int32_t* match_col_of_row, int32_t* det_matched, int32_t* status, void* ws, size_t ws_bytes,
                                  void* stream) {
  if (T == 0 || M == 0 || nrc == 0) return REIDB200_OK;
  RB_REQUIRE(cost && row_cand && match_col_of_row && det_matched && status && T > 0 && M > 0 && nrc > 0 && ld >= (size_t)M,
             "lap_gated: bad argument");
  RB_REQUIRE(nrc <= LAP_MAX_SIDE && M <= LAP_MAX_SIDE && nrc <= T, "lap_gated: at most %d rows / columns per solve", LAP_MAX_SIDE);
  RB_REQUIRE(ws && ws_bytes >= reidb200_lap_workspace_bytes(T, M), "lap_gated: workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  const size_t n = (size_t)T + M;
  uint8_t* p = (uint8_t*)ws;
  auto take = [&](size_t bytes) { uint8_t* q = p; p += lap_align(bytes); return q; };
  int32_t* rows_sorted = (int32_t*)take(n * 4);
  int32_t* cols_sorted = (int32_t*)take(n * 4);
  int32_t* comps = (int32_t*)take(n * 16 + 64);
  int32_t* info = comps + 4 * n;
  LapWork w;
  w.u = (double*)take(n * 8); w.v = (double*)take(n * 8); w.sp = (double*)take(n * 8);
  w.path = (int32_t*)take(n * 4); w.a4b = (int32_t*)take(n * 4); w.b4a = (int32_t*)take(n * 4);
  w.SR = (int32_t*)take(n * 4); w.SC = (int32_t*)take(n * 4);
  const size_t smem = ((size_t)3 * (nrc + M) + (size_t)(nrc > M ? nrc : M)) * sizeof(int);
  static DeviceOnce attr;
  if (attr.pending()) {
    RB_CUDA(cudaFuncSetAttribute(lap_components_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)(((size_t)3 * 2 * LAP_MAX_SIDE + LAP_MAX_SIDE) * sizeof(int))));
  }
  static const int debug_stop = getenv("REIDB200_LAP_DEBUG") ? atoi(getenv("REIDB200_LAP_DEBUG")) : 0;  // timing experiments only
  {
    ProfScope _ps("lap_components_kernel", st);
    lap_components_kernel<<<1, 1024, smem, st>>>(cost, ld, row_cand, nrc, M, max_distance, match_col_of_row, det_matched, rows_sorted,
                                                 cols_sorted, comps, info, debug_stop);
  }
  RB_LAUNCH_CHECK("lap_components_kernel");
  const int max_comps = nrc < M ? nrc : M;
  {
    ProfScope _ps("lap_solve_kernel", st);
    lap_solve_kernel<<<(max_comps + LAP_SOLVE_WARPS - 1) / LAP_SOLVE_WARPS, 32 * LAP_SOLVE_WARPS, 0, st>>>(cost, ld, max_distance, rows_sorted, cols_sorted, comps, info, w,
                                                          match_col_of_row, det_matched, status);
  }
  RB_LAUNCH_CHECK("lap_solve_kernel");
  return REIDB200_OK;
}
