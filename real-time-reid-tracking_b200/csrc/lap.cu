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

// ------------------------------------------------------------------------------------ exact mode
// The SAME assignment scipy returns, ties included -- and with it the ORDER of the reference's unmatched lists (the pairs
// scipy forms through gated entries, all of equal cost, decide where a detection lands in `unmatched_detections`, hence
// which id a new track gets: linear_assignment.py:58-77, tracker.py:175-233).  One warp runs scipy's _lsap.cpp on the whole
// clamped matrix of the candidate rows x candidate columns, statement by statement:
//   * the smaller side is augmented (scipy transposes when there are more rows than columns);
//   * reduced costs as ((min_val + c) - u[i]) - v[j] in float64;
//   * the scan runs over `remaining` (columns not yet in the tree: filled in DESCENDING column order at the start of every
//     augmentation, a visited column replaced by the last one) and resolves equal shortest-path costs the way the scalar
//     loop does: the LAST unassigned column of the scan if there is one, else the FIRST column of the scan.
// The sequential part never waits for global memory: the whole CTA first gathers the ADMISSIBLE entries (cost <=
// max_distance; everything else costs the same constant) into a CSR by augmented-side row in shared memory; a step of the
// search drops the current row's few entries into a dense row of constants, scans it, and takes them out again.  (When the
// admissible entries do not fit, the scan reads the matrix directly.)
// Checked against scipy on tie-heavy matrices: oracle/reid_oracle.py::lsap_scipy_order (CPU), tests/test_gpu_lap.py.
constexpr int LX_THREADS = 256;

struct LxSmem {  // sizes of the dynamic shared-memory tables, computed on the host
  int side, cap;
};

__global__ void __launch_bounds__(LX_THREADS)
lap_exact_kernel(const double* __restrict__ cost, size_t ld, const int32_t* __restrict__ row_cand, int nrc,
                 const int32_t* __restrict__ col_cand, const int32_t* __restrict__ col_count, int max_cols, double max_distance,
                 int32_t* __restrict__ match_col_of_row, int32_t* __restrict__ det_matched, int32_t* __restrict__ um_dets,
                 int32_t* __restrict__ um_count, int32_t* __restrict__ um_rows, int32_t* __restrict__ um_rows_count,
                 int32_t* __restrict__ status, LxSmem lay) {
  extern __shared__ __align__(16) unsigned char lx_smem[];
  __shared__ int s_nr, s_nc, s_total;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int ncc = min(col_count ? col_count[0] : max_cols, max_cols);
  const int side = lay.side, cap = lay.cap;
  double* u = reinterpret_cast<double*>(lx_smem);
  double* v = u + side;
  double* sp = v + side;
  double* rowval = sp + side;
  double* csr_val = rowval + side;
  double* trip_val = csr_val + cap;
  int32_t* path = reinterpret_cast<int32_t*>(trip_val + cap);
  int32_t* a4b = path + side;
  int32_t* b4a = a4b + side;
  int32_t* remaining = b4a + side;
  int32_t* R = remaining + side;
  int32_t* C = R + side;
  int32_t* off = C + side;          // [side + 1]
  int32_t* fillc = off + side + 1;  // [side]
  int32_t* stamp = fillc + side;    // [side] id of the step that wrote rowval[b] last
  int32_t* csr_col = stamp + side;  // [cap]
  int32_t* trip_a = csr_col + cap;  // [cap]
  int32_t* trip_b = trip_a + cap;   // [cap]
  int32_t* vis_col = trip_b + cap;   // [side] column chosen by step s of the current augmentation
  int32_t* vis_pos = vis_col + side; // [side] its position in `remaining`

  // ---- candidate rows / columns still unmatched, in the order given (warp 0)
  if (warp == 0) {
    int nr = 0, nc = 0;
    for (int t0 = 0; t0 < nrc; t0 += 32) {
      const int t = t0 + lane;
      const int r = t < nrc ? row_cand[t] : -1;
      const bool keep = r >= 0 && match_col_of_row[r] < 0;
      const unsigned m = __ballot_sync(0xffffffffu, keep);
      if (keep) R[nr + __popc(m & ((1u << lane) - 1))] = r;
      nr += __popc(m);
    }
    for (int t0 = 0; t0 < ncc; t0 += 32) {
      const int t = t0 + lane;
      const int c = t < ncc ? col_cand[t] : -1;
      const bool keep = c >= 0 && det_matched[c] == 0;
      const unsigned m = __ballot_sync(0xffffffffu, keep);
      if (keep) C[nc + __popc(m & ((1u << lane) - 1))] = c;
      nc += __popc(m);
    }
    if (lane == 0) { s_nr = nr; s_nc = nc; s_total = 0; }
  }
  __syncthreads();
  const int nr = s_nr, nc = s_nc;
  if (nr == 0 || nc == 0) {  // nothing to match: every candidate column stays unmatched, in its order
    for (int t = tid; t < nc; t += LX_THREADS) um_dets[t] = C[t];
    if (tid == 0) um_count[0] = nc;
    if (um_rows) {  // (the reference returns track_indices as it got them)
      for (int t = tid; t < nr; t += LX_THREADS) um_rows[t] = R[t];
      if (tid == 0) um_rows_count[0] = nr;
    }
    return;
  }
  const bool tr = nc < nr;
  const int na = tr ? nc : nr, nb = tr ? nr : nc;
  const int32_t* A = tr ? C : R;
  const int32_t* B = tr ? R : C;
  const double gated = max_distance + 1e-5;
  // ---- admissible entries -> CSR by augmented-side row (whole CTA)
  for (int a = tid; a <= na; a += LX_THREADS) off[a] = 0;
  for (int a = tid; a < na; a += LX_THREADS) fillc[a] = 0;
  for (int b = tid; b < nb; b += LX_THREADS) rowval[b] = gated;
  __syncthreads();
  if (cap > 0) {
    for (int p = warp; p < nr; p += LX_THREADS / 32) {
      const double* crow = cost + (size_t)R[p] * ld;
#pragma unroll 4
      for (int q = lane; q < nc; q += 32) {
        const double x = crow[C[q]];
        if (x <= max_distance) {
          const int e = atomicAdd(&s_total, 1);
          const int a = tr ? q : p, b = tr ? p : q;
          if (e < cap) { trip_val[e] = x; trip_a[e] = a; trip_b[e] = b; }
          atomicAdd(&off[a + 1], 1);
        }
      }
    }
  }
  __syncthreads();
  const int total = s_total;
  const bool use_csr = cap > 0 && total <= cap;
  if (use_csr) {
    if (warp == 0) {  // inclusive scan of the counts in off[1 .. na] -> row extents
      const int per = (na + 31) >> 5;
      const int b0 = 1 + lane * per, e0 = min(na + 1, b0 + per);
      int sum = 0;
      for (int t = b0; t < e0; ++t) sum += off[t];
      int incl = sum;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int x = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += x;
      }
      int run = incl - sum;
      for (int t = b0; t < e0; ++t) { run += off[t]; off[t] = run; }
    }
    __syncthreads();
    for (int e = tid; e < total; e += LX_THREADS) {
      const int a = trip_a[e];
      const int pos = off[a] + atomicAdd(&fillc[a], 1);
      csr_val[pos] = trip_val[e];
      csr_col[pos] = trip_b[e];
    }
  }
  __syncthreads();

  // ---- the search: scipy's loop, every scan spread over the CTA (a lone warp runs each dependent shared-memory access at
  // full latency: 6700 cycles per augmentation measured, 3700 in this form; two barriers per step)
  auto cst = [&](int a, int b) {  // direct read (the admissible entries did not fit)
    const double x = tr ? cost[(size_t)B[b] * ld + A[a]] : cost[(size_t)A[a] * ld + B[b]];
    return x <= max_distance ? x : gated;
  };
  __shared__ unsigned w_hi[LX_THREADS / 32], w_lo[LX_THREADS / 32];
  __shared__ int w_free[LX_THREADS / 32], w_first[LX_THREADS / 32], w_jfree[LX_THREADS / 32], w_jfirst[LX_THREADS / 32];
  // Per augmentation only what the search TOUCHED is reset or updated: the first step writes every shortest-path cost
  // (no reset of sp[]), the visited columns are kept as a list (vis_col / vis_pos, in visit order) for the dual
  // updates, and `remaining` is put back to its canonical descending order at the visited positions only.
  for (int a = tid; a < na; a += LX_THREADS) { u[a] = 0.0; b4a[a] = -1; }
  for (int b = tid; b < nb; b += LX_THREADS) { v[b] = 0.0; a4b[b] = -1; stamp[b] = -1; remaining[b] = nb - b - 1; }
  int step_id = 0;
  __syncthreads();
  for (int cur = 0; cur < na; ++cur) {
    int num_remaining = nb, sink = -1, i = cur, nsteps = 0;
    double min_val = 0.0;
    while (sink < 0) {
      // row i of the clamped matrix = one constant + its few admissible entries, stamped with this step's id
      if (use_csr)
        for (int e = off[i] + tid; e < off[i + 1]; e += LX_THREADS) { rowval[csr_col[e]] = csr_val[e]; stamp[csr_col[e]] = step_id; }
      __syncthreads();  // (also: the previous step's bookkeeping, the previous augmentation's updates)
      const double ui = u[i];
      const bool first_step = nsteps == 0;
      double best = INFINITY;
      int first_it = INT_MAX, lastfree_it = -1;
      for (int it = tid; it < num_remaining; it += LX_THREADS) {
        const int j = remaining[it];
        const double c = use_csr ? (stamp[j] == step_id ? rowval[j] : gated) : cst(i, j);
        const double r = min_val + c - ui - v[j];
        double sv = first_step ? INFINITY : sp[j];
        if (r < sv) { sv = r; sp[j] = r; path[j] = i; }
        else if (first_step) sp[j] = sv;  // (r is not below infinity: an infinite reduced cost stays recorded as such)
        const bool fr = a4b[j] < 0;
        if (sv < best) { best = sv; first_it = it; lastfree_it = fr ? it : -1; }
        else if (sv == best && fr) lastfree_it = it;
      }
      // the minimum (32-bit warp reductions over the ordered bit pattern of the double; + 0.0 folds -0.0 into +0.0 so that
      // bit equality is numeric equality); among the threads that hold it, the last unassigned column of the scan, else
      // the first column -- first inside each warp, then over the warps' results (each warp publishes the columns too, so
      // that nobody reads `remaining` while thread 0 updates it)
      best = best + 0.0;
      unsigned long long key = (unsigned long long)__double_as_longlong(best);
      key = (key >> 63) ? ~key : (key | 0x8000000000000000ull);
      const unsigned khi = (unsigned)(key >> 32), klo = (unsigned)key;
      unsigned mhi = __reduce_min_sync(0xffffffffu, khi);
      unsigned mlo = __reduce_min_sync(0xffffffffu, khi == mhi ? klo : 0xffffffffu);
      const bool mine = khi == mhi && klo == mlo;
      int gfree = __reduce_max_sync(0xffffffffu, mine ? lastfree_it : -1);
      int gfirst = __reduce_min_sync(0xffffffffu, mine ? first_it : INT_MAX);
      if (lane == 0) {
        w_hi[warp] = mhi; w_lo[warp] = mlo; w_free[warp] = gfree; w_first[warp] = gfirst;
        w_jfree[warp] = gfree >= 0 ? remaining[gfree] : -1;
        w_jfirst[warp] = gfirst != INT_MAX ? remaining[gfirst] : -1;
      }
      __syncthreads();
      mhi = 0xffffffffu; mlo = 0xffffffffu;
#pragma unroll
      for (int x = 0; x < LX_THREADS / 32; ++x)
        if (w_hi[x] < mhi || (w_hi[x] == mhi && w_lo[x] < mlo)) { mhi = w_hi[x]; mlo = w_lo[x]; }
      gfree = -1; gfirst = INT_MAX;
      int jfree = -1, jfirst = -1;
#pragma unroll
      for (int x = 0; x < LX_THREADS / 32; ++x)
        if (w_hi[x] == mhi && w_lo[x] == mlo) {
          if (w_free[x] > gfree) { gfree = w_free[x]; jfree = w_jfree[x]; }
          if (w_first[x] < gfirst) { gfirst = w_first[x]; jfirst = w_jfirst[x]; }
        }
      unsigned long long gk = ((unsigned long long)mhi << 32) | mlo;
      gk = (gk >> 63) ? (gk & 0x7fffffffffffffffull) : ~gk;
      const double gmin = __longlong_as_double((long long)gk);
      if (!(gmin < INFINITY)) {  // NaN / inf costs: scipy raises; never spin
        if (tid == 0) { atomicExch(status, 1); um_count[0] = 0; if (um_rows) um_rows_count[0] = 0; }
        return;
      }
      const int index = gfree >= 0 ? gfree : gfirst;
      const int j = gfree >= 0 ? jfree : jfirst;
      min_val = gmin;
      const int owner = a4b[j];
      if (tid == 0) {  // (the next scan starts behind the next barrier)
        vis_col[nsteps] = j;
        vis_pos[nsteps] = index;
        remaining[index] = remaining[num_remaining - 1];
      }
      --num_remaining;
      ++step_id;
      ++nsteps;
      if (owner < 0) sink = j; else i = owner;
    }
    __syncthreads();  // the visit lists are complete
    // dual updates (scipy: u[cur] += minVal; u[i] += minVal - sp[col4row[i]] for the other visited rows; v[j] -= minVal -
    // sp[j] for the visited columns): step s >= 1 scanned the row that owns the column of step s - 1
    for (int sidx = tid; sidx < nsteps; sidx += LX_THREADS) {
      const int jc = vis_col[sidx];
      const double d = min_val - sp[jc];
      if (sidx + 1 < nsteps) { const int a = a4b[jc]; u[a] = u[a] + d; }  // the row scanned by step sidx + 1
      v[jc] = v[jc] - d;
      remaining[vis_pos[sidx]] = nb - vis_pos[sidx] - 1;  // back to the canonical order
    }
    if (tid == 0) u[cur] = u[cur] + min_val;
    __syncthreads();  // a4b is read above and rewritten below
    if (tid == 0) {  // augment along the path back to cur
      int j = sink;
      for (int hop = 0; hop <= na; ++hop) {
        const int a = path[j];
        a4b[j] = a;
        const int t = b4a[a];
        b4a[a] = j;
        j = t;
        if (a == cur) break;
      }
    }
  }
  __syncthreads();
  if (warp != 0) return;
  // ---- the reference's lists (linear_assignment.py:61-76): columns scipy left unassigned, in column order; then, pair by
  // pair in ROW order, matches (cost <= max_distance) and the gated pairs, whose detection joins the unmatched list
  int cnt = 0, rcnt = 0;
  if (!tr) {
    for (int q0 = 0; q0 < nc; q0 += 32) {
      const int q = q0 + lane;
      const bool keep = q < nc && a4b[q] < 0;
      const unsigned m = __ballot_sync(0xffffffffu, keep);
      if (keep) um_dets[cnt + __popc(m & ((1u << lane) - 1))] = C[q];
      cnt += __popc(m);
    }
  } else if (um_rows) {  // rows scipy left unassigned (more rows than columns), in row order
    for (int p0 = 0; p0 < nr; p0 += 32) {
      const int p = p0 + lane;
      const bool keep = p < nr && a4b[p] < 0;
      const unsigned m = __ballot_sync(0xffffffffu, keep);
      if (keep) um_rows[rcnt + __popc(m & ((1u << lane) - 1))] = R[p];
      rcnt += __popc(m);
    }
  }
  for (int p0 = 0; p0 < nr; p0 += 32) {
    const int p = p0 + lane;
    const int q = p < nr ? (tr ? a4b[p] : b4a[p]) : -1;
    bool gated_pair = false;
    if (q >= 0) {
      const int row = R[p], col = C[q];
      if (cost[(size_t)row * ld + col] <= max_distance) { match_col_of_row[row] = col; det_matched[col] = 1; }
      else gated_pair = true;
    }
    const unsigned m = __ballot_sync(0xffffffffu, gated_pair);
    const int at = __popc(m & ((1u << lane) - 1));
    if (gated_pair) {
      um_dets[cnt + at] = C[q];
      if (um_rows) um_rows[rcnt + at] = R[p];
    }
    cnt += __popc(m);
    rcnt += __popc(m);
  }
  if (lane == 0) {
    um_count[0] = cnt;
    if (um_rows) um_rows_count[0] = rcnt;
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

// Exact mode (see lap_exact_kernel): rows = row_cand (in that order, those still unmatched), columns = col_cand[0 ..
// *col_count) (device list + device count, in that order, those still unmatched; col_count == NULL: all max_cols entries).
// Writes the matches into match_col_of_row / det_matched like reidb200_lap_gated and the reference's
// `unmatched_detections` list, in the reference's order, into um_dets[0 .. *um_count), and (optional) its `unmatched_tracks`
// list into um_rows[0 .. *um_rows_count).
extern "C" int reidb200_lap_exact(const double* cost, int T, int M, size_t ld, const int32_t* row_cand, int nrc,
                                  const int32_t* col_cand, const int32_t* col_count, int max_cols, double max_distance,
                                  int32_t* match_col_of_row, int32_t* det_matched, int32_t* um_dets, int32_t* um_count,
                                  int32_t* um_rows, int32_t* um_rows_count, int32_t* status, void* stream) {
  RB_REQUIRE(cost && col_cand && match_col_of_row && det_matched && um_dets && um_count && status && T > 0 && M > 0 && ld >= (size_t)M,
             "lap_exact: bad argument");
  RB_REQUIRE(nrc >= 0 && nrc <= T && max_cols >= 0 && max_cols <= M && (nrc == 0 || row_cand), "lap_exact: bad candidate lists");
  RB_REQUIRE((um_rows == nullptr) == (um_rows_count == nullptr), "lap_exact: um_rows and um_rows_count go together");
  RB_REQUIRE(nrc <= LAP_MAX_SIDE && max_cols <= LAP_MAX_SIDE, "lap_exact: at most %d rows / columns per solve", LAP_MAX_SIDE);
  cudaStream_t st = (cudaStream_t)stream;
  // shared memory: per-side arrays (4 f64 + 11 i32 per element) + the admissible-entry tables (2 f64 + 3 i32 per
  // entry); the tables get what is left of ~200 KB, at most 8192 entries, none below 256 (direct reads then)
  const size_t side = (size_t)(nrc > max_cols ? nrc : max_cols) + 1;
  const size_t per_side = 4 * sizeof(double) + 11 * sizeof(int32_t), per_entry = 2 * sizeof(double) + 3 * sizeof(int32_t);
  const size_t budget = 200 * 1024;
  size_t cap = budget > side * per_side + 64 ? (budget - side * per_side - 64) / per_entry : 0;
  if (cap > 8192) cap = 8192;
  if (cap < 256 || getenv("REIDB200_LAP_NO_CSR")) cap = 0;
  cap &= ~(size_t)1;  // keeps the i32 tables behind the f64 tables 8-byte aligned
  const size_t smem = side * per_side + cap * per_entry + 64;
  static DeviceOnce attr;
  if (attr.pending()) {
    RB_CUDA(cudaFuncSetAttribute(lap_exact_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(budget + 4096)));
  }
  LxSmem lay;
  lay.side = (int)side; lay.cap = (int)cap;
  {
    ProfScope _ps("lap_exact_kernel", st);
    lap_exact_kernel<<<1, LX_THREADS, smem, st>>>(cost, ld, row_cand, nrc, col_cand, col_count, max_cols, max_distance,
                                                  match_col_of_row, det_matched, um_dets, um_count, um_rows, um_rows_count, status, lay);
  }
  RB_LAUNCH_CHECK("lap_exact_kernel");
  return REIDB200_OK;
}
