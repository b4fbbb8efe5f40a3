// N4: the steps around compute_jaccard_distance in reid/image_reid_inference.py:252-310, device resident so
// that the N x N Jaccard matrix (1.49 GB at Market size) never leaves HBM:
//   flip-TTA merge     (e1 + e2) / 2 -> F.normalize                      image_reid_inference.py:252-253,268-269
//   dists[dists<0]=0 ; dists += attribute_dist                            image_reid_inference.py:287-290
//   DBSCAN(eps, min_samples, metric="precomputed").fit_predict(dists)     image_reid_inference.py:291-301 (sklearn)
//   smooth_tracklets   e*0.1 + mean(group)*0.9 per (sequence, cluster)    reid/inference_utils.py:19-27
// DBSCAN follows sklearn's definition on a dense precomputed matrix: neighbourhood(i) = {j : d[i,j] <= eps} (self
// included, compared in float32), core = |neighbourhood| >= min_samples, clusters = connected components of the core
// points numbered by their lowest core index (sklearn seeds its depth-first search at ascending i), a border point
// takes the lowest-numbered cluster that has a core point in its neighbourhood (the first search that reaches it),
// everything else is noise (-1).  The matrix is read once, row by row (HBM bound): the degree pass keeps each row's
// neighbour list (up to 64 entries; longer rows are scanned again), the union and border passes walk the lists.  d is
// taken as symmetric (the Jaccard matrix is, bit for bit): row i stands for column i.
#include "common.cuh"

namespace reidb200 {

// Calls hit(j) for every j in [0, n) with p[j] <= eps.  A warp walks the row with eight float4 loads in flight per lane;
// rows need not be 16-byte aligned (N is odd at Market size): the first few elements up to the next 16-byte boundary
// and the last few are read as scalars.  Hits are rare, so a lane gathers them in a bit mask and serves them in one place.
template <typename F>
__device__ __forceinline__ void row_hits(const float* __restrict__ p, int n, float eps, int lane, F hit) {
  int head = (int)(((16u - (uint32_t)((uintptr_t)p & 15u)) & 15u) >> 2);
  head = head < n ? head : n;
  if (lane < head && p[lane] <= eps) hit(lane);
  const float* q = p + head;  // 16-byte aligned
  const int n4 = (n - head) & ~3;
  for (int c0 = lane * 4; c0 < n4; c0 += 1024) {
    float4 v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int c = c0 + u * 128;
      v[u] = c < n4 ? __ldcs(reinterpret_cast<const float4*>(q + c)) : make_float4(INFINITY, INFINITY, INFINITY, INFINITY);
    }
    uint32_t m = 0;
#pragma unroll
    for (int u = 0; u < 8; ++u)
      m |= ((v[u].x <= eps ? 1u : 0u) | (v[u].y <= eps ? 2u : 0u) | (v[u].z <= eps ? 4u : 0u) | (v[u].w <= eps ? 8u : 0u)) << (4 * u);
    while (m) {
      const int b = __ffs(m) - 1;
      m &= m - 1;
      hit(head + c0 + (b >> 2) * 128 + (b & 3));
    }
  }
  for (int c = head + n4 + lane; c < n; c += 32)
    if (p[c] <= eps) hit(c);
}

// Hits of a row are kept (up to DB_CAP per row, any order) so that the later passes need not read the matrix again;
// a row with more hits than that is simply scanned again where it is needed.
constexpr int DB_CAP = 64;

// One CTA per row (its 256 threads stream the row: consecutive warps read consecutive 512-byte pieces).
__global__ void __launch_bounds__(256)
dbscan_degree_kernel(const float* __restrict__ d, int N, size_t ld, float eps, int min_samples, int32_t* __restrict__ core,
                     int32_t* __restrict__ parent, int32_t* __restrict__ degree, int32_t* __restrict__ hits) {
  __shared__ int s_cnt;
  const int row = blockIdx.x, tid = threadIdx.x;
  if (tid == 0) s_cnt = 0;
  __syncthreads();
  int32_t* h = hits + (size_t)row * DB_CAP;
  const float* p = d + (size_t)row * ld;
  auto hit = [&](int j) {
    const int pos = atomicAdd(&s_cnt, 1);
    if (pos < DB_CAP) h[pos] = j;
  };
  int head = (int)(((16u - (uint32_t)((uintptr_t)p & 15u)) & 15u) >> 2);
  head = head < N ? head : N;
  if (tid < head && p[tid] <= eps) hit(tid);
  const float* q = p + head;  // 16-byte aligned
  const int n4 = (N - head) & ~3;
  for (int c0 = tid * 4; c0 < n4; c0 += 4096) {
    float4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int c = c0 + u * 1024;
      v[u] = c < n4 ? __ldcs(reinterpret_cast<const float4*>(q + c)) : make_float4(INFINITY, INFINITY, INFINITY, INFINITY);
    }
    uint32_t m = 0;
#pragma unroll
    for (int u = 0; u < 4; ++u)
      m |= ((v[u].x <= eps ? 1u : 0u) | (v[u].y <= eps ? 2u : 0u) | (v[u].z <= eps ? 4u : 0u) | (v[u].w <= eps ? 8u : 0u)) << (4 * u);
    while (m) {
      const int b = __ffs(m) - 1;
      m &= m - 1;
      hit(head + c0 + (b >> 2) * 1024 + (b & 3));
    }
  }
  for (int c = head + n4 + tid; c < N; c += 256)
    if (p[c] <= eps) hit(c);
  __syncthreads();
  if (tid == 0) {
    const int cnt = s_cnt;
    degree[row] = cnt;
    core[row] = cnt >= min_samples ? 1 : 0;
    parent[row] = row;
  }
}

// Calls hit(j) for the row's neighbours j < limit: from the stored list, or from the matrix when the list overflowed.
template <typename F>
__device__ __forceinline__ void row_neighbours(const float* __restrict__ d, size_t ld, int row, int limit, float eps, int lane,
                                               const int32_t* __restrict__ degree, const int32_t* __restrict__ hits, F hit) {
  const int cnt = degree[row];
  if (cnt <= DB_CAP) {
    for (int e = lane; e < cnt; e += 32) {
      const int j = hits[(size_t)row * DB_CAP + e];
      if (j < limit) hit(j);
    }
  } else {
    row_hits(d + (size_t)row * ld, limit, eps, lane, hit);
  }
}

// Union-find over the core points; a parent always has the smaller index, so the root of a component is its lowest
// core index (the point sklearn's search starts the cluster from).
__device__ __forceinline__ int uf_find(int32_t* parent, int x) {
  volatile int32_t* vp = parent;
  int r = x, p;
  while ((p = vp[r]) != r) r = p;
  while (x > r) {  // path compression (indices only fall along a path; a concurrent hook may lead below r: stop there)
    p = vp[x];
    if (p > r) atomicMin(parent + x, r);
    x = p;
  }
  return r;
}

__device__ __noinline__ void uf_unite(int32_t* parent, int a, int b) {
  while (true) {
    a = uf_find(parent, a);
    b = uf_find(parent, b);
    if (a == b) return;
    if (a < b) { const int t = a; a = b; b = t; }
    if (atomicCAS(parent + a, a, b) == a) return;  // else a was hooked elsewhere meanwhile: look again
  }
}

__global__ void __launch_bounds__(256)
dbscan_link_kernel(const float* __restrict__ d, int N, size_t ld, float eps, const int32_t* __restrict__ core, int32_t* parent,
                   const int32_t* __restrict__ degree, const int32_t* __restrict__ hits) {
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= N || !core[row]) return;
  row_neighbours(d, ld, row, row, eps, lane, degree, hits, [&](int j) {  // lower triangle: every core pair once
    if (core[j]) uf_unite(parent, row, j);
  });
}

// One CTA of 1024 threads: cluster number = rank of the component's root among the roots; core points get their
// label, the others -1 for now.  Dynamic shared memory: N ints.
__global__ void __launch_bounds__(1024)
dbscan_number_kernel(int N, const int32_t* __restrict__ core, int32_t* parent, int32_t* __restrict__ flags,
                     int32_t* __restrict__ labels, int32_t* __restrict__ n_clusters) {
  extern __shared__ int sbuf[];
  __shared__ int wsum[32];
  const int tid = threadIdx.x;
  for (int i = tid; i < N; i += 1024) flags[i] = (core[i] && parent[i] == i) ? 1 : 0;
  __syncthreads();
  const int total = cta_exclusive_scan_smem(flags, N, sbuf, wsum);
  for (int i = tid; i < N; i += 1024) labels[i] = core[i] ? sbuf[uf_find(parent, i)] : -1;
  if (tid == 0 && n_clusters) *n_clusters = total;
}

__global__ void __launch_bounds__(256)
dbscan_border_kernel(const float* __restrict__ d, int N, size_t ld, float eps, const int32_t* __restrict__ core,
                     int32_t* labels, const int32_t* __restrict__ degree, const int32_t* __restrict__ hits) {
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= N || core[row]) return;
  int best = INT_MAX;
  // labels of core points are final (written by the previous kernel); only non-core rows are written here
  row_neighbours(d, ld, row, N, eps, lane, degree, hits, [&](int j) {
    if (core[j]) best = min(best, *(volatile int32_t*)(labels + j));
  });
  best = __reduce_min_sync(0xffffffffu, best);
  if (lane == 0) labels[row] = best == INT_MAX ? -1 : best;
}

// dist = max(dist, 0) [+ other], row by row (float4 when the rows allow it)
__global__ void __launch_bounds__(256)
clamp_add_kernel(float* __restrict__ dist, int rows, int cols, size_t ld, const float* __restrict__ other, size_t ldo, int vec) {
  const int row = blockIdx.y;
  float* p = dist + (size_t)row * ld;
  const float* q = other ? other + (size_t)row * ldo : nullptr;
  if (vec) {
    const int c = (blockIdx.x * 256 + threadIdx.x) * 4;
    if (c + 3 < cols) {
      float4 v = *reinterpret_cast<float4*>(p + c);
      v.x = v.x < 0.f ? 0.f : v.x; v.y = v.y < 0.f ? 0.f : v.y; v.z = v.z < 0.f ? 0.f : v.z; v.w = v.w < 0.f ? 0.f : v.w;
      if (q) {
        const float4 o = __ldcs(reinterpret_cast<const float4*>(q + c));
        v.x = v.x + o.x; v.y = v.y + o.y; v.z = v.z + o.z; v.w = v.w + o.w;
      }
      *reinterpret_cast<float4*>(p + c) = v;
      return;
    }
    for (int e = c; e < cols; ++e) {
      float v = p[e];
      v = v < 0.f ? 0.f : v;
      p[e] = q ? v + q[e] : v;
    }
  } else {
    for (int u = 0; u < 4; ++u) {
      const int e = blockIdx.x * 1024 + u * 256 + threadIdx.x;
      if (e < cols) {
        float v = p[e];
        v = v < 0.f ? 0.f : v;
        p[e] = q ? v + q[e] : v;
      }
    }
  }
}

// the same over fully contiguous matrices, as one flat array (rows of odd length are not 16-byte aligned one by one)
__global__ void __launch_bounds__(256)
clamp_add_flat_kernel(float* __restrict__ dist, const float* __restrict__ other, size_t n4, size_t n) {
  const size_t stride = (size_t)gridDim.x * 256;
  for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n4; i += stride) {
    float4 v = reinterpret_cast<float4*>(dist)[i];
    v.x = v.x < 0.f ? 0.f : v.x; v.y = v.y < 0.f ? 0.f : v.y; v.z = v.z < 0.f ? 0.f : v.z; v.w = v.w < 0.f ? 0.f : v.w;
    if (other) {
      const float4 o = __ldcs(reinterpret_cast<const float4*>(other) + i);
      v.x = v.x + o.x; v.y = v.y + o.y; v.z = v.z + o.z; v.w = v.w + o.w;
    }
    reinterpret_cast<float4*>(dist)[i] = v;
  }
  if (blockIdx.x == 0)
    for (size_t e = n4 * 4 + threadIdx.x; e < n; e += 256) {
      float v = dist[e];
      v = v < 0.f ? 0.f : v;
      dist[e] = other ? v + other[e] : v;
    }
}

// out[r] = normalize((a[r] + b[r]) / 2): one warp per row, x / max(||x||_2, 1e-12) as F.normalize
__global__ void __launch_bounds__(256)
tta_merge_normalize_kernel(const float* __restrict__ a, size_t lda, const float* __restrict__ b, size_t ldb, int rows, int D,
                           float* __restrict__ out, size_t ldo) {
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= rows) return;
  const float* pa = a + (size_t)row * lda;
  const float* pb = b ? b + (size_t)row * ldb : nullptr;
  float ss = 0.f;
  for (int c = lane; c < D; c += 32) {
    const float v = pb ? (pa[c] + pb[c]) / 2.0f : pa[c];
    ss = ss + v * v;
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) ss = ss + __shfl_xor_sync(0xffffffffu, ss, o);
  const float nrm = fmaxf(sqrtf(ss), 1e-12f);
  float* po = out + (size_t)row * ldo;
  for (int c = lane; c < D; c += 32) {
    const float v = pb ? (pa[c] + pb[c]) / 2.0f : pa[c];
    po[c] = v / nrm;
  }
}

// One CTA per group g: rows order[seg[g] .. seg[g+1]) of emb become emb * keep + mean(rows) * avg_w, column by column
// (thread = column: coalesced across the row, the rows of a group summed in list order).
__global__ void __launch_bounds__(256)
smooth_groups_kernel(float* __restrict__ emb, int D, size_t lde, const int32_t* __restrict__ order,
                     const int32_t* __restrict__ seg, float keep, float avg_w) {
  const int g = blockIdx.x;
  const int b = seg[g], e = seg[g + 1];
  if (e <= b) return;
  const float n_rows = (float)(e - b);
  for (int c = threadIdx.x; c < D; c += 256) {
    float s = 0.f;
    for (int r = b; r < e; ++r) s = s + emb[(size_t)order[r] * lde + c];
    const float avg = s / n_rows;
    for (int r = b; r < e; ++r) {
      float* p = emb + (size_t)order[r] * lde + c;
      *p = *p * keep + avg * avg_w;
    }
  }
}

// out = x - column mean(x) (diminish_camera_bias, reid/inference_utils.py:9-10): a CTA owns 32 columns; 8 row lanes per
// column sum every 8th row, a shared-memory tree joins them (fixed order), then the CTA subtracts.
__global__ void __launch_bounds__(256)
center_columns_kernel(const float* __restrict__ x, int rows, int D, size_t ldx, float* __restrict__ out, size_t ldo,
                      float* __restrict__ mean_out) {
  __shared__ float part[8][33];
  const int cl = threadIdx.x & 31, rl = threadIdx.x >> 5, c = blockIdx.x * 32 + cl;
  float s = 0.f;
  if (c < D)
    for (int r = rl; r < rows; r += 8) s = s + x[(size_t)r * ldx + c];
  part[rl][cl] = s;
  __syncthreads();
  if (rl == 0) {
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) t = t + part[k][cl];
    part[0][cl] = t / (float)rows;
    if (mean_out && c < D) mean_out[c] = part[0][cl];
  }
  __syncthreads();
  const float m = part[0][cl];
  if (c < D)
    for (int r = rl; r < rows; r += 8) out[(size_t)r * ldo + c] = x[(size_t)r * ldx + c] - m;
}

}  // namespace reidb200

using namespace reidb200;

extern "C" size_t reidb200_dbscan_workspace_bytes(int N) { return (size_t)(N > 0 ? N : 0) * (4 + DB_CAP) * sizeof(int32_t) + 256; }

extern "C" int reidb200_dbscan_precomputed(const float* dist, int N, size_t ld, float eps, int min_samples, int32_t* labels,
                                           int32_t* n_clusters, int32_t* core_out, void* ws, size_t ws_bytes, void* stream) {
  if (N == 0) return REIDB200_OK;
  RB_REQUIRE(dist && labels && N > 0 && ld >= (size_t)N && min_samples >= 1, "dbscan: bad argument");
  RB_REQUIRE((size_t)N * 4 + 256 <= 232448 - 256, "dbscan: N = %d exceeds the %d points the cluster-numbering pass holds", N,
             (232448 - 512) / 4);
  RB_REQUIRE(ws && ws_bytes >= reidb200_dbscan_workspace_bytes(N), "dbscan: workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  int32_t* core = core_out ? core_out : (int32_t*)ws;
  int32_t* parent = (int32_t*)ws + N;
  int32_t* flags = (int32_t*)ws + 2 * (size_t)N;
  int32_t* degree = (int32_t*)ws + 3 * (size_t)N;
  int32_t* hits = (int32_t*)ws + 4 * (size_t)N;
  const int grid = (N + 7) / 8;
  {
    ProfScope _ps("dbscan_degree_kernel", st);
    dbscan_degree_kernel<<<N, 256, 0, st>>>(dist, N, ld, eps, min_samples, core, parent, degree, hits);
  }
  RB_LAUNCH_CHECK("dbscan_degree_kernel");
  {
    ProfScope _ps("dbscan_link_kernel", st);
    dbscan_link_kernel<<<grid, 256, 0, st>>>(dist, N, ld, eps, core, parent, degree, hits);
  }
  RB_LAUNCH_CHECK("dbscan_link_kernel");
  {
    static DeviceOnce attr;
    if (attr.pending()) {
      RB_CUDA(cudaFuncSetAttribute(dbscan_number_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448 - 256));
    }
    ProfScope _ps("dbscan_number_kernel", st);
    dbscan_number_kernel<<<1, 1024, (size_t)N * 4, st>>>(N, core, parent, flags, labels, n_clusters);
  }
  RB_LAUNCH_CHECK("dbscan_number_kernel");
  {
    ProfScope _ps("dbscan_border_kernel", st);
    dbscan_border_kernel<<<grid, 256, 0, st>>>(dist, N, ld, eps, core, labels, degree, hits);
  }
  RB_LAUNCH_CHECK("dbscan_border_kernel");
  return REIDB200_OK;
}

extern "C" int reidb200_clamp_add(float* dist, int rows, int cols, size_t ld, const float* other, size_t ldo, void* stream) {
  if (rows == 0 || cols == 0) return REIDB200_OK;
  RB_REQUIRE(dist && rows > 0 && cols > 0 && ld >= (size_t)cols && (!other || ldo >= (size_t)cols), "clamp_add: bad argument");
  RB_REQUIRE(rows <= 65535 * 1024, "clamp_add: too many rows");
  cudaStream_t st = (cudaStream_t)stream;
  if (ld == (size_t)cols && (!other || ldo == (size_t)cols) && (uintptr_t)dist % 16 == 0 && (!other || (uintptr_t)other % 16 == 0)) {
    const size_t n = (size_t)rows * cols, n4 = n / 4;
    const size_t want = (n4 + 255) / 256;
    const int grid = (int)(want < (size_t)num_sms() * 16 ? (want ? want : 1) : (size_t)num_sms() * 16);
    { ProfScope _ps("clamp_add_kernel", st); clamp_add_flat_kernel<<<grid, 256, 0, st>>>(dist, other, n4, n); }
    RB_LAUNCH_CHECK("clamp_add_kernel");
    return REIDB200_OK;
  }
  const int vec = (ld % 4 == 0) && ((uintptr_t)dist % 16 == 0) && (!other || ((ldo % 4 == 0) && ((uintptr_t)other % 16 == 0)));
  // grid.y is limited to 65535 rows per launch
  for (int r0 = 0; r0 < rows; r0 += 65535) {
    const int nr = rows - r0 < 65535 ? rows - r0 : 65535;
    dim3 grid((cols + 1023) / 1024, nr);
    ProfScope _ps("clamp_add_kernel", st);
    clamp_add_kernel<<<grid, 256, 0, st>>>(dist + (size_t)r0 * ld, nr, cols, ld, other ? other + (size_t)r0 * ldo : nullptr, ldo, vec);
  }
  RB_LAUNCH_CHECK("clamp_add_kernel");
  return REIDB200_OK;
}

extern "C" int reidb200_tta_merge_normalize(const float* a, size_t lda, const float* b, size_t ldb, int rows, int D, float* out,
                                            size_t ldo, void* stream) {
  if (rows == 0) return REIDB200_OK;
  RB_REQUIRE(a && out && rows > 0 && D > 0 && lda >= (size_t)D && ldo >= (size_t)D && (!b || ldb >= (size_t)D),
             "tta_merge_normalize: bad argument");
  cudaStream_t st = (cudaStream_t)stream;
  { ProfScope _ps("tta_merge_normalize_kernel", st); tta_merge_normalize_kernel<<<(rows + 7) / 8, 256, 0, st>>>(a, lda, b, ldb, rows, D, out, ldo); }
  RB_LAUNCH_CHECK("tta_merge_normalize_kernel");
  return REIDB200_OK;
}

extern "C" int reidb200_smooth_groups(float* emb, int D, size_t lde, const int32_t* order, const int32_t* seg_start, int n_groups,
                                      float keep, float avg_w, void* stream) {
  if (n_groups == 0) return REIDB200_OK;
  RB_REQUIRE(emb && order && seg_start && n_groups > 0 && D > 0 && lde >= (size_t)D, "smooth_groups: bad argument");
  cudaStream_t st = (cudaStream_t)stream;
  { ProfScope _ps("smooth_groups_kernel", st); smooth_groups_kernel<<<n_groups, 256, 0, st>>>(emb, D, lde, order, seg_start, keep, avg_w); }
  RB_LAUNCH_CHECK("smooth_groups_kernel");
  return REIDB200_OK;
}

extern "C" int reidb200_center_columns(const float* x, int rows, int D, size_t ldx, float* out, size_t ldo, float* mean_out,
                                       void* stream) {
  if (rows == 0 || D == 0) return REIDB200_OK;
  RB_REQUIRE(x && out && rows > 0 && D > 0 && ldx >= (size_t)D && ldo >= (size_t)D, "center_columns: bad argument");
  cudaStream_t st = (cudaStream_t)stream;
  { ProfScope _ps("center_columns_kernel", st); center_columns_kernel<<<(D + 31) / 32, 256, 0, st>>>(x, rows, D, ldx, out, ldo, mean_out); }
  RB_LAUNCH_CHECK("center_columns_kernel");
  return REIDB200_OK;
}
