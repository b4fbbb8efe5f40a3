// (5) DeepSort / StrongSort association cost.
//   gallery   boxmot/utils/matching.py:363-409  NearestNeighborDistanceMetric (device-resident ring buffer)
//   cosine    boxmot/utils/matching.py:278-339  _cosine_distance / _nn_cosine_distance
//   gate      boxmot/motion/adapters/strongsort_kf_adapter.py:137-169 -> kalman_filter.py:577-586 and
//             boxmot/trackers/strongsort/sort/linear_assignment.py:147-198 (gate_cost_matrix), :58-59 (clamp)
//   iou       boxmot/trackers/strongsort/sort/iou_matching.py:9-86, modification_deepsort/iou_matching.py:5-47
// One kernel per frame: CTA t owns track t: per-track gallery x detections cosine minimum (fp32),
// Mahalanobis gate + blend (fp64), IoU / DIoU row (fp64).
#include <stdlib.h>
#include <string.h>

#include "gemm.cuh"
#include "ptx.cuh"

struct reidb200_gallery {
  int slots, budget, d, dp;
  float* feats;      // [slots, budget, d] L2-normalised rows (ring per slot)
  __half* packed;    // the same rows as fp16 hi|lo planes in TILE-IMAGE layout (img_* below): A operand of the frame kernels
  int rb;            // rows of one (slot, k-block, plane) region of `packed`: budget rounded up to 16
  int32_t* count;    // [slots] valid rows
  int32_t* head;     // [slots] next write position
  int32_t* scratch;  // [2*scratch_rows] per-row destination / rows-after
  int scratch_rows;
  __half* det_packed;  // L2-normalised detection features of the current frame, tile-image layout, regions of det_cap rows
  int det_cap;         // multiple of 16
  int32_t* sched;             // frame schedule written by assoc_schedule (tracks packed into 128-row tiles by 32-row quadrants)
  int sched_cap;              // tracks the schedule buffer holds
};

namespace reidb200 {

constexpr double INFTY_COST = 1e5;        // sort/linear_assignment.py:9
constexpr double CHI2INV95_4 = 9.4877;    // boxmot/utils/matching.py:12-22
// L2-normalised rows (|v| <= 1) are scaled by 2^14 before the fp16 hi|lo split (both planes stay in fp16's normal range,
// 22 significand bits survive); the frame kernel multiplies the accumulator by 2^-28.
constexpr float UNIT_SCALE = 16384.0f;
constexpr float UNIT_UNSCALE2 = 1.0f / (16384.0f * 16384.0f);

// TILE-IMAGE layout of a packed operand (gallery ring, frame detections).  For every k-block of 64 features and each of
// the two fp16 planes there is one contiguous REGION of `region_rows` rows x 128 bytes, and inside a region the 16-byte
// chunks of row r are XOR-swizzled by (r & 7): exactly the bytes TMA's SWIZZLE_128B would leave in shared memory for a
// K-major operand tile.  Any run of rows that starts at a multiple of 8 is therefore ONE contiguous piece of global memory
// that a single cp.async.bulk drops into an operand stage -- measured: a 128-row x 64-column TMA tensor box is 128 separate
// 128-byte requests (row pitch 2 KB), and at two tracks + 208 detections per stage the frame kernel spent ~2 us per
// k-block waiting for ~670 of them.
//   region index: gallery (slot * nkb + kb) * 2 + plane, detections kb * 2 + plane;  half index inside a region:
__host__ __device__ __forceinline__ size_t img_half(int row, int kk) {  // kk = feature index inside the k-block (0..63)
  return (size_t)row * 64 + (size_t)((((kk >> 3) ^ (row & 7)) << 3) | (kk & 7));
}

// ------------------------------------------------------------------------------------ gallery
// dst[i] = ring position of row i (or -1 when a later row of the same call overwrites it)
__global__ void pf_assign_kernel(const int32_t* __restrict__ slot_of_row, int n, int budget,
                                 const int32_t* __restrict__ head, int32_t* __restrict__ dst,
                                 int32_t* __restrict__ after) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int s = slot_of_row[i];
  int before = 0, aft = 0;
  for (int o = 0; o < n; ++o) {
    const bool same = slot_of_row[o] == s;
    before += same && o < i;
    aft += same && o > i;
  }
  after[i] = aft;
  dst[i] = aft < budget ? (head[s] + before) % budget : -1;
}

// warp per row: a / ||a|| in fp32 (matching.py:294-296), written to its ring position
__global__ void __launch_bounds__(256)
pf_write_kernel(const float* __restrict__ feats, int64_t ldf, const int32_t* __restrict__ slot_of_row,
                const int32_t* __restrict__ dst, int n, int budget, int d, int dp, int rb, float* __restrict__ gal,
                __half* __restrict__ packed) {
  const int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (i >= n) return;
  const int p = dst[i];
  if (p < 0) return;
  const float* src = feats + (size_t)i * ldf;
  float s = 0.f;
  for (int k = lane; k < d; k += 32) { const float v = src[k]; s = fmaf(v, v, s); }
  s = sqrtf(warp_sum(s));
  const int slot = slot_of_row[i];
  float* o = gal + ((size_t)slot * budget + p) * d;
  const int nkb = dp / 64;
  const size_t region = (size_t)rb * 64;
  for (int k = lane; k < dp; k += 32) {
    const float v = k < d ? __fdiv_rn(src[k], s) : 0.f;
    if (k < d) o[k] = v;
    const float vs = __fmul_rn(v, UNIT_SCALE);
    const __half h = __float2half_rn(vs);
    __half* hi = packed + ((size_t)(slot * nkb + (k >> 6)) * 2) * region + img_half(p, k & 63);
    hi[0] = h;
    hi[region] = __float2half_rn(__fsub_rn(vs, __half2float(h)));
  }
}

// warp per detection: f / ||f|| in fp32 (matching.py:294-296) split into fp16 hi|lo planes (B operand of the frame kernel)
__device__ void assoc_schedule_block(const int32_t* __restrict__ trk_slot, const int32_t* __restrict__ gcount, int T,
                                     int32_t* __restrict__ sched);

__global__ void __launch_bounds__(256)
det_pack_kernel(const float* __restrict__ feats, int64_t ldf, int n, int d, int dp, int det_rows, __half* __restrict__ packed,
                const int32_t* __restrict__ trk_slot, const int32_t* __restrict__ gcount, int T, int32_t* __restrict__ sched) {
  if (sched != nullptr && blockIdx.x == gridDim.x - 1) {  // the extra block: frame schedule of the packed-tile kernel
    assoc_schedule_block(trk_slot, gcount, T, sched);
    return;
  }
  const int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (i >= n) return;
  const float* src = feats + (size_t)i * ldf;
  float s = 0.f;
  for (int k = lane; k < d; k += 32) { const float v = src[k]; s = fmaf(v, v, s); }
  s = sqrtf(warp_sum(s));
  const size_t region = (size_t)det_rows * 64;
  for (int k = lane; k < dp; k += 32) {
    const float v = k < d ? __fmul_rn(__fdiv_rn(src[k], s), UNIT_SCALE) : 0.f;
    const __half h = __float2half_rn(v);
    __half* hi = packed + ((size_t)(k >> 6) * 2) * region + img_half(i, k & 63);
    hi[0] = h;
    hi[region] = __float2half_rn(__fsub_rn(v, __half2float(h)));
  }
}

__global__ void pf_update_kernel(const int32_t* __restrict__ slot_of_row, const int32_t* __restrict__ after, int n,
                                 int budget, int32_t* __restrict__ head, int32_t* __restrict__ count) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n || after[i] != 0) return;  // the last row of each slot updates the slot
  const int s = slot_of_row[i];
  int total = 0;
  for (int o = 0; o < n; ++o) total += slot_of_row[o] == s;
  head[s] = (head[s] + total) % budget;
  count[s] = min(budget, count[s] + total);
}

__global__ void gal_reset_kernel(const int32_t* __restrict__ slots, int n, int32_t* __restrict__ head,
                                 int32_t* __restrict__ count) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) { head[slots[i]] = 0; count[slots[i]] = 0; }
}

// rows of one slot, oldest first
__global__ void gal_read_kernel(const float* __restrict__ gal, const int32_t* __restrict__ head,
                                const int32_t* __restrict__ count, int slot, int budget, int d, float* __restrict__ out) {
  const int c = count[slot], h = head[slot];
  for (int r = blockIdx.x; r < budget; r += gridDim.x) {
    for (int k = threadIdx.x; k < d; k += blockDim.x) {
      float v = 0.f;
      if (r < c) v = gal[((size_t)slot * budget + ((h - c + r) % budget + budget) % budget) * d + k];
      out[(size_t)r * d + k] = v;
    }
  }
}

// copy of one gallery into a larger one (more slots and / or a longer ring): the rows of every slot move, oldest first,
// to ring positions 0..count-1 of the new layout (head = count mod new budget), bit for bit
__global__ void gal_resize_kernel(const float* __restrict__ src, const __half* __restrict__ src_packed,
                                  const int32_t* __restrict__ src_head, const int32_t* __restrict__ src_count, int old_budget,
                                  int old_rb, int new_budget, int new_rb, int d, int dp, float* __restrict__ dst,
                                  __half* __restrict__ dst_packed, int32_t* __restrict__ dst_head, int32_t* __restrict__ dst_count) {
  const int slot = blockIdx.x;
  const int c = src_count[slot], h = src_head[slot];
  const int nkb = dp / 64;
  const size_t reg_o = (size_t)old_rb * 64, reg_n = (size_t)new_rb * 64;
  for (int r = blockIdx.y; r < c; r += gridDim.y) {
    const int from = ((h - c + r) % old_budget + old_budget) % old_budget;
    for (int k = threadIdx.x; k < d; k += blockDim.x)
      dst[((size_t)slot * new_budget + r) * d + k] = src[((size_t)slot * old_budget + from) * d + k];
    for (int k = threadIdx.x; k < 2 * dp; k += blockDim.x) {  // both planes: the row moves to ring position r (re-swizzled)
      const int plane = k >= dp ? 1 : 0, kf = k - plane * dp;
      const size_t ro = (size_t)(slot * nkb + (kf >> 6)) * 2 + plane;
      dst_packed[ro * reg_n + img_half(r, kf & 63)] = src_packed[ro * reg_o + img_half(from, kf & 63)];
    }
  }
  if (blockIdx.y == 0 && threadIdx.x == 0) { dst_count[slot] = c; dst_head[slot] = c % new_budget; }
}

// ------------------------------------------------------------------------------------ box / gate math
// Detection.to_xyah in float32 (sort/detection.py:42-50), promoted to double by z - Hx
__device__ __forceinline__ void tlwh_to_xyah_f32(const float* t, double z[4]) {
  const float x = __fadd_rn(t[0], __fdiv_rn(t[2], 2.0f));
  const float y = __fadd_rn(t[1], __fdiv_rn(t[3], 2.0f));
  const float a = __fdiv_rn(t[2], t[3]);
  z[0] = (double)x; z[1] = (double)y; z[2] = (double)a; z[3] = (double)t[3];
}

// sqrt(y^T SI y) with y = z - x[:4] (kalman_filter.py:577-586)
__device__ __forceinline__ double maha(const double z[4], const double* x4, const double* SI) {
  double y[4];
#pragma unroll
  for (int r = 0; r < 4; ++r) y[r] = z[r] - x4[r];
  double acc = 0.0;
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    double v = 0.0;
#pragma unroll
    for (int r = 0; r < 4; ++r) v += y[r] * SI[r * 4 + c];
    acc += v * y[c];
  }
  return sqrt(acc);
}

// Track.to_tlwh from the KF mean (sort/track.py:107-120), fp64
__device__ __forceinline__ void mean_to_tlwh(const double* mean, double b[4]) {
  b[2] = mean[2] * mean[3];
  b[3] = mean[3];
  b[0] = mean[0] - b[2] / 2;
  b[1] = mean[1] - b[3] / 2;
}

// iou / DIoU of a f64 tlwh box against one candidate whose own arithmetic runs in its dtype CT
// (numpy: candidates[:, :2] + candidates[:, 2:] and the area product stay float32 for float32 input).
template <typename CT>
__device__ __forceinline__ double iou_one(const double b[4], const CT* c, int mode) {
  const double b_tlx = b[0], b_tly = b[1], b_brx = b[0] + b[2], b_bry = b[1] + b[3];
  const CT cbx = c[0] + c[2], cby = c[1] + c[3];
  const CT carea = c[2] * c[3];
  const double c_tlx = (double)c[0], c_tly = (double)c[1], c_brx = (double)cbx, c_bry = (double)cby;
  const double tlx = fmax(b_tlx, c_tlx), tly = fmax(b_tly, c_tly);
  const double brx = fmin(b_brx, c_brx), bry = fmin(b_bry, c_bry);
  const double w = fmax(0.0, brx - tlx), h = fmax(0.0, bry - tly);
  const double inter = w * h;
  const double area_b = b[2] * b[3];
  const double iou = inter / (area_b + (double)carea - inter);
  if (mode == REIDB200_IOU_PLAIN) return iou;
  // DIoU: centre distance over outer diagonal (modification_deepsort/iou_matching.py:26-46)
  const double bcx = (b_tly + b_bry) / 2, bcy = (b_tlx + b_brx) / 2;
  const double ccx = (c_tly + c_bry) / 2, ccy = (c_tlx + c_brx) / 2;
  const double dd = (bcx - ccx) * (bcx - ccx) + (bcy - ccy) * (bcy - ccy);
  const double otlx = fmin(b_tlx, c_tlx), otly = fmin(b_tly, c_tly);
  const double obrx = fmax(b_brx, c_brx), obry = fmax(b_bry, c_bry);
  const double rou = (otlx - obrx) * (otlx - obrx) + (otly - obry) * (otly - obry);
  return iou - dd / rou;
}

// ------------------------------------------------------------------------------------ fused frame kernel
// CTA = one track.  Appearance: 64 gallery rows x 64 detections per tile, 4x4 register micro-tile,
// k-chunks of 16 staged k-major in shared memory; the running per-detection maximum of the dot product
// stays in registers over the row tiles (min(1-dot) == 1-max(dot) exactly).
constexpr int AS_THREADS = 256;
constexpr int AS_TR = 64, AS_TD = 64, AS_TK = 16;

struct AssocArgs {
  const float* gal; const int32_t* gcount; const int32_t* ghead; int budget, d;
  const float* det_feat; int64_t ldf; const float* det_tlwh; int M;
  const int32_t* trk_slot; const double* trk_mean; const double* trk_SI; const int32_t* trk_tsu; int T;
  double mc_lambda, max_dist, max_iou_dist; int iou_mode, flags;
  double* out_app; double* out_raw; double* out_iou;
};

__global__ void __launch_bounds__(AS_THREADS)
assoc_frame_simt_kernel(AssocArgs a) {
  __shared__ float Gs[AS_TK][AS_TR + 4];
  __shared__ float Fs[AS_TK][AS_TD + 4];
  __shared__ float red[16][AS_TD + 1];
  __shared__ float fsq[AS_TD];
  __shared__ double s_x4[4], s_SI[16], s_box[4];

  const int t = blockIdx.x;
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;  // tx -> detections, ty -> gallery rows
  const int slot = a.trk_slot ? a.trk_slot[t] : -1;
  const int cnt = (slot >= 0 && a.gal) ? a.gcount[slot] : 0;
  const float* G = a.gal + (size_t)(slot >= 0 ? slot : 0) * a.budget * a.d;
  if (tid < 4) s_x4[tid] = a.trk_mean[(size_t)t * 8 + tid];
  if (tid < 16) s_SI[tid] = a.trk_SI ? a.trk_SI[(size_t)t * 16 + tid] : 0.0;
  if (tid == 0) mean_to_tlwh(a.trk_mean + (size_t)t * 8, s_box);
  __syncthreads();
  const bool tsu_far = a.trk_tsu[t] > 1;

  for (int d0 = 0; d0 < a.M; d0 += AS_TD) {
    float best[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
    if (cnt > 0 && a.out_app) {
      for (int r0 = 0; r0 < cnt; r0 += AS_TR) {
        float acc[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
        float sq = 0.f;
        for (int k0 = 0; k0 < a.d; k0 += AS_TK) {
          // 64 rows x 16 k = 1024 elements per operand: 4 per thread (row = tid/4, k = (tid%4)*4..+3)
          const int lr = tid >> 2, lk = (tid & 3) * 4;
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int k = k0 + lk + e;
            const int gr = r0 + lr, dm = d0 + lr;
            Gs[lk + e][lr] = (gr < cnt && k < a.d) ? G[(size_t)gr * a.d + k] : 0.f;
            const float fv = (dm < a.M && k < a.d) ? a.det_feat[(size_t)dm * a.ldf + k] : 0.f;
            Fs[lk + e][lr] = fv;
            sq = fmaf(fv, fv, sq);
          }
          __syncthreads();
#pragma unroll
          for (int k = 0; k < AS_TK; ++k) {
            float g[4], f[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) g[i] = Gs[k][ty * 4 + i];
#pragma unroll
            for (int j = 0; j < 4; ++j) f[j] = Fs[k][tx * 4 + j];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
              for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(g[i], f[j], acc[i][j]);
          }
          __syncthreads();
        }
        if (r0 == 0) {  // |f|^2 of detection d0 + tid/4: partial sums sit in 4 consecutive lanes
          sq += __shfl_xor_sync(0xffffffffu, sq, 1);
          sq += __shfl_xor_sync(0xffffffffu, sq, 2);
          if ((tid & 3) == 0) fsq[tid >> 2] = sq;
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          if (r0 + ty * 4 + i < cnt) {
#pragma unroll
            for (int j = 0; j < 4; ++j) best[j] = fmaxf(best[j], acc[i][j]);
          }
        }
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) red[ty][tx * 4 + j] = best[j];
    __syncthreads();
    if (tid < AS_TD) {
      const int m = d0 + tid;
      if (m < a.M) {
        const size_t o = (size_t)t * a.M + m;
        // ---- appearance + gate (linear_assignment.py:147-198)
        if (a.out_app) {
          double cost = INFTY_COST, raw = INFTY_COST;
          if (cnt > 0) {
            float mx = red[0][tid];
#pragma unroll
            for (int r = 1; r < 16; ++r) mx = fmaxf(mx, red[r][tid]);
            const float dist = __fsub_rn(1.0f, __fdiv_rn(mx, sqrtf(fsq[tid])));
            raw = (double)dist;
            cost = raw;
          }
          if (a.out_raw) a.out_raw[o] = raw;
          if (!(a.flags & REIDB200_ASSOC_NO_GATE)) {
            double z[4];
            tlwh_to_xyah_f32(a.det_tlwh + (size_t)m * 4, z);
            const double gd = maha(z, s_x4, s_SI);
            if (gd > CHI2INV95_4) cost = INFTY_COST;
            cost = a.mc_lambda * cost + (1 - a.mc_lambda) * gd;
          }
          if ((a.flags & REIDB200_ASSOC_CLAMP_APP) && cost > a.max_dist) cost = a.max_dist + 1e-5;
          a.out_app[o] = cost;
        }
        // ---- IoU row (iou_matching.py:77-86)
        if (a.out_iou) {
          double c = INFTY_COST;
          if (!tsu_far) c = 1.0 - iou_one<float>(s_box, a.det_tlwh + (size_t)m * 4, a.iou_mode);
          if ((a.flags & REIDB200_ASSOC_CLAMP_IOU) && c > a.max_iou_dist) c = a.max_iou_dist + 1e-5;
          a.out_iou[o] = c;
        }
      }
    }
    __syncthreads();
  }
}


// ------------------------------------------------------------------------------------ fused frame kernel, tcgen05
// Persistent, one CTA per SM, work item = (track, 256-detection tile).  The track's ring rows (<= 128 per row tile,
// rows >= count masked) are the A operand, the frame's normalised detections the B operand, both as fp16 hi|lo planes
// (fp16, three UMMAs per k-step: lo.hi + hi.lo + hi.hi, fp32 accumulation in TMEM).  Warp 0 = TMA producer, warp 1 = MMA
// issuer, warps 2-5 = epilogue: TMEM -> registers (thread = gallery row), masked column maximum by a transposing
// butterfly (31 shuffles per 32 x 32 block), then -- thread = detection -- 1 - max, Mahalanobis gate + blend + clamp and
// the IoU / DIoU row in fp64, exactly as the SIMT kernel above does.
constexpr int AT_BM = 128, AT_BN = 256, AT_BK = 64, AT_STAGES = 2;
constexpr int AT_PLANE = AT_BM * AT_BK * 2;                             // 16 KB: one fp16 plane of an A stage
constexpr int AT_STAGE_BYTES = 2 * AT_PLANE + 2 * AT_BN * AT_BK * 2;     // 96 KB (B planes sized for 256 detections)
constexpr int AT_THREADS = 192;
constexpr size_t AT_SMEM = 1024 + (size_t)AT_STAGES * AT_STAGE_BYTES + 2 * 4 * AT_BN * sizeof(float) + 2 * 32 * sizeof(double) + 128;

// v[j] = value of column j in this lane's row  ->  returns max over the 32 lanes (rows) of column `lane`
__device__ __forceinline__ float colmax_transpose(float (&v)[32], int lane) {
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    const bool up = (lane & off) != 0;
#pragma unroll
    for (int i = 0; i < off; ++i) {
      const float send = up ? v[i] : v[i + off];
      const float keep = up ? v[i + off] : v[i];
      v[i] = fmaxf(keep, __shfl_xor_sync(0xffffffffu, send, off));
    }
  }
  return v[0];
}

__global__ void __launch_bounds__(AT_THREADS, 1)
assoc_frame_kernel(const __half* __restrict__ gpacked, int rb, const __half* __restrict__ dpacked, int det_rows, AssocArgs a, int dp,
                   int n_dtiles, int bn_box) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  float* part = reinterpret_cast<float*>(smem + (size_t)AT_STAGES * AT_STAGE_BYTES);  // [2][4][AT_BN] per-quadrant column maxima
  double* trk = reinterpret_cast<double*>(part + 2 * 4 * AT_BN);                       // [2][32] x4 | SI | tlwh box
  uint64_t* bars = reinterpret_cast<uint64_t*>(trk + 2 * 32);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * AT_STAGES + 4);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;
  const uint32_t bar_full = ptx::smem_u32(bars);
  const uint32_t bar_empty = ptx::smem_u32(bars + AT_STAGES);
  const uint32_t bar_tfull = ptx::smem_u32(bars + 2 * AT_STAGES);
  const uint32_t bar_tempty = ptx::smem_u32(bars + 2 * AT_STAGES + 2);
  const uint32_t smem_base = ptx::smem_u32(smem);
  const int num_kb = dp / AT_BK;
  const uint32_t b_plane = (uint32_t)bn_box * AT_BK * 2;

  if (threadIdx.x == 0) {
    for (int s = 0; s < AT_STAGES; ++s) {
      ptx::mbar_init(bar_full + 8 * s, 1);
      ptx::mbar_init(bar_empty + 8 * s, 1);
    }
    for (int s = 0; s < 2; ++s) {
      ptx::mbar_init(bar_tfull + 8 * s, 1);
      ptx::mbar_init(bar_tempty + 8 * s, 4);
    }
    ptx::fence_barrier_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(ptx::smem_u32(tmem_slot), 512);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int total = a.T * n_dtiles;
  const size_t greg = (size_t)rb * 64, dreg = (size_t)det_rows * 64;  // halfs per (k-block, plane) region

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int w = blockIdx.x; w < total; w += gridDim.x) {
        const int t = w / n_dtiles, dt = w - t * n_dtiles;
        const int slot = a.trk_slot[t];
        const int cnt = slot >= 0 ? a.gcount[slot] : 0;
        const int rts = (cnt + AT_BM - 1) / AT_BM;
        const int nb_rows = min(bn_box, (a.M - dt * AT_BN + 15) & ~15);  // detection rows of this tile (multiple of 16)
        for (int rt = 0; rt < rts; ++rt) {
          const int na_rows = min(AT_BM, (cnt - rt * AT_BM + 15) & ~15);  // ring rows fetched: the valid ones, rounded up to 16
          const uint32_t a_bytes = (uint32_t)na_rows * 128u, b_bytes = (uint32_t)nb_rows * 128u;
          for (int kb = 0; kb < num_kb; ++kb) {
            ptx::mbar_wait(bar_empty + 8 * stage, phase ^ 1);
            const uint32_t full = bar_full + 8 * stage;
            const uint32_t sb = smem_base + stage * AT_STAGE_BYTES;
            ptx::mbar_arrive_expect_tx(full, 2 * a_bytes + 2 * b_bytes);
            const __half* ga = gpacked + ((size_t)(slot * num_kb + kb) * 2) * greg + (size_t)rt * AT_BM * 64;
            const __half* gd = dpacked + ((size_t)kb * 2) * dreg + (size_t)dt * AT_BN * 64;
            ptx::bulk_load(sb, ga, a_bytes, full);
            ptx::bulk_load(sb + AT_PLANE, ga + greg, a_bytes, full);
            ptx::bulk_load(sb + 2 * AT_PLANE, gd, b_bytes, full);
            ptx::bulk_load(sb + 2 * AT_PLANE + b_plane, gd + dreg, b_bytes, full);
            if (++stage == AT_STAGES) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      int stage = 0, as = 0;
      uint32_t phase = 0, aphase = 0;
      for (int w = blockIdx.x; w < total; w += gridDim.x) {
        const int t = w / n_dtiles, dt = w - t * n_dtiles;
        const int slot = a.trk_slot[t];
        const int cnt = slot >= 0 ? a.gcount[slot] : 0;
        const int rts = (cnt + AT_BM - 1) / AT_BM;
        const int n_tile = min(bn_box, (a.M - dt * AT_BN + 15) & ~15);
        const uint32_t idesc = ptx::umma_idesc_f16(AT_BM, n_tile);
        for (int rt = 0; rt < rts; ++rt) {
          ptx::mbar_wait(bar_tempty + 8 * as, aphase ^ 1);
          ptx::tc_fence_after();
          const uint32_t d_tmem = tmem_base + as * AT_BN;
          for (int kb = 0; kb < num_kb; ++kb) {
            ptx::mbar_wait(bar_full + 8 * stage, phase);
            ptx::tc_fence_after();
            const uint32_t sb = smem_base + stage * AT_STAGE_BYTES;
#pragma unroll
            for (int kk = 0; kk < AT_BK / 16; ++kk) {
              const uint64_t ahi = ptx::umma_desc_k_sw128(sb + kk * 32);
              const uint64_t alo = ptx::umma_desc_k_sw128(sb + AT_PLANE + kk * 32);
              const uint64_t bhi = ptx::umma_desc_k_sw128(sb + 2 * AT_PLANE + kk * 32);
              const uint64_t blo = ptx::umma_desc_k_sw128(sb + 2 * AT_PLANE + b_plane + kk * 32);
              ptx::mma_f16_ss(d_tmem, alo, bhi, idesc, (kb | kk) != 0);
              ptx::mma_f16_ss(d_tmem, ahi, blo, idesc, 1);
              ptx::mma_f16_ss(d_tmem, ahi, bhi, idesc, 1);
            }
            ptx::mma_commit(bar_empty + 8 * stage);
            if (++stage == AT_STAGES) { stage = 0; phase ^= 1; }
          }
          ptx::mma_commit(bar_tfull + 8 * as);
          as ^= 1;
          if (as == 0) aphase ^= 1;
        }
      }
    }
  } else {
    const int q = warp & 3;               // TMEM lane quadrant of this warp
    const int e = threadIdx.x - 64;       // 0..127: finalize index (detection column within the tile)
    int as = 0, it = 0;
    uint32_t aphase = 0;
    for (int w = blockIdx.x; w < total; w += gridDim.x, ++it) {
      const int t = w / n_dtiles, dt = w - t * n_dtiles;
      const int slot = a.trk_slot[t];
      const int cnt = slot >= 0 ? a.gcount[slot] : 0;
      const int rts = (cnt + AT_BM - 1) / AT_BM;
      const int n_cols = min(AT_BN, a.M - dt * AT_BN);
      float* mypart = part + ((it & 1) * 4 + q) * AT_BN;
      double* mytrk = trk + (it & 1) * 32;
      if (e < 4) mytrk[e] = a.trk_mean[(size_t)t * 8 + e];
      else if (e < 20) mytrk[e] = a.trk_SI ? a.trk_SI[(size_t)t * 16 + (e - 4)] : 0.0;
      else if (e == 20) mean_to_tlwh(a.trk_mean + (size_t)t * 8, mytrk + 20);
      for (int rt = 0; rt < rts; ++rt) {
        ptx::mbar_wait(bar_tfull + 8 * as, aphase);
        ptx::tc_fence_after();
        const bool valid = rt * AT_BM + q * 32 + lane < cnt;
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + as * AT_BN;
#pragma unroll 1
        for (int c = 0; c * 32 < n_cols; ++c) {
          uint32_t r[32];
          ptx::tmem_ld_32x32b_x32(taddr + c * 32, r);
          ptx::tmem_ld_wait();
          float v[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = valid ? __uint_as_float(r[j]) : -INFINITY;
          const float mx = colmax_transpose(v, lane);
          float* dst = mypart + c * 32 + lane;
          *dst = rt == 0 ? mx : fmaxf(*dst, mx);
        }
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(bar_tempty + 8 * as);
        as ^= 1;
        if (as == 0) aphase ^= 1;
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");  // the four epilogue warps: partial maxima + track state visible
      const bool tsu_far = a.trk_tsu[t] > 1;
      const float* p0 = part + (it & 1) * 4 * AT_BN;
      for (int col = e; col < n_cols; col += 128) {
        const int m = dt * AT_BN + col;
        const size_t o = (size_t)t * a.M + m;
        if (a.out_app) {
          double cost = INFTY_COST, raw = INFTY_COST;
          if (cnt > 0) {
            const float mx = fmaxf(fmaxf(p0[col], p0[AT_BN + col]), fmaxf(p0[2 * AT_BN + col], p0[3 * AT_BN + col]));
            raw = (double)__fsub_rn(1.0f, __fmul_rn(mx, UNIT_UNSCALE2));
            cost = raw;
          }
          if (a.out_raw) a.out_raw[o] = raw;
          if (!(a.flags & REIDB200_ASSOC_NO_GATE)) {
            double z[4];
            tlwh_to_xyah_f32(a.det_tlwh + (size_t)m * 4, z);
            const double gd = maha(z, mytrk, mytrk + 4);
            if (gd > CHI2INV95_4) cost = INFTY_COST;
            cost = a.mc_lambda * cost + (1 - a.mc_lambda) * gd;
          }
          if ((a.flags & REIDB200_ASSOC_CLAMP_APP) && cost > a.max_dist) cost = a.max_dist + 1e-5;
          a.out_app[o] = cost;
        }
        if (a.out_iou) {
          double c = INFTY_COST;
          if (!tsu_far) c = 1.0 - iou_one<float>(mytrk + 20, a.det_tlwh + (size_t)m * 4, a.iou_mode);
          if ((a.flags & REIDB200_ASSOC_CLAMP_IOU) && c > a.max_iou_dist) c = a.max_iou_dist + 1e-5;
          a.out_iou[o] = c;
        }
      }
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, 512);
  }
}

// ------------------------------------------------------------------------------------ fused frame kernel, packed tiles
// Second form of the tcgen05 frame kernel, for rings of at most 128 rows (StrongSORT's nn_budget is 100): a frame's tracks
// are PACKED into 128-row A tiles by 32-row quadrants (a track with `count` rows takes ceil(count / 32) consecutive
// quadrants = TMEM lane quadrants = epilogue warps), and only ceil16(count) rows of every track are fetched -- TMA boxes
// of 16 / 32 / 64 / 128 rows written at the track's row offset inside the stage; the MMA still runs M = 128 over whatever
// the untouched rows hold, and those accumulator rows are masked.  Against one track per tile this halves the tiles (and
// with them the re-streaming of the detection operand from L2) and reads the ring's valid rows only: DRAM traffic ~ sum of
// count x 2 KB instead of the whole [slots, budget] ring.
// Schedule (device side, assoc_schedule_block): sched[0] = tiles in total, sched[1] = tiles with rows;
// sched[4 + 4 * tile + q] = track row index of quadrant q (or -1).  Tiles >= sched[1] list the tracks WITHOUT rows
// (unknown slot / empty ring: every cost is 1e5 before the gate), four per tile, no MMA.
constexpr int A2_MAX_STAGES = 4;
constexpr size_t A2_STAGE_BUDGET = 196 * 1024;  // operand stages share this much shared memory
constexpr size_t A2_SMEM = 1024 + A2_STAGE_BUDGET + 2 * 4 * AT_BN * sizeof(float) + 2 * 4 * 32 * sizeof(double) + 128;

// The schedule is computed by ONE extra block of the detection-packing launch (det_pack_kernel), concurrently with the
// packing blocks, fully in parallel: a track's place is a closed form of its CLASS (quadrants needed: 4, 3, 2, 1, none)
// and its rank inside the class --
//   class 4: a tile each;  class 3: a tile each, quadrants 0-2, the fourth quadrant taken by one class-1 track while there
//   are any;  class 2: two per tile;  remaining class 1: four per tile;  tracks without rows: four per row-less tile --
// so one block-wide exclusive scan of the five class counts places every track (no serial pass: a single thread walking
// 300 tracks took 55 us).
constexpr int SCHED_THREADS = 256;

__device__ void assoc_schedule_block(const int32_t* __restrict__ trk_slot, const int32_t* __restrict__ gcount, int T,
                                     int32_t* __restrict__ sched) {
  __shared__ int s_warp[SCHED_THREADS / 32][5];
  __shared__ int s_tot[5];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  auto cls = [&](int t) {  // quadrants of track t: 0 (no rows) .. 4
    const int slot = trk_slot ? trk_slot[t] : -1;
    const int cnt = slot >= 0 ? gcount[slot] : 0;
    return min((cnt + 31) >> 5, 4);
  };
  // tracks are dealt out in contiguous chunks: thread `tid` owns [t0, t1)
  const int per = (T + SCHED_THREADS - 1) / SCHED_THREADS;
  const int t0 = min(T, tid * per), t1 = min(T, t0 + per);
  int c[5] = {0, 0, 0, 0, 0};
  for (int t = t0; t < t1; ++t) {
    const int k = cls(t);
#pragma unroll
    for (int j = 0; j < 5; ++j) c[j] += (k == j) ? 1 : 0;
  }
  int base[5];
#pragma unroll
  for (int j = 0; j < 5; ++j) {
    int incl = c[j];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    base[j] = incl - c[j];
    if (lane == 31) s_warp[warp][j] = incl;
  }
  __syncthreads();
  if (tid < 5) {
    int tot = 0;
    for (int w = 0; w < SCHED_THREADS / 32; ++w) { const int v = s_warp[w][tid]; s_warp[w][tid] = tot; tot += v; }
    s_tot[tid] = tot;
  }
  __syncthreads();
  const int n0 = s_tot[0], n1 = s_tot[1], n2 = s_tot[2], n3 = s_tot[3], n4 = s_tot[4];
  const int rest1 = max(n1 - n3, 0);
  const int tile2 = n4 + n3, tile1 = tile2 + (n2 + 1) / 2, n_real = tile1 + (rest1 + 3) / 4, n_tiles = n_real + (n0 + 3) / 4;
  int32_t* seg = sched + 4;
  for (int i = tid; i < 4 * n_tiles; i += SCHED_THREADS) seg[i] = -1;
  __syncthreads();
  int r[5];
#pragma unroll
  for (int j = 0; j < 5; ++j) r[j] = base[j] + s_warp[warp][j];
  for (int t = t0; t < t1; ++t) {
    const int k = cls(t);
    if (k == 4) {
      const int tl = r[4]++;
#pragma unroll
      for (int q = 0; q < 4; ++q) seg[tl * 4 + q] = t;
    } else if (k == 3) {
      const int tl = n4 + r[3]++;
#pragma unroll
      for (int q = 0; q < 3; ++q) seg[tl * 4 + q] = t;
    } else if (k == 2) {
      const int i = r[2]++;
      const int tl = tile2 + (i >> 1), q = (i & 1) * 2;
      seg[tl * 4 + q] = t;
      seg[tl * 4 + q + 1] = t;
    } else if (k == 1) {
      const int i = r[1]++;
      if (i < n3) seg[(n4 + i) * 4 + 3] = t;   // the free quadrant of a class-3 tile
      else seg[(tile1 + ((i - n3) >> 2)) * 4 + ((i - n3) & 3)] = t;
    } else {
      const int i = r[0]++;
      seg[(n_real + (i >> 2)) * 4 + (i & 3)] = t;
    }
  }
  if (tid == 0) { sched[0] = n_tiles; sched[1] = n_real; }
}

constexpr int A2_THREADS = 320;  // TMA producer warp, MMA warp, 4 column-max warps, 4 finaliser warps

__global__ void __launch_bounds__(A2_THREADS, 1)
assoc_frame2_kernel(const __half* __restrict__ gpacked, int rb, const __half* __restrict__ dpacked, int det_rows, AssocArgs a, int dp,
                    int n_dtiles, int bn_box, int nst, const int32_t* __restrict__ sched, int debug) {
  // bn_box = detection columns of one work item (multiple of 16, <= AT_BN): a frame's detections are split into n_dtiles
  // column ranges so that every SM gets two or three items of which the epilogue of one overlaps the k-loop of the next
  // (one 128 x 208 item per SM left its epilogue and the launch ramp fully exposed); nst = operand stages (2..4) that fit.
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const uint32_t b_plane = (uint32_t)bn_box * AT_BK * 2;
  const uint32_t stage_bytes = 2 * AT_PLANE + 2 * b_plane;  // multiple of 1024: bn_box is a multiple of 16
  float* part = reinterpret_cast<float*>(smem + (size_t)nst * stage_bytes);  // [2][4][AT_BN] per-quadrant column maxima
  double* trk = reinterpret_cast<double*>(part + 2 * 4 * AT_BN);                       // [2][4][32] x4 | SI | tlwh box per segment
  uint64_t* bars = reinterpret_cast<uint64_t*>(trk + 2 * 4 * 32);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * A2_MAX_STAGES + 4);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;
  const uint32_t bar_full = ptx::smem_u32(bars);
  const uint32_t bar_empty = ptx::smem_u32(bars + A2_MAX_STAGES);
  const uint32_t bar_tfull = ptx::smem_u32(bars + 2 * A2_MAX_STAGES);
  const uint32_t bar_tempty = ptx::smem_u32(bars + 2 * A2_MAX_STAGES + 2);
  const uint32_t smem_base = ptx::smem_u32(smem);
  const int num_kb = dp / AT_BK;

  if (threadIdx.x == 0) {
    for (int s = 0; s < nst; ++s) {
      ptx::mbar_init(bar_full + 8 * s, 1);
      ptx::mbar_init(bar_empty + 8 * s, 1);
    }
    for (int s = 0; s < 2; ++s) {
      ptx::mbar_init(bar_tfull + 8 * s, 1);
      ptx::mbar_init(bar_tempty + 8 * s, 4);
    }
    ptx::fence_barrier_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(ptx::smem_u32(tmem_slot), 512);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int n_tiles = sched[0], n_real = sched[1];
  const int32_t* seg = sched + 4;
  const int total = n_tiles * n_dtiles;
  const size_t greg = (size_t)rb * 64, dreg = (size_t)det_rows * 64;  // halfs per (k-block, plane) region

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int w = blockIdx.x; w < total; w += gridDim.x) {
        const int tile = w / n_dtiles, dt = w - tile * n_dtiles;
        if (tile >= n_real) continue;
        // the tile's tracks: gallery slot, rows to fetch (ceil16(count)) and row offset inside the tile
        int gslot[4], nrow[4], roff[4], nseg = 0;
        uint32_t a_bytes = 0;
        for (int q = 0; q < 4; ++q) {
          const int t = seg[tile * 4 + q];
          if (t < 0 || (q > 0 && seg[tile * 4 + q - 1] == t)) continue;
          const int slot = a.trk_slot[t];
          const int r16 = (a.gcount[slot] + 15) & ~15;
          gslot[nseg] = slot; nrow[nseg] = r16; roff[nseg] = q * 32;
          a_bytes += (uint32_t)r16 * 128u;
          ++nseg;
        }
        const int nb_rows = min(bn_box, (a.M - dt * bn_box + 15) & ~15);
        const uint32_t b_bytes = (uint32_t)nb_rows * 128u;
        // (debug: timing experiments only, REIDB200_ASSOC_DEBUG -- 1 = no gallery loads, 2 = no detection loads)
        const uint32_t stage_tx = ((debug & 1) ? 0u : 2 * a_bytes) + ((debug & 2) ? 0u : 2 * b_bytes);
        for (int kb = 0; kb < num_kb; ++kb) {
          ptx::mbar_wait(bar_empty + 8 * stage, phase ^ 1);
          const uint32_t full = bar_full + 8 * stage;
          const uint32_t sb = smem_base + stage * stage_bytes;
          if (stage_tx) ptx::mbar_arrive_expect_tx(full, stage_tx); else ptx::mbar_arrive(full);
          for (int s = 0; s < nseg && !(debug & 1); ++s) {  // one contiguous piece per track and plane (tile-image layout)
            const __half* ga = gpacked + ((size_t)(gslot[s] * num_kb + kb) * 2) * greg;
            const uint32_t dst = sb + (uint32_t)roff[s] * 128u, bytes = (uint32_t)nrow[s] * 128u;
            ptx::bulk_load(dst, ga, bytes, full);
            ptx::bulk_load(dst + AT_PLANE, ga + greg, bytes, full);
          }
          const __half* gd = dpacked + ((size_t)kb * 2) * dreg + (size_t)dt * bn_box * 64;
          if (!(debug & 2)) {
            ptx::bulk_load(sb + 2 * AT_PLANE, gd, b_bytes, full);
            ptx::bulk_load(sb + 2 * AT_PLANE + b_plane, gd + dreg, b_bytes, full);
          }
          if (++stage == nst) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      int stage = 0, as = 0;
      uint32_t phase = 0, aphase = 0;
      for (int w = blockIdx.x; w < total; w += gridDim.x) {
        const int tile = w / n_dtiles, dt = w - tile * n_dtiles;
        if (tile >= n_real) continue;
        const int n_tile = min(bn_box, (a.M - dt * bn_box + 15) & ~15);
        const uint32_t idesc = ptx::umma_idesc_f16(AT_BM, n_tile);
        ptx::mbar_wait(bar_tempty + 8 * as, aphase ^ 1);
        ptx::tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * AT_BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          ptx::mbar_wait(bar_full + 8 * stage, phase);
          ptx::tc_fence_after();
          const uint32_t sb = smem_base + stage * stage_bytes;
#pragma unroll
          for (int kk = 0; kk < AT_BK / 16; ++kk) {
            const uint64_t ahi = ptx::umma_desc_k_sw128(sb + kk * 32);
            const uint64_t alo = ptx::umma_desc_k_sw128(sb + AT_PLANE + kk * 32);
            const uint64_t bhi = ptx::umma_desc_k_sw128(sb + 2 * AT_PLANE + kk * 32);
            const uint64_t blo = ptx::umma_desc_k_sw128(sb + 2 * AT_PLANE + b_plane + kk * 32);
            ptx::mma_f16_ss(d_tmem, alo, bhi, idesc, (kb | kk) != 0);
            ptx::mma_f16_ss(d_tmem, ahi, blo, idesc, 1);
            ptx::mma_f16_ss(d_tmem, ahi, bhi, idesc, 1);
          }
          ptx::mma_commit(bar_empty + 8 * stage);
          if (++stage == nst) { stage = 0; phase ^= 1; }
        }
        ptx::mma_commit(bar_tfull + 8 * as);
        as ^= 1;
        if (as == 0) aphase ^= 1;
      }
    }
  } else if (warp < 6) {
    // ===== column-max warps (2-5): TMEM lane quadrant q = quadrant of the tile.  They reduce the accumulator to per-
    // quadrant column maxima in `part[buf]`, stage the quadrant's track state in `trk[buf]`, release the TMEM stage and
    // hand the buffer to the finaliser warps (named barriers 1+buf "filled", 3+buf "drained"; two buffers).
    const int q = warp & 3;
    int as = 0, it = 0;
    uint32_t aphase = 0;
    for (int w = blockIdx.x; w < total; w += gridDim.x, ++it) {
      const int tile = w / n_dtiles, dt = w - tile * n_dtiles;
      const bool real = tile < n_real;
      const int n_cols = min(bn_box, a.M - dt * bn_box);
      const int4 sg4 = *reinterpret_cast<const int4*>(seg + tile * 4);
      auto SG = [&](int i) { return i == 0 ? sg4.x : (i == 1 ? sg4.y : (i == 2 ? sg4.z : sg4.w)); };  // (no local array)
      const int t_mine = SG(q);
      const int buf = it & 1;
      float* mypart = part + (buf * 4 + q) * AT_BN;
      double* mytrk = trk + (buf * 4 + q) * 32;
      if (it >= 2) {  // the finalisers have drained this buffer (its previous occupant was item it - 2)
        if (buf == 0) asm volatile("bar.sync 3, 256;" ::: "memory"); else asm volatile("bar.sync 4, 256;" ::: "memory");
      }
      if (t_mine >= 0) {
        if (lane < 4) mytrk[lane] = a.trk_mean[(size_t)t_mine * 8 + lane];
        else if (lane < 20) mytrk[lane] = a.trk_SI ? a.trk_SI[(size_t)t_mine * 16 + (lane - 4)] : 0.0;
        else if (lane == 20) mean_to_tlwh(a.trk_mean + (size_t)t_mine * 8, mytrk + 20);
      }
      if (real) {
        int chunk = 0, cnt = 0;
        if (t_mine >= 0) {
#pragma unroll
          for (int p = 0; p < 4; ++p) chunk += (p < q && SG(p) == t_mine) ? 1 : 0;
          cnt = a.gcount[a.trk_slot[t_mine]];
        }
        ptx::mbar_wait(bar_tfull + 8 * as, aphase);
        ptx::tc_fence_after();
        const bool valid = t_mine >= 0 && chunk * 32 + lane < cnt;
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + as * AT_BN;
        if (t_mine >= 0) {  // warp-uniform
          const int nch = (n_cols + 31) >> 5;
          uint32_t rn[32];
          ptx::tmem_ld_32x32b_x32(taddr, rn);
#pragma unroll 1
          for (int c = 0; c < nch; ++c) {  // the next chunk's accumulator columns are requested before this one is reduced
            float v[32];
            ptx::tmem_ld_wait(rn);
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = valid ? __uint_as_float(rn[j]) : -INFINITY;
            if (c + 1 < nch) ptx::tmem_ld_32x32b_x32(taddr + (c + 1) * 32, rn);
            mypart[c * 32 + lane] = colmax_transpose(v, lane);
          }
        }
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(bar_tempty + 8 * as);
        as ^= 1;
        if (as == 0) aphase ^= 1;
      }
      if (buf == 0) asm volatile("bar.arrive 1, 256;" ::: "memory"); else asm volatile("bar.arrive 2, 256;" ::: "memory");
    }
  } else {
    // ===== finaliser warps (6-9): every (track of the tile, detection column) pair -> 1 - max, Mahalanobis gate + blend +
    // clamp and the IoU / DIoU row in fp64, while the column-max warps already reduce the next item
    const int e = threadIdx.x - 192;       // 0..127
    int it = 0;
    for (int w = blockIdx.x; w < total; w += gridDim.x, ++it) {
      const int tile = w / n_dtiles, dt = w - tile * n_dtiles;
      const bool real = tile < n_real;
      const int n_cols = min(bn_box, a.M - dt * bn_box);
      const int4 sg4 = *reinterpret_cast<const int4*>(seg + tile * 4);
      auto SG = [&](int i) { return i == 0 ? sg4.x : (i == 1 ? sg4.y : (i == 2 ? sg4.z : sg4.w)); };
      const int buf = it & 1;
      if (buf == 0) asm volatile("bar.sync 1, 256;" ::: "memory"); else asm volatile("bar.sync 2, 256;" ::: "memory");
      const float* p0 = part + buf * 4 * AT_BN;
      const double* t0 = trk + buf * 4 * 32;
      for (int item = e; item < 4 * n_cols && !(debug & 4); item += 128) {
        const int sq = item / n_cols, col = item - sq * n_cols;   // sq = a quadrant that STARTS a track
        const int t = SG(sq);
        if (t < 0 || (real && sq > 0 && SG(sq - 1) == t)) continue;
        const int m = dt * bn_box + col;
        const size_t o = (size_t)t * a.M + m;
        const double* st = t0 + sq * 32;
        if (a.out_app) {
          double cost = INFTY_COST, raw = INFTY_COST;
          if (real) {
            float mx = p0[sq * AT_BN + col];
#pragma unroll
            for (int p = 1; p < 4; ++p)
              if (p > sq && SG(p) == t) mx = fmaxf(mx, p0[p * AT_BN + col]);
            raw = (double)__fsub_rn(1.0f, __fmul_rn(mx, UNIT_UNSCALE2));
            cost = raw;
          }
          if (a.out_raw) a.out_raw[o] = raw;
          if (!(a.flags & REIDB200_ASSOC_NO_GATE)) {
            double z[4];
            tlwh_to_xyah_f32(a.det_tlwh + (size_t)m * 4, z);
            const double gd = maha(z, st, st + 4);
            if (gd > CHI2INV95_4) cost = INFTY_COST;
            cost = a.mc_lambda * cost + (1 - a.mc_lambda) * gd;
          }
          if ((a.flags & REIDB200_ASSOC_CLAMP_APP) && cost > a.max_dist) cost = a.max_dist + 1e-5;
          a.out_app[o] = cost;
        }
        if (a.out_iou) {
          double c = INFTY_COST;
          if (!(a.trk_tsu[t] > 1)) c = 1.0 - iou_one<float>(st + 20, a.det_tlwh + (size_t)m * 4, a.iou_mode);
          if ((a.flags & REIDB200_ASSOC_CLAMP_IOU) && c > a.max_iou_dist) c = a.max_iou_dist + 1e-5;
          a.out_iou[o] = c;
        }
      }
      if (buf == 0) asm volatile("bar.arrive 3, 256;" ::: "memory"); else asm volatile("bar.arrive 4, 256;" ::: "memory");
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, 512);
  }
}

// ------------------------------------------------------------------------------------ stand-alone pieces
template <typename CT>
__global__ void iou_matrix_kernel(const double* __restrict__ bbox, int T, const CT* __restrict__ cand, int M, int mode,
                                  double* __restrict__ out) {
  const int m = blockIdx.x * blockDim.x + threadIdx.x, t = blockIdx.y;
  if (m >= M) return;
  double b[4] = {bbox[t * 4 + 0], bbox[t * 4 + 1], bbox[t * 4 + 2], bbox[t * 4 + 3]};
  out[(size_t)t * M + m] = iou_one<CT>(b, cand + (size_t)m * 4, mode);
}

__global__ void gating_kernel(const double* __restrict__ x4, const double* __restrict__ SI,
                              const double* __restrict__ meas, int T, int M, double* __restrict__ out) {
  const int m = blockIdx.x * blockDim.x + threadIdx.x, t = blockIdx.y;
  if (m >= M) return;
  double z[4] = {meas[m * 4 + 0], meas[m * 4 + 1], meas[m * 4 + 2], meas[m * 4 + 3]};
  out[(size_t)t * M + m] = maha(z, x4 + (size_t)t * 4, SI + (size_t)t * 16);
}

// warp per row: out = a / ||a||
__global__ void normalize_rows_kernel(const float* __restrict__ x, int n, int64_t ld, int d, float* __restrict__ out) {
  const int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (i >= n) return;
  float s = 0.f;
  for (int k = lane; k < d; k += 32) { const float v = x[(size_t)i * ld + k]; s = fmaf(v, v, s); }
  s = sqrtf(warp_sum(s));
  for (int k = lane; k < d; k += 32) out[(size_t)i * d + k] = __fdiv_rn(x[(size_t)i * ld + k], s);
}

__global__ void one_minus_kernel(float* __restrict__ x, size_t n) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) x[i] = __fsub_rn(1.0f, x[i]);
}

}  // namespace reidb200

using namespace reidb200;

// ints of a frame schedule for T tracks: header (4) + 4 per tile, at most T tiles with rows + ceil(T / 4) without
static size_t sched_ints(int T) { return 4 + 4 * ((size_t)T + (size_t)T / 4 + 2); }

// bytes of the tile-image gallery: slots x k-blocks x 2 planes regions of rb rows x 128 bytes, plus one tile of slack so
// that a fetch rounded up to 16 rows never leaves the allocation
static size_t gallery_packed_bytes(int slots, int rb, int dp) {
  return (size_t)slots * (dp / 64) * 2 * rb * 128 + (size_t)AT_BM * 128;
}

extern "C" int reidb200_gallery_create(int slots, int budget, int d, reidb200_gallery** out) {
  RB_REQUIRE(out && slots > 0 && budget > 0 && d > 0, "gallery_create: bad argument");
  reidb200_gallery* g = new reidb200_gallery();
  g->slots = slots; g->budget = budget; g->d = d; g->dp = padded_dim(d);
  g->rb = (budget + 15) & ~15;
  g->feats = nullptr; g->packed = nullptr; g->count = nullptr; g->head = nullptr; g->scratch = nullptr;
  g->det_packed = nullptr; g->det_cap = 1024;
  g->sched = nullptr; g->sched_cap = slots > 1024 ? slots : 1024;
  g->scratch_rows = slots > 1024 ? slots : 1024;
  const size_t packed_sz = gallery_packed_bytes(slots, g->rb, g->dp);
  cudaError_t e = cudaMalloc(&g->feats, (size_t)slots * budget * d * sizeof(float));
  if (e == cudaSuccess) e = cudaMalloc(&g->packed, packed_sz);
  if (e == cudaSuccess) e = cudaMemset(g->packed, 0, packed_sz);
  if (e == cudaSuccess) e = cudaMalloc(&g->det_packed, (size_t)g->det_cap * 2 * g->dp * sizeof(__half) + (size_t)AT_BN * 128);
  if (e == cudaSuccess) e = cudaMalloc(&g->count, (size_t)slots * sizeof(int32_t));
  if (e == cudaSuccess) e = cudaMalloc(&g->head, (size_t)slots * sizeof(int32_t));
  if (e == cudaSuccess) e = cudaMalloc(&g->scratch, (size_t)2 * g->scratch_rows * sizeof(int32_t));
  if (e == cudaSuccess) e = cudaMalloc(&g->sched, sched_ints(g->sched_cap) * sizeof(int32_t));
  if (e == cudaSuccess) e = cudaMemset(g->count, 0, (size_t)slots * sizeof(int32_t));
  if (e == cudaSuccess) e = cudaMemset(g->head, 0, (size_t)slots * sizeof(int32_t));
  if (e != cudaSuccess) {
    reidb200_gallery_destroy(g);
    return cuda_fail(e, "gallery_create", __FILE__, __LINE__);
  }
  *out = g;
  return REIDB200_OK;
}

// Grow a gallery in place (the handle stays valid): more slots and / or a longer ring per slot, contents preserved bit
// for bit (NearestNeighborDistanceMetric is unbounded in the reference: budget=None keeps every sample and any number of
// targets, boxmot/utils/matching.py:363-389).  Synchronises `stream`; not capturable in a CUDA graph.
extern "C" int reidb200_gallery_resize(reidb200_gallery* g, int new_slots, int new_budget, void* stream) {
  RB_REQUIRE(g && new_slots >= g->slots && new_budget >= g->budget, "gallery_resize: a gallery can only grow");
  if (new_slots == g->slots && new_budget == g->budget) return REIDB200_OK;
  cudaStream_t st = (cudaStream_t)stream;
  float* feats = nullptr; __half* packed = nullptr; int32_t *count = nullptr, *head = nullptr, *scratch = nullptr;
  const int new_rb = (new_budget + 15) & ~15;
  const size_t packed_sz = gallery_packed_bytes(new_slots, new_rb, g->dp);
  const int scratch_rows = new_slots > 1024 ? new_slots : 1024;
  cudaError_t e = cudaMalloc(&feats, (size_t)new_slots * new_budget * g->d * sizeof(float));
  if (e == cudaSuccess) e = cudaMalloc(&packed, packed_sz);
  if (e == cudaSuccess) e = cudaMemsetAsync(packed, 0, packed_sz, st);
  if (e == cudaSuccess) e = cudaMalloc(&count, (size_t)new_slots * sizeof(int32_t));
  if (e == cudaSuccess) e = cudaMalloc(&head, (size_t)new_slots * sizeof(int32_t));
  if (e == cudaSuccess) e = cudaMalloc(&scratch, (size_t)2 * scratch_rows * sizeof(int32_t));
  if (e == cudaSuccess) e = cudaMemsetAsync(count, 0, (size_t)new_slots * sizeof(int32_t), st);
  if (e == cudaSuccess) e = cudaMemsetAsync(head, 0, (size_t)new_slots * sizeof(int32_t), st);
  if (e != cudaSuccess) {
    cudaFree(feats); cudaFree(packed); cudaFree(count); cudaFree(head); cudaFree(scratch);
    return cuda_fail(e, "gallery_resize", __FILE__, __LINE__);
  }
  gal_resize_kernel<<<dim3(g->slots, g->budget < 16 ? g->budget : 16), 128, 0, st>>>(g->feats, g->packed, g->head, g->count, g->budget,
                                                                                  g->rb, new_budget, new_rb, g->d, g->dp, feats, packed, head, count);
  RB_LAUNCH_CHECK("gal_resize_kernel");
  RB_CUDA(cudaStreamSynchronize(st));
  cudaFree(g->feats); cudaFree(g->packed); cudaFree(g->count); cudaFree(g->head); cudaFree(g->scratch);
  g->feats = feats; g->packed = packed; g->count = count; g->head = head; g->scratch = scratch;
  g->slots = new_slots; g->budget = new_budget; g->rb = new_rb; g->scratch_rows = scratch_rows;
  return REIDB200_OK;
}

extern "C" int reidb200_gallery_destroy(reidb200_gallery* g) {
  if (!g) return REIDB200_OK;
  cudaFree(g->feats); cudaFree(g->packed); cudaFree(g->det_packed); cudaFree(g->count); cudaFree(g->head); cudaFree(g->scratch);
  cudaFree(g->sched);
  delete g;
  return REIDB200_OK;
}

extern "C" int reidb200_gallery_partial_fit(reidb200_gallery* g, const float* feats, int64_t ldf,
                                            const int32_t* slot_of_row, int n, void* stream) {
  if (n == 0) return REIDB200_OK;
  RB_REQUIRE(g && feats && slot_of_row && n > 0 && ldf >= g->d, "gallery_partial_fit: bad argument");
  cudaStream_t st = (cudaStream_t)stream;
  for (int b0 = 0; b0 < n; b0 += g->scratch_rows) {  // batches keep the append order
    const int nb = n - b0 < g->scratch_rows ? n - b0 : g->scratch_rows;
    int32_t* dst = g->scratch;
    int32_t* after = g->scratch + g->scratch_rows;
    { ProfScope _ps("pf_assign_kernel", st); pf_assign_kernel<<<ceil_div(nb, 128), 128, 0, st>>>(slot_of_row + b0, nb, g->budget, g->head, dst, after); }
    RB_LAUNCH_CHECK("pf_assign_kernel");
    { ProfScope _ps("pf_write_kernel", st); pf_write_kernel<<<ceil_div(nb, 8), 256, 0, st>>>(feats + (size_t)b0 * ldf, ldf, slot_of_row + b0, dst, nb, g->budget,
                                                    g->d, g->dp, g->rb, g->feats, g->packed); }
    RB_LAUNCH_CHECK("pf_write_kernel");
    { ProfScope _ps("pf_update_kernel", st); pf_update_kernel<<<ceil_div(nb, 128), 128, 0, st>>>(slot_of_row + b0, after, nb, g->budget, g->head, g->count); }
    RB_LAUNCH_CHECK("pf_update_kernel");
  }
  return REIDB200_OK;
}

extern "C" int reidb200_gallery_reset_slots(reidb200_gallery* g, const int32_t* slots, int n, void* stream) {
  if (n == 0) return REIDB200_OK;
  RB_REQUIRE(g && slots && n > 0, "gallery_reset_slots: bad argument");
  { ProfScope _ps("gal_reset_kernel", (cudaStream_t)stream); gal_reset_kernel<<<ceil_div(n, 128), 128, 0, (cudaStream_t)stream>>>(slots, n, g->head, g->count); }
  RB_LAUNCH_CHECK("gal_reset_kernel");
  return REIDB200_OK;
}

extern "C" int reidb200_gallery_counts(reidb200_gallery* g, int32_t* out_counts, void* stream) {
  RB_REQUIRE(g && out_counts, "gallery_counts: bad argument");
  RB_CUDA(cudaMemcpyAsync(out_counts, g->count, (size_t)g->slots * sizeof(int32_t), cudaMemcpyDeviceToDevice,
                          (cudaStream_t)stream));
  return REIDB200_OK;
}

extern "C" int reidb200_gallery_read_slot(reidb200_gallery* g, int slot, float* out_rows, void* stream) {
  RB_REQUIRE(g && out_rows && slot >= 0 && slot < g->slots, "gallery_read_slot: bad argument");
  { ProfScope _ps("gal_read_kernel", (cudaStream_t)stream); gal_read_kernel<<<g->budget, 128, 0, (cudaStream_t)stream>>>(g->feats, g->head, g->count, slot, g->budget, g->d,
                                                              out_rows); }
  RB_LAUNCH_CHECK("gal_read_kernel");
  return REIDB200_OK;
}

extern "C" int reidb200_assoc_frame(reidb200_gallery* g, const float* det_feat, int64_t ldf, const float* det_tlwh, int M,
                                    const int32_t* trk_slot, const double* trk_mean, const double* trk_SI,
                                    const int32_t* trk_tsu, int T, double mc_lambda, double max_dist,
                                    double max_iou_dist, int iou_mode, int flags, double* out_app, double* out_raw,
                                    double* out_iou, void* stream) {
  if (T == 0 || M == 0) return REIDB200_OK;  // the reference returns early for empty inputs (linear_assignment.py:55-56)
  RB_REQUIRE(T > 0 && M > 0 && det_tlwh && trk_mean && trk_tsu, "assoc_frame: bad argument");
  RB_REQUIRE(out_app || out_iou, "assoc_frame: no output requested");
  RB_REQUIRE(!out_app || (g && det_feat && trk_slot && ldf >= g->d), "assoc_frame: appearance cost needs gallery, features and slots");
  RB_REQUIRE(!out_app || (flags & REIDB200_ASSOC_NO_GATE) || trk_SI, "assoc_frame: gate needs trk_SI");
  RB_REQUIRE(iou_mode == REIDB200_IOU_PLAIN || iou_mode == REIDB200_IOU_DIOU, "assoc_frame: bad iou_mode");
  AssocArgs a;
  a.gal = g ? g->feats : nullptr; a.gcount = g ? g->count : nullptr; a.ghead = g ? g->head : nullptr;
  a.budget = g ? g->budget : 0; a.d = g ? g->d : 0;
  a.det_feat = det_feat; a.ldf = ldf; a.det_tlwh = det_tlwh; a.M = M;
  a.trk_slot = trk_slot; a.trk_mean = trk_mean; a.trk_SI = trk_SI; a.trk_tsu = trk_tsu; a.T = T;
  a.mc_lambda = mc_lambda; a.max_dist = max_dist; a.max_iou_dist = max_iou_dist; a.iou_mode = iou_mode; a.flags = flags;
  a.out_app = out_app; a.out_raw = out_app ? out_raw : nullptr; a.out_iou = out_iou;
  cudaStream_t st = (cudaStream_t)stream;
  if (!out_app || (flags & REIDB200_ASSOC_SIMT)) {  // IoU-only frames and the fp32 cross-check of the tests
    { ProfScope _ps("assoc_frame_simt_kernel", st); assoc_frame_simt_kernel<<<T, AS_THREADS, 0, st>>>(a); }
    RB_LAUNCH_CHECK("assoc_frame_simt_kernel");
    return REIDB200_OK;
  }
  if (M > g->det_cap) {  // grows outside the steady state only (synchronises: not capturable in a CUDA graph)
    RB_CUDA(cudaStreamSynchronize(st));
    cudaFree(g->det_packed);
    g->det_packed = nullptr;
    g->det_cap = (M + M / 2 + 15) & ~15;
    RB_CUDA(cudaMalloc(&g->det_packed, (size_t)g->det_cap * 2 * g->dp * sizeof(__half) + (size_t)AT_BN * 128));
  }
  const bool packed_tiles = g->budget <= AT_BM && !getenv("REIDB200_ASSOC_V1");
  if (packed_tiles && T > g->sched_cap) {  // grows outside the steady state only (synchronises)
    RB_CUDA(cudaStreamSynchronize(st));
    cudaFree(g->sched);
    g->sched = nullptr;
    g->sched_cap = T + T / 2;
    RB_CUDA(cudaMalloc(&g->sched, sched_ints(g->sched_cap) * sizeof(int32_t)));
  }
  { ProfScope _ps("det_pack_kernel", st); det_pack_kernel<<<ceil_div(M, 8) + (packed_tiles ? 1 : 0), 256, 0, st>>>(
      det_feat, ldf, M, g->d, g->dp, g->det_cap, g->det_packed, trk_slot, g->count, T, packed_tiles ? g->sched : nullptr); }
  RB_LAUNCH_CHECK("det_pack_kernel");
  if (packed_tiles) {
    // rings of at most 128 rows: tracks packed into 128-row tiles by quadrants, count-aware loads (assoc_frame2_kernel);
    // the detections are split into column ranges of about 112 so that an SM gets several items to pipeline
    const int width = getenv("REIDB200_ASSOC_BN") ? atoi(getenv("REIDB200_ASSOC_BN")) : 112;  // (experiments)
    const int nd = ceil_div(M, width < 16 ? 16 : (width > AT_BN ? AT_BN : width));
    const int bn = (ceil_div(M, nd) + 15) & ~15;
    const int n_dt = ceil_div(M, bn);
    const size_t stage_bytes = 2 * (size_t)AT_PLANE + 2 * (size_t)bn * AT_BK * 2;
    int nst = (int)(A2_STAGE_BUDGET / stage_bytes);
    nst = nst > A2_MAX_STAGES ? A2_MAX_STAGES : nst;
    if (getenv("REIDB200_ASSOC_STAGES")) nst = nst < atoi(getenv("REIDB200_ASSOC_STAGES")) ? nst : atoi(getenv("REIDB200_ASSOC_STAGES"));
    const int debug = getenv("REIDB200_ASSOC_DEBUG") ? atoi(getenv("REIDB200_ASSOC_DEBUG")) : 0;
    static DeviceOnce attr2;
    if (attr2.pending()) {
      RB_CUDA(cudaFuncSetAttribute(assoc_frame2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)A2_SMEM));
    }
    const long long most = ((long long)T + T / 4 + 2) * n_dt;  // upper bound of the work items (the count lives on the device)
    const int grid = (int)(most < num_sms() ? most : num_sms());
    { ProfScope _ps("assoc_frame_kernel", st); assoc_frame2_kernel<<<grid, A2_THREADS, A2_SMEM, st>>>(g->packed, g->rb, g->det_packed, g->det_cap, a, g->dp, n_dt, bn, nst, g->sched, debug); }
    RB_LAUNCH_CHECK("assoc_frame2_kernel");
    return REIDB200_OK;
  }
  const int bn_box = M >= AT_BN ? AT_BN : ((M + 15) & ~15);
  const int n_dtiles = ceil_div(M, AT_BN);
  static DeviceOnce attr_done;
  if (attr_done.pending()) {
    RB_CUDA(cudaFuncSetAttribute(assoc_frame_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)AT_SMEM));
  }
  const long long total = (long long)T * n_dtiles;
  const int grid = (int)(total < num_sms() ? total : num_sms());
  { ProfScope _ps("assoc_frame_kernel", st); assoc_frame_kernel<<<grid, AT_THREADS, AT_SMEM, st>>>(g->packed, g->rb, g->det_packed, g->det_cap, a, g->dp, n_dtiles, bn_box); }
  RB_LAUNCH_CHECK("assoc_frame_kernel");
  return REIDB200_OK;
}

extern "C" int reidb200_iou_matrix(const double* bbox_tlwh, int T, const void* cand_tlwh, int cand_is_f64, int M,
                                   int iou_mode, double* out, void* stream) {
  if (T == 0 || M == 0) return REIDB200_OK;
  RB_REQUIRE(bbox_tlwh && cand_tlwh && out && T > 0 && M > 0, "iou_matrix: bad argument");
  dim3 grid(ceil_div(M, 128), T);
  if (cand_is_f64)
    iou_matrix_kernel<double><<<grid, 128, 0, (cudaStream_t)stream>>>(bbox_tlwh, T, (const double*)cand_tlwh, M, iou_mode, out);
  else
    iou_matrix_kernel<float><<<grid, 128, 0, (cudaStream_t)stream>>>(bbox_tlwh, T, (const float*)cand_tlwh, M, iou_mode, out);
  RB_LAUNCH_CHECK("iou_matrix_kernel");
  return REIDB200_OK;
}

extern "C" int reidb200_gating_distance(const double* x4, const double* SI, const double* meas, int T, int M, double* out,
                                        void* stream) {
  if (T == 0 || M == 0) return REIDB200_OK;
  RB_REQUIRE(x4 && SI && meas && out && T > 0 && M > 0, "gating_distance: bad argument");
  dim3 grid(ceil_div(M, 128), T);
  { ProfScope _ps("gating_kernel", (cudaStream_t)stream); gating_kernel<<<grid, 128, 0, (cudaStream_t)stream>>>(x4, SI, meas, T, M, out); }
  RB_LAUNCH_CHECK("gating_kernel");
  return REIDB200_OK;
}

extern "C" size_t reidb200_cosine_distance_workspace_bytes(int na, int nb, int d) {
  if (na <= 0 || nb <= 0 || d <= 0) return 0;
  CarveSize c;
  c.take<float>((size_t)na * d);
  c.take<float>((size_t)nb * d);
  c.take<char>(packed_bytes(na, d));
  c.take<char>(packed_bytes(nb, d));
  c.take<float>(na);
  c.take<float>(nb);
  return align_up(c.off, 256);
}

extern "C" int reidb200_cosine_distance(const float* a, int na, int64_t lda, const float* b, int nb, int64_t ldb, int d,
                                        int normalized, float* out, void* workspace, size_t workspace_bytes,
                                        void* stream) {
  if (na == 0 || nb == 0) return REIDB200_OK;
  RB_REQUIRE(a && b && out && na > 0 && nb > 0 && d > 0, "cosine_distance: bad argument");
  RB_REQUIRE(workspace && workspace_bytes >= reidb200_cosine_distance_workspace_bytes(na, nb, d),
             "cosine_distance: workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  Carver c(workspace, workspace_bytes);
  float* an = c.take<float>((size_t)na * d);
  float* bn = c.take<float>((size_t)nb * d);
  char* pa = c.take<char>(packed_bytes(na, d));
  char* pb = c.take<char>(packed_bytes(nb, d));
  float* isa = c.take<float>(na);
  float* isb = c.take<float>(nb);
  const float* ax = a; const float* bx = b;
  int64_t la = lda, lb = ldb;
  if (!normalized) {
    { ProfScope _ps("normalize_rows_kernel", st); normalize_rows_kernel<<<ceil_div(na, 8), 256, 0, st>>>(a, na, lda, d, an); }
    RB_LAUNCH_CHECK("normalize_rows_kernel");
    { ProfScope _ps("normalize_rows_kernel", st); normalize_rows_kernel<<<ceil_div(nb, 8), 256, 0, st>>>(b, nb, ldb, d, bn); }
    RB_LAUNCH_CHECK("normalize_rows_kernel");
    ax = an; bx = bn; la = d; lb = d;
  }
  int rc = pack_rows_launch(ax, na, la, d, pa, nullptr, isa, st);
  if (rc) return rc;
  rc = pack_rows_launch(bx, nb, lb, d, pb, nullptr, isb, st);
  if (rc) return rc;
  rc = gemm_split_launch(pa, nullptr, isa, na, pb, nullptr, isb, nb, d, REIDB200_DIST_DOT, out, nb, st);
  if (rc) return rc;
  const size_t n = (size_t)na * nb;
  { ProfScope _ps("one_minus_kernel", st); one_minus_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(out, n); }
  RB_LAUNCH_CHECK("one_minus_kernel");
  return REIDB200_OK;
}

// ------------------------------------------------------------------------------------ one frame from host arrays
// The per-frame host <-> device traffic of a tracker loop in two calls (crops before the backbone, costs
// after it), each ONE pinned host -> device copy of the small per-frame tables, the kernels, and -- for the costs --
// ONE device -> host copy of both float64 matrices into pinned memory the caller reads in place.  No per-frame
// allocation, no Python-side marshalling: plain host pointers and sizes in, pointers into the pinned result out.
struct reidb200_frame {
  reidb200_gallery* g;
  int max_dets, max_tracks;
  size_t in_cap;            // bytes of one input staging buffer
  uint8_t* h_in[2];         // pinned
  uint8_t* d_in[2];
  cudaEvent_t in_done[2];   // every consumer of staging pair i (copy and kernels) has finished
  cudaEvent_t in_ready[2];  // the side-stream copy of staging pair i has landed
  cudaStream_t copy;        // side stream of the cost tables' H2D: it overlaps whatever the caller's stream still runs
  double* h_out[2];         // pinned + mapped [2][max_tracks][max_dets]
  double* h_out_dev[2];     // the same buffers as the device sees them (zero-copy target of the finaliser warps)
  double* d_out;
  int turn_in, turn_out;
  int zero_copy;            // 1: the kernel writes the cost blocks straight into h_out (no D2H copy behind the kernel)
};

static size_t frame_in_bytes(int max_dets, int max_tracks) {
  const size_t crops = (size_t)max_dets * (8 + 12);                                    // crop_off i64 | crop_geom i32[3]
  const size_t costs = (size_t)max_dets * 16 + (size_t)max_tracks * (4 + 4 + 64 + 128);  // tlwh | slot | tsu | mean | SI
  return align_up((crops > costs ? crops : costs) + 256, 256);
}

extern "C" int reidb200_frame_create(reidb200_gallery* g, int max_dets, int max_tracks, reidb200_frame** out) {
  RB_REQUIRE(g && out && max_dets > 0 && max_tracks > 0, "frame_create: bad argument");
  reidb200_frame* f = new reidb200_frame();
  memset(f, 0, sizeof(*f));
  f->g = g; f->max_dets = max_dets; f->max_tracks = max_tracks;
  f->in_cap = frame_in_bytes(max_dets, max_tracks);
  const size_t out_bytes = (size_t)2 * max_tracks * max_dets * sizeof(double);
  cudaError_t e = cudaSuccess;
  for (int i = 0; i < 2 && e == cudaSuccess; ++i) {
    e = cudaMallocHost(&f->h_in[i], f->in_cap);
    if (e == cudaSuccess) e = cudaMalloc(&f->d_in[i], f->in_cap);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&f->in_done[i], cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&f->in_ready[i], cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaHostAlloc(&f->h_out[i], out_bytes, cudaHostAllocMapped);
    if (e == cudaSuccess) e = cudaHostGetDevicePointer(reinterpret_cast<void**>(&f->h_out_dev[i]), f->h_out[i], 0);
  }
  if (e == cudaSuccess) e = cudaMalloc(&f->d_out, out_bytes);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&f->copy, cudaStreamNonBlocking);
  { const char* m = getenv("REIDB200_FRAME_MODE"); f->zero_copy = m ? atoi(m) : 3; }  // measured: profiles/r02_frame_modes.log
  if (e != cudaSuccess) {
    reidb200_frame_destroy(f);
    return cuda_fail(e, "frame_create", __FILE__, __LINE__);
  }
  *out = f;
  return REIDB200_OK;
}

extern "C" int reidb200_frame_destroy(reidb200_frame* f) {
  if (!f) return REIDB200_OK;
  for (int i = 0; i < 2; ++i) {
    if (f->h_in[i]) cudaFreeHost(f->h_in[i]);
    if (f->d_in[i]) cudaFree(f->d_in[i]);
    if (f->in_done[i]) cudaEventDestroy(f->in_done[i]);
    if (f->in_ready[i]) cudaEventDestroy(f->in_ready[i]);
    if (f->h_out[i]) cudaFreeHost(f->h_out[i]);
  }
  if (f->d_out) cudaFree(f->d_out);
  if (f->copy) cudaStreamDestroy(f->copy);
  delete f;
  return REIDB200_OK;
}

// next input staging pair (whatever consumed it two calls ago must have finished: normally long ago)
static int frame_next_in(reidb200_frame* f) {
  const int i = f->turn_in;
  f->turn_in ^= 1;
  cudaEventSynchronize(f->in_done[i]);  // an event that was never recorded is complete
  return i;
}

extern "C" int reidb200_frame_crops(reidb200_frame* f, const uint8_t* frame, int H, int W, const int32_t* xyxy_host, int n,
                                    int out_h, int out_w, const float mean_host[3], const float std_host[3], float* out,
                                    void* stream) {
  if (n == 0) return REIDB200_OK;
  RB_REQUIRE(f && frame && xyxy_host && out && n > 0 && H > 0 && W > 0, "frame_crops: bad argument");
  RB_REQUIRE(n <= f->max_dets, "frame_crops: %d boxes exceed the context's %d", n, f->max_dets);
  cudaStream_t st = (cudaStream_t)stream;
  const int bi = frame_next_in(f);
  int64_t* off = reinterpret_cast<int64_t*>(f->h_in[bi]);
  int32_t* geom = reinterpret_cast<int32_t*>(f->h_in[bi] + (size_t)n * 8);
  for (int i = 0; i < n; ++i) {
    // img[y1:y2, x1:x2] (strong_sort.py:153-155): numpy clips the slice end at the frame border
    const int x1 = xyxy_host[4 * i], y1 = xyxy_host[4 * i + 1];
    const int x2 = xyxy_host[4 * i + 2] < W ? xyxy_host[4 * i + 2] : W, y2 = xyxy_host[4 * i + 3] < H ? xyxy_host[4 * i + 3] : H;
    RB_REQUIRE(x1 >= 0 && y1 >= 0, "frame_crops: box %d has a negative origin", i);
    RB_REQUIRE(x2 > x1 && y2 > y1, "frame_crops: empty crop: box %d has zero area after clamping", i);
    off[i] = ((int64_t)y1 * W + x1) * 3;
    geom[3 * i] = y2 - y1; geom[3 * i + 1] = x2 - x1; geom[3 * i + 2] = W * 3;
  }
  const size_t bytes = (size_t)n * 20;
  RB_CUDA(cudaMemcpyAsync(f->d_in[bi], f->h_in[bi], bytes, cudaMemcpyHostToDevice, st));
  const int rc = reidb200_crop_resize_norm(frame, reinterpret_cast<const int64_t*>(f->d_in[bi]),
                                           reinterpret_cast<const int32_t*>(f->d_in[bi] + (size_t)n * 8), n, out_h, out_w,
                                           mean_host, std_host, REIDB200_RESIZE_CV_LINEAR, out, stream);
  RB_CUDA(cudaEventRecord(f->in_done[bi], st));
  return rc;
}

extern "C" int reidb200_frame_costs(reidb200_frame* f, const float* det_feat, int64_t ldf, const float* det_tlwh_host, int M,
                                    const int32_t* trk_slot_host, const double* trk_mean_host, const double* trk_SI_host,
                                    const int32_t* trk_tsu_host, int T, double mc_lambda, double max_dist, double max_iou_dist,
                                    int iou_mode, int flags, const double** out_app_host, const double** out_iou_host,
                                    void* stream) {
  RB_REQUIRE(f && out_app_host && out_iou_host, "frame_costs: bad argument");
  *out_app_host = *out_iou_host = nullptr;
  if (T == 0 || M == 0) return REIDB200_OK;
  RB_REQUIRE(det_feat && det_tlwh_host && trk_slot_host && trk_mean_host && trk_SI_host && trk_tsu_host && T > 0 && M > 0,
             "frame_costs: bad argument");
  RB_REQUIRE(M <= f->max_dets && T <= f->max_tracks, "frame_costs: %d detections / %d tracks exceed the context's %d / %d", M, T,
             f->max_dets, f->max_tracks);
  cudaStream_t st = (cudaStream_t)stream;
  const int bi = frame_next_in(f);
  // layout (every piece 8-byte aligned): mean f64[T,8] | SI f64[T,16] | tlwh f32[M,4] | slot i32[T] | tsu i32[T]
  uint8_t* h = f->h_in[bi];
  const size_t o_mean = 0, o_si = o_mean + (size_t)T * 64, o_tlwh = o_si + (size_t)T * 128, o_slot = o_tlwh + (size_t)M * 16,
               o_tsu = o_slot + align_up((size_t)T * 4, 8), total = o_tsu + (size_t)T * 4;
  memcpy(h + o_mean, trk_mean_host, (size_t)T * 64);
  memcpy(h + o_si, trk_SI_host, (size_t)T * 128);
  memcpy(h + o_tlwh, det_tlwh_host, (size_t)M * 16);
  memcpy(h + o_slot, trk_slot_host, (size_t)T * 4);
  memcpy(h + o_tsu, trk_tsu_host, (size_t)T * 4);
  uint8_t* d = f->d_in[bi];
  // mode bit 1: the tables go up on the side stream (they depend on nothing the caller's stream is doing: the crop kernel
  // or the embedding network may still be running there), the kernels wait for them by event
  // (worth its event pair only when the stream is still busy with a large crop launch: 93 vs 105 us at 200 boxes, 75 vs 62 us
  // at 30; REIDB200_FRAME_MODE=7 forces it for every frame)
  if ((f->zero_copy & 2) && (M >= 96 || (f->zero_copy & 4))) {
    RB_CUDA(cudaMemcpyAsync(d, h, total, cudaMemcpyHostToDevice, f->copy));
    RB_CUDA(cudaEventRecord(f->in_ready[bi], f->copy));
    RB_CUDA(cudaStreamWaitEvent(st, f->in_ready[bi], 0));
  } else {
    RB_CUDA(cudaMemcpyAsync(d, h, total, cudaMemcpyHostToDevice, st));
  }
  const int bo = f->turn_out;
  f->turn_out ^= 1;
  // mode bit 0 (zero copy): the finaliser warps store the two [T, M] blocks straight into mapped pinned memory (write-only,
  // row pieces of consecutive columns), so the 2*T*M*8 bytes cross PCIe WHILE the kernel runs instead of in a copy behind it
  double* app = (f->zero_copy & 1) ? f->h_out_dev[bo] : f->d_out;
  double* iou = app + (size_t)T * M;
  int rc = reidb200_assoc_frame(f->g, det_feat, ldf, reinterpret_cast<const float*>(d + o_tlwh), M,
                                reinterpret_cast<const int32_t*>(d + o_slot), reinterpret_cast<const double*>(d + o_mean),
                                reinterpret_cast<const double*>(d + o_si), reinterpret_cast<const int32_t*>(d + o_tsu), T, mc_lambda,
                                max_dist, max_iou_dist, iou_mode, flags, app, nullptr, iou, stream);
  RB_CUDA(cudaEventRecord(f->in_done[bi], st));
  if (rc) return rc;
  if (!(f->zero_copy & 1))
    RB_CUDA(cudaMemcpyAsync(f->h_out[bo], f->d_out, (size_t)2 * T * M * sizeof(double), cudaMemcpyDeviceToHost, st));
  RB_CUDA(cudaStreamSynchronize(st));
  *out_app_host = f->h_out[bo];
  *out_iou_host = f->h_out[bo] + (size_t)T * M;
  return REIDB200_OK;
}

