// N2: the pairwise box costs of the other boxmot trackers (ByteTrack / BoT-SORT / OC-SORT / DeepOCSORT).
//   iou_batch / giou_batch / diou_batch / ciou_batch   yolov8_tracking/boxmot/utils/iou.py:4-164   (xyxy, input dtype)
//   bbox_ious ("+1" pixel IoU), iou_distance            yolov8_tracking/boxmot/utils/matching.py:71-114, 216-248
//   fuse_score, fuse_iou                                yolov8_tracking/boxmot/utils/matching.py:189-213
// One thread per (row box, column box) pair; every operation is the reference's, in the reference's order and dtype
// (no fused multiply-add: the library is built with -fmad=false), so the fp32 and fp64 results are bit-identical to
// numpy's except for arctan in the CIoU aspect term (library function, last-bit differences).
#include "common.cuh"

namespace reidb200 {

template <typename T>
struct Box4 {
  T x1, y1, x2, y2;
};

template <typename T>
__device__ __forceinline__ Box4<T> load_box(const T* p, int i) {
  Box4<T> b;
  b.x1 = p[4 * i + 0]; b.y1 = p[4 * i + 1]; b.x2 = p[4 * i + 2]; b.y2 = p[4 * i + 3];
  return b;
}

template <typename T> __device__ __forceinline__ T tmax(T a, T b) { return a > b ? a : b; }
template <typename T> __device__ __forceinline__ T tmin(T a, T b) { return a < b ? a : b; }
__device__ __forceinline__ float tatan(float x) { return atanf(x); }
__device__ __forceinline__ double tatan(double x) { return atan(x); }

// np.maximum / np.minimum propagate NaN; boxes are finite here, plain comparisons are equivalent.
template <typename T>
__device__ __forceinline__ T pair_value(const Box4<T>& a, const Box4<T>& b, int mode) {
  const T iw = tmax((T)0, tmin(a.x2, b.x2) - tmax(a.x1, b.x1));
  const T ih = tmax((T)0, tmin(a.y2, b.y2) - tmax(a.y1, b.y1));
  const T inter = iw * ih;
  const T iou = inter / (((a.x2 - a.x1) * (a.y2 - a.y1) + (b.x2 - b.x1) * (b.y2 - b.y1)) - inter);
  if (mode == REIDB200_BOX_IOU) return iou;
  const T ex1 = tmin(a.x1, b.x1), ey1 = tmin(a.y1, b.y1), ex2 = tmax(a.x2, b.x2), ey2 = tmax(a.y2, b.y2);
  if (mode == REIDB200_BOX_GIOU) {
    const T enclose = (ex2 - ex1) * (ey2 - ey1);
    return ((iou - (enclose - inter) / enclose) + (T)1) / (T)2;
  }
  const T cx1 = (a.x1 + a.x2) / (T)2, cy1 = (a.y1 + a.y2) / (T)2;
  const T cx2 = (b.x1 + b.x2) / (T)2, cy2 = (b.y1 + b.y2) / (T)2;
  const T dx = cx1 - cx2, dy = cy1 - cy2;
  const T inner = dx * dx + dy * dy;
  const T ox = ex2 - ex1, oy = ey2 - ey1;
  const T outer = ox * ox + oy * oy;
  if (mode == REIDB200_BOX_DIOU) return ((iou - inner / outer) + (T)1) / (T)2;
  // CIoU: aspect-ratio term with the reference's one-pixel shift of both heights
  const T w1 = a.x2 - a.x1, h1 = (a.y2 - a.y1) + (T)1;
  const T w2 = b.x2 - b.x1, h2 = (b.y2 - b.y1) + (T)1;
  const T at = tatan(w2 / h2) - tatan(w1 / h1);
  const T v = (T)(4.0 / (3.141592653589793 * 3.141592653589793)) * (at * at);
  const T alpha = v / (((T)1 - iou) + v);
  return (((iou - inner / outer) - alpha * v) + (T)1) / (T)2;
}

// "+1" IoU on float32 boxes: float32 extents and overlap, union promoted to double, quotient stored as float32
__device__ __forceinline__ float plus_one_iou(const Box4<float>& a, const Box4<float>& q) {
  const float iw = (tmin(a.x2, q.x2) - tmax(a.x1, q.x1)) + 1.0f;
  if (!(iw > 0.f)) return 0.f;
  const float ih = (tmin(a.y2, q.y2) - tmax(a.y1, q.y1)) + 1.0f;
  if (!(ih > 0.f)) return 0.f;
  const float qa = ((q.x2 - q.x1) + 1.0f) * ((q.y2 - q.y1) + 1.0f);
  const float ba = ((a.x2 - a.x1) + 1.0f) * ((a.y2 - a.y1) + 1.0f);
  const float inter = iw * ih;
  const double ua = (double)((ba + qa) - inter);
  return (float)((double)inter / ua);
}

template <typename T>
__global__ void __launch_bounds__(256)
box_pairs_kernel(const T* __restrict__ a, int n, const T* __restrict__ b, int k, int mode, T* __restrict__ out) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)n * k) return;
  const int r = (int)(idx / k), c = (int)(idx - (long long)r * k);
  out[idx] = pair_value<T>(load_box(a, r), load_box(b, c), mode);
}

__global__ void __launch_bounds__(256)
box_plus_one_kernel(const float* __restrict__ a, int n, const float* __restrict__ b, int k, int as_distance,
                    float* __restrict__ out) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)n * k) return;
  const int r = (int)(idx / k), c = (int)(idx - (long long)r * k);
  const float v = plus_one_iou(load_box(a, r), load_box(b, c));
  out[idx] = as_distance ? 1.0f - v : v;
}

// fuse_score: 1 - (1 - cost) * score[c]      (cost f32 or f64 -> out f64)
// fuse_iou  : 1 - (1 - cost) * (1 + (1 - iou_distance)) / 2   (cost f32 -> f32, cost f64 -> f64; IoU part in f32)
template <typename T>
__global__ void __launch_bounds__(256)
fuse_kernel(const T* __restrict__ cost, int n, int k, const double* __restrict__ score, const float* __restrict__ a,
            const float* __restrict__ b, void* __restrict__ out) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)n * k) return;
  const int r = (int)(idx / k), c = (int)(idx - (long long)r * k);
  if (score) {
    const T sim = (T)1 - cost[idx];
    reinterpret_cast<double*>(out)[idx] = 1.0 - (double)sim * score[c];
  } else {
    const float iou_dist = 1.0f - plus_one_iou(load_box(a, r), load_box(b, c));
    const float iou_sim = 1.0f - iou_dist;
    const T reid_sim = (T)1 - cost[idx];
    const float one_plus = 1.0f + iou_sim;  // float32 in the reference whatever the cost dtype
    reinterpret_cast<T*>(out)[idx] = (T)1 - (reid_sim * (T)one_plus) / (T)2;
  }
}

// OC-SORT / DeepOCSORT association cost (boxmot/utils/association.py:6-16 speed_direction_batch, :107-140 associate):
//   iou[d,t]   = iou_batch(detections, trackers)
//   angle[d,t] = valid[t] * (pi/2 - |acos(clip(vx[t]*X + vy[t]*Y, -1, 1))|) / pi * vdc_weight * score[d]
// with (Y, X) the unit direction from the track's previous observation centre to the detection centre.  float64.
__global__ void __launch_bounds__(256)
ocsort_cost_kernel(const double* __restrict__ det, int D, const double* __restrict__ trk, int T,
                   const double* __restrict__ vel, const double* __restrict__ prev, double vdc_weight,
                   double* __restrict__ out_iou, double* __restrict__ out_angle) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)D * T) return;
  const int d = (int)(idx / T), t = (int)(idx - (long long)d * T);
  Box4<double> a, b;
  a.x1 = det[5 * d + 0]; a.y1 = det[5 * d + 1]; a.x2 = det[5 * d + 2]; a.y2 = det[5 * d + 3];
  b.x1 = trk[5 * t + 0]; b.y1 = trk[5 * t + 1]; b.x2 = trk[5 * t + 2]; b.y2 = trk[5 * t + 3];
  out_iou[idx] = pair_value<double>(a, b, REIDB200_BOX_IOU);
  const double cx1 = (a.x1 + a.x2) / 2.0, cy1 = (a.y1 + a.y2) / 2.0;
  const double cx2 = (prev[5 * t + 0] + prev[5 * t + 2]) / 2.0, cy2 = (prev[5 * t + 1] + prev[5 * t + 3]) / 2.0;
  double dx = cx1 - cx2, dy = cy1 - cy2;
  const double norm = sqrt(dx * dx + dy * dy) + 1e-6;
  dx = dx / norm;
  dy = dy / norm;
  // velocities[:, 0] pairs with Y (= dy), velocities[:, 1] with X (= dx)
  double c = vel[2 * t + 1] * dx + vel[2 * t + 0] * dy;
  c = c < -1.0 ? -1.0 : (c > 1.0 ? 1.0 : c);
  const double diff = (1.5707963267948966 - fabs(acos(c))) / 3.141592653589793;
  const double valid = prev[5 * t + 4] < 0.0 ? 0.0 : 1.0;
  out_angle[idx] = ((valid * diff) * vdc_weight) * det[5 * d + 4];
}

}  // namespace reidb200

using namespace reidb200;

extern "C" int reidb200_box_pairs(const void* boxes_a, int n, const void* boxes_b, int k, int is_f64, int mode, void* out,
                                  void* stream) {
  if (n == 0 || k == 0) return REIDB200_OK;
  RB_REQUIRE(boxes_a && boxes_b && out && n > 0 && k > 0, "box_pairs: bad argument");
  RB_REQUIRE(mode >= REIDB200_BOX_IOU && mode <= REIDB200_BOX_IOU_DISTANCE, "box_pairs: unknown mode %d", mode);
  cudaStream_t st = (cudaStream_t)stream;
  const long long total = (long long)n * k;
  const int grid = (int)((total + 255) / 256);
  if (mode >= REIDB200_BOX_PLUS_ONE) {
    RB_REQUIRE(!is_f64, "box_pairs: the +1 IoU of matching.py is defined on float32 boxes");
    { ProfScope _ps("box_plus_one_kernel", st); box_plus_one_kernel<<<grid, 256, 0, st>>>(
        static_cast<const float*>(boxes_a), n, static_cast<const float*>(boxes_b), k, mode == REIDB200_BOX_IOU_DISTANCE,
        static_cast<float*>(out)); }
    RB_LAUNCH_CHECK("box_plus_one_kernel");
    return REIDB200_OK;
  }
  if (is_f64) {
    { ProfScope _ps("box_pairs_kernel", st); box_pairs_kernel<double><<<grid, 256, 0, st>>>(
        static_cast<const double*>(boxes_a), n, static_cast<const double*>(boxes_b), k, mode, static_cast<double*>(out)); }
  } else {
    { ProfScope _ps("box_pairs_kernel", st); box_pairs_kernel<float><<<grid, 256, 0, st>>>(
        static_cast<const float*>(boxes_a), n, static_cast<const float*>(boxes_b), k, mode, static_cast<float*>(out)); }
  }
  RB_LAUNCH_CHECK("box_pairs_kernel");
  return REIDB200_OK;
}

extern "C" int reidb200_fuse_cost(const void* cost, int cost_is_f64, int n, int k, const double* det_scores,
                                  const float* tlbr_a, const float* tlbr_b, void* out, void* stream) {
  if (n == 0 || k == 0) return REIDB200_OK;
  RB_REQUIRE(cost && out && n > 0 && k > 0, "fuse_cost: bad argument");
  RB_REQUIRE((det_scores != nullptr) != (tlbr_a != nullptr && tlbr_b != nullptr),
             "fuse_cost: pass either the detection scores (fuse_score) or both box sets (fuse_iou)");
  cudaStream_t st = (cudaStream_t)stream;
  const long long total = (long long)n * k;
  const int grid = (int)((total + 255) / 256);
  if (cost_is_f64) {
    { ProfScope _ps("fuse_kernel", st); fuse_kernel<double><<<grid, 256, 0, st>>>(static_cast<const double*>(cost), n, k, det_scores, tlbr_a, tlbr_b, out); }
  } else {
    { ProfScope _ps("fuse_kernel", st); fuse_kernel<float><<<grid, 256, 0, st>>>(static_cast<const float*>(cost), n, k, det_scores, tlbr_a, tlbr_b, out); }
  }
  RB_LAUNCH_CHECK("fuse_kernel");
  return REIDB200_OK;
}

extern "C" int reidb200_ocsort_cost(const double* detections, int D, const double* trackers, int T, const double* velocities,
                                    const double* previous_obs, double vdc_weight, double* out_iou, double* out_angle,
                                    void* stream) {
  if (D == 0 || T == 0) return REIDB200_OK;
  RB_REQUIRE(detections && trackers && velocities && previous_obs && out_iou && out_angle && D > 0 && T > 0,
             "ocsort_cost: bad argument");
  cudaStream_t st = (cudaStream_t)stream;
  const long long total = (long long)D * T;
  { ProfScope _ps("ocsort_cost_kernel", st); ocsort_cost_kernel<<<(int)((total + 255) / 256), 256, 0, st>>>(
      detections, D, trackers, T, velocities, previous_obs, vdc_weight, out_iou, out_angle); }
  RB_LAUNCH_CHECK("ocsort_cost_kernel");
  return REIDB200_OK;
}
