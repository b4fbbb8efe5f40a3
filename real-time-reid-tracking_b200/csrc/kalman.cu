// N3: batched constant-velocity Kalman filter of StrongSORT, device resident.
//   StrongSortKalmanFilterAdapter.initiate / predict / update   boxmot/motion/adapters/strongsort_kf_adapter.py:25-135
//   KalmanFilter.predict / update (x, P, S, SI, K, Joseph form)   boxmot/motion/kalman_filter.py:340-379, 486-575
// State per track slot: x f64[8] (xyah + velocities), P f64[8,8], SI f64[4,4] (inverse innovation covariance of the
// LAST update: the gate of the association kernel reads it stale, exactly like the reference).
// One CTA of 64 threads per listed track, thread = (row i, column j) of the 8 x 8 covariance.  F = I + dt * shift and
// H = [I4 | 0] are applied structurally: every product with a 0 or 1 entry of F / H is exact in the reference's BLAS
// too, so predict is bit-identical for dt = 1 and update differs only by the summation order of the 4- and 8-term dot
// products and by the 4 x 4 inverse (LU with partial pivoting here, LAPACK there).
#include "common.cuh"

namespace reidb200 {

__global__ void __launch_bounds__(64)
kf_initiate_kernel(double* __restrict__ x, double* __restrict__ P, double* __restrict__ SI, const int32_t* __restrict__ slots,
                   const double* __restrict__ meas, int n, double w_pos, double w_vel) {
  const int t = blockIdx.x;
  if (t >= n) return;
  const int s = slots[t], tid = threadIdx.x, i = tid >> 3, j = tid & 7;
  const double* m = meas + 4 * t;
  double std;
  switch (i) {
    case 0: std = 2 * w_pos * m[0]; break;
    case 1: std = 2 * w_pos * m[1]; break;
    case 2: std = 1 * m[2]; break;
    case 3: std = 2 * w_pos * m[3]; break;
    case 4: std = 10 * w_vel * m[0]; break;
    case 5: std = 10 * w_vel * m[1]; break;
    case 6: std = 0.1 * m[2]; break;
    default: std = 10 * w_vel * m[3]; break;
  }
  P[(size_t)s * 64 + tid] = i == j ? std * std : 0.0;
  if (tid < 8) x[(size_t)s * 8 + tid] = tid < 4 ? m[tid] : 0.0;
  if (tid < 16) SI[(size_t)s * 16 + tid] = 0.0;  // KalmanFilter.__init__: SI = zeros until the first update
}

__global__ void __launch_bounds__(64)
kf_predict_kernel(double* __restrict__ x, double* __restrict__ P, const int32_t* __restrict__ slots, int n, double dt,
                  double w_pos, double w_vel) {
  __shared__ double sP[64], sFP[64], sx[8];
  const int t = blockIdx.x;
  if (t >= n) return;
  const int s = slots ? slots[t] : t, tid = threadIdx.x, i = tid >> 3, j = tid & 7;
  sP[tid] = P[(size_t)s * 64 + tid];
  if (tid < 8) sx[tid] = x[(size_t)s * 8 + tid];
  __syncthreads();
  // FP = F P: rows 0-3 pick up dt * rows 4-7
  sFP[tid] = i < 4 ? sP[tid] + dt * sP[tid + 32] : sP[tid];
  __syncthreads();
  // (FP) F^T: columns 0-3 pick up dt * columns 4-7; + Q (diagonal, from the mean BEFORE the prediction)
  double v = j < 4 ? sFP[tid] + dt * sFP[tid + 4] : sFP[tid];
  if (i == j) {
    double std;
    switch (i) {
      case 0: std = w_pos * sx[0]; break;
      case 1: std = w_pos * sx[1]; break;
      case 2: std = 1 * sx[2]; break;
      case 3: std = w_pos * sx[3]; break;
      case 4: std = w_vel * sx[0]; break;
      case 5: std = w_vel * sx[1]; break;
      case 6: std = 0.1 * sx[2]; break;
      default: std = w_vel * sx[3]; break;
    }
    v = v + std * std;
  }
  P[(size_t)s * 64 + tid] = v;
  if (tid < 8) x[(size_t)s * 8 + tid] = tid < 4 ? sx[tid] + dt * sx[tid + 4] : sx[tid];
}

// 4 x 4 inverse by LU with partial pivoting, solving for the identity (one thread)
__device__ void inv4(const double* S, double* out) {
  double a[4][4];
  int piv[4];
  for (int r = 0; r < 4; ++r)
    for (int c = 0; c < 4; ++c) a[r][c] = S[4 * r + c];
  for (int k = 0; k < 4; ++k) {
    int p = k;
    double best = fabs(a[k][k]);
    for (int r = k + 1; r < 4; ++r)
      if (fabs(a[r][k]) > best) { best = fabs(a[r][k]); p = r; }
    piv[k] = p;
    if (p != k)
      for (int c = 0; c < 4; ++c) { const double tmp = a[k][c]; a[k][c] = a[p][c]; a[p][c] = tmp; }
    for (int r = k + 1; r < 4; ++r) {
      a[r][k] = a[r][k] / a[k][k];
      for (int c = k + 1; c < 4; ++c) a[r][c] = a[r][c] - a[r][k] * a[k][c];
    }
  }
  for (int col = 0; col < 4; ++col) {
    double b[4] = {0.0, 0.0, 0.0, 0.0};
    b[col] = 1.0;
    for (int k = 0; k < 4; ++k)
      if (piv[k] != k) { const double tmp = b[k]; b[k] = b[piv[k]]; b[piv[k]] = tmp; }
    for (int r = 1; r < 4; ++r)
      for (int c = 0; c < r; ++c) b[r] = b[r] - a[r][c] * b[c];
    for (int r = 3; r >= 0; --r) {
      for (int c = r + 1; c < 4; ++c) b[r] = b[r] - a[r][c] * b[c];
      b[r] = b[r] / a[r][r];
    }
    for (int r = 0; r < 4; ++r) out[4 * r + col] = b[r];
  }
}

__global__ void __launch_bounds__(64)
kf_update_kernel(double* __restrict__ x, double* __restrict__ P, double* __restrict__ SI, const int32_t* __restrict__ slots,
                 const double* __restrict__ meas, const double* __restrict__ conf, int n, double w_pos) {
  __shared__ double sP[64], sA[64], sB[64], sK[32], sKR[32], sS[16], sSI[16], sx[8], sy[4], sR[4];
  const int t = blockIdx.x;
  if (t >= n) return;
  const int s = slots[t], tid = threadIdx.x, i = tid >> 3, j = tid & 7;
  sP[tid] = P[(size_t)s * 64 + tid];
  if (tid < 8) sx[tid] = x[(size_t)s * 8 + tid];
  __syncthreads();
  if (tid < 4) {
    const double c = conf ? conf[t] : 0.0;
    const double base = tid == 2 ? 1e-1 : w_pos * sx[3];
    const double std = (1 - c) * base;
    sR[tid] = std * std;
    sy[tid] = meas[4 * t + tid] - sx[tid];  // y = z - H x
  }
  __syncthreads();
  if (tid < 16) {  // S = H P H^T + R
    const int r = tid >> 2, c = tid & 3;
    sS[tid] = sP[8 * r + c] + (r == c ? sR[r] : 0.0);
  }
  __syncthreads();
  if (tid == 0) inv4(sS, sSI);
  __syncthreads();
  if (tid < 16) SI[(size_t)s * 16 + tid] = sSI[tid];
  if (tid < 32) {  // K = P H^T SI  (8 x 4)
    const int r = tid >> 2, c = tid & 3;
    double acc = 0.0;
    for (int k = 0; k < 4; ++k) acc = acc + sP[8 * r + k] * sSI[4 * k + c];
    sK[tid] = acc;
  }
  __syncthreads();
  if (tid < 8) {  // x = x + K y
    double acc = 0.0;
    for (int k = 0; k < 4; ++k) acc = acc + sK[4 * tid + k] * sy[k];
    x[(size_t)s * 8 + tid] = sx[tid] + acc;
  }
  if (tid < 32) sKR[tid] = sK[tid] * sR[tid & 3];  // K R (R diagonal)
  // A = I - K H
  sA[tid] = (i == j ? 1.0 : 0.0) - (j < 4 ? sK[4 * i + j] : 0.0);
  __syncthreads();
  {  // B = A P
    double acc = 0.0;
    for (int k = 0; k < 8; ++k) acc = acc + sA[8 * i + k] * sP[8 * k + j];
    sB[tid] = acc;
  }
  __syncthreads();
  {  // P = B A^T + (K R) K^T
    double acc = 0.0;
    for (int k = 0; k < 8; ++k) acc = acc + sB[8 * i + k] * sA[8 * j + k];
    double krk = 0.0;
    for (int k = 0; k < 4; ++k) krk = krk + sKR[4 * i + k] * sK[4 * j + k];
    P[(size_t)s * 64 + tid] = acc + krk;
  }
}

}  // namespace reidb200

using namespace reidb200;

extern "C" int reidb200_kf_initiate(double* x, double* P, double* SI, const int32_t* slots, const double* measurements, int n,
                                    double w_pos, double w_vel, void* stream) {
  if (n == 0) return REIDB200_OK;
  RB_REQUIRE(x && P && SI && slots && measurements && n > 0, "kf_initiate: bad argument");
  cudaStream_t st = (cudaStream_t)stream;
  { ProfScope _ps("kf_initiate_kernel", st); kf_initiate_kernel<<<n, 64, 0, st>>>(x, P, SI, slots, measurements, n, w_pos, w_vel); }
  RB_LAUNCH_CHECK("kf_initiate_kernel");
  return REIDB200_OK;
}

extern "C" int reidb200_kf_predict(double* x, double* P, const int32_t* slots, int n, double dt, double w_pos, double w_vel,
                                   void* stream) {
  if (n == 0) return REIDB200_OK;
  RB_REQUIRE(x && P && n > 0, "kf_predict: bad argument");
  cudaStream_t st = (cudaStream_t)stream;
  { ProfScope _ps("kf_predict_kernel", st); kf_predict_kernel<<<n, 64, 0, st>>>(x, P, slots, n, dt, w_pos, w_vel); }
  RB_LAUNCH_CHECK("kf_predict_kernel");
  return REIDB200_OK;
}

extern "C" int reidb200_kf_update(double* x, double* P, double* SI, const int32_t* slots, const double* measurements,
                                  const double* confidences, int n, double w_pos, void* stream) {
  if (n == 0) return REIDB200_OK;
  RB_REQUIRE(x && P && SI && slots && measurements && n > 0, "kf_update: bad argument");
  cudaStream_t st = (cudaStream_t)stream;
  { ProfScope _ps("kf_update_kernel", st); kf_update_kernel<<<n, 64, 0, st>>>(x, P, SI, slots, measurements, confidences, n, w_pos); }
  RB_LAUNCH_CHECK("kf_update_kernel");
  return REIDB200_OK;
}
