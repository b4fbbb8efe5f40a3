// (2) Distance matrix as a tcgen05 / TMA tensor-core GEMM with fp32-accurate operands.
//
// Every fp32 operand row is split into two bf16 planes, x = hi + lo (|x - hi - lo| <= 2^-17 |x|), and
// each 64-wide k-block issues three UMMAs into one fp32 TMEM accumulator:
//        acc += Ahi.Bhi^T + Ahi.Blo^T + Alo.Bhi^T            (the dropped lo.lo term is ~2^-18 relative)
// Epilogues restate reid/losses/utils.py:12-35 (euclidean_dist, cosine_dist), reid/evaluate.py:58
// (dot score) and faiss METRIC_L2 (reid/faiss_utils.py:97-111).
//
// Kernel shape: persistent, 1 CTA / SM, 6 warps: warp 0 = TMA producer, warp 1 = MMA issuer (+ TMEM
// owner), warps 2-5 = epilogue (TMEM -> registers -> smem transpose -> coalesced global stores).
// CTA tile 128 x 256, k-block 64 (one SWIZZLE_128B atom), 2 smem stages of 96 KB
// (Ahi, Alo, Bhi, Blo), 2 TMEM accumulator stages of 256 columns.
#include <cuda.h>

#include "gemm.cuh"
#include "ptx.cuh"

namespace reidb200 {

// ------------------------------------------------------------------------------------ packing
__global__ void __launch_bounds__(256)
pack_rows_kernel(const float* __restrict__ x, int rows, int64_t ldx, int d, int dp,
                 __nv_bfloat16* __restrict__ packed, float* __restrict__ sqnorm) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= rows) return;
  const float* src = x + (size_t)warp * ldx;
  __nv_bfloat16* hi = packed + (size_t)warp * 2 * dp;
  __nv_bfloat16* lo = hi + dp;
  float s = 0.f;
  for (int k = lane; k < dp; k += 32) {
    const float v = k < d ? __ldg(src + k) : 0.f;
    const __nv_bfloat16 h = __float2bfloat16_rn(v);
    const __nv_bfloat16 l = __float2bfloat16_rn(v - __bfloat162float(h));
    hi[k] = h;
    lo[k] = l;
    s = fmaf(v, v, s);
  }
  s = warp_sum(s);
  if (lane == 0 && sqnorm) sqnorm[warp] = s;
}

int pack_rows_launch(const float* x, int rows, int64_t ldx, int d, void* packed, float* sqnorm, cudaStream_t st) {
  if (rows == 0) return REIDB200_OK;
  const int dp = padded_dim(d);
  const int warps_per_block = 8;
  { ProfScope _ps("pack_rows_kernel", st); pack_rows_kernel<<<ceil_div(rows, warps_per_block), warps_per_block * 32, 0, st>>>(
      x, rows, ldx, d, dp, static_cast<__nv_bfloat16*>(packed), sqnorm); }
  RB_LAUNCH_CHECK("pack_rows_kernel");
  return REIDB200_OK;
}

// ------------------------------------------------------------------------------------ epilogue math
__device__ __forceinline__ float dist_epilogue(int kind, float acc, float xx, float yy) {
  switch (kind) {
    case REIDB200_DIST_DOT:
      return acc;
    case REIDB200_DIST_SQL2:
      return fmaf(-2.0f, acc, __fadd_rn(xx, yy));
    case REIDB200_DIST_EUCLID:
      return sqrtf(fmaxf(fmaf(-2.0f, acc, __fadd_rn(xx, yy)), 1e-12f));
    default: {  // COSINE_HALF: (1 - xy / (|x| |y|)) / 2
      const float den = __fmul_rn(sqrtf(xx), sqrtf(yy));
      return __fmul_rn(__fsub_rn(1.0f, __fdiv_rn(acc, den)), 0.5f);
    }
  }
}

// ------------------------------------------------------------------------------------ tcgen05 kernel
constexpr int BM = 128, BN = 256, BK = 64;
constexpr int STAGES = 2;
constexpr int A_BYTES = BM * BK * 2;                      // 16 KB
constexpr int B_BYTES = BN * BK * 2;                      // 32 KB
constexpr int STAGE_BYTES = 2 * A_BYTES + 2 * B_BYTES;    // 96 KB
constexpr int EPI_STAGE_FLOATS = 32 * 33;                 // per epilogue warp
constexpr int GEMM_THREADS = 192;
constexpr int TMEM_COLS = 512;
constexpr size_t GEMM_SMEM = 1024 + (size_t)STAGES * STAGE_BYTES + 4 * EPI_STAGE_FLOATS * 4 + 128;

struct GemmArgs {
  int m, n, dp, num_kb, tiles_m, tiles_n, kind;
  const float* xx;
  const float* yy;
  float* out;
  int64_t ldo;
};

__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_split_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, GemmArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  float* epi_buf = reinterpret_cast<float*>(smem + (size_t)STAGES * STAGE_BYTES);
  uint64_t* bars = reinterpret_cast<uint64_t*>(epi_buf + 4 * EPI_STAGE_FLOATS);
  // bars: [0,STAGES) full, [STAGES,2*STAGES) empty, then tmem_full[2], tmem_empty[2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;
  const uint32_t bar_full = ptx::smem_u32(bars);
  const uint32_t bar_empty = ptx::smem_u32(bars + STAGES);
  const uint32_t bar_tfull = ptx::smem_u32(bars + 2 * STAGES);
  const uint32_t bar_tempty = ptx::smem_u32(bars + 2 * STAGES + 2);
  const uint32_t smem_base = ptx::smem_u32(smem);

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) {
      ptx::mbar_init(bar_full + 8 * s, 1);
      ptx::mbar_init(bar_empty + 8 * s, 1);
    }
    for (int s = 0; s < 2; ++s) {
      ptx::mbar_init(bar_tfull + 8 * s, 1);
      ptx::mbar_init(bar_tempty + 8 * s, 4);  // one arrival per epilogue warp
    }
    ptx::fence_barrier_init();
    ptx::prefetch_tensormap(&tmA);
    ptx::prefetch_tensormap(&tmB);
  }
  if (warp == 1) {
    ptx::tmem_alloc(ptx::smem_u32(tmem_slot), TMEM_COLS);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int total_tiles = a.tiles_m * a.tiles_n;

  if (warp == 0) {
    // ===== TMA producer (one thread)
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int tm = tile % a.tiles_m, tn = tile / a.tiles_m;
        for (int kb = 0; kb < a.num_kb; ++kb) {
          ptx::mbar_wait(bar_empty + 8 * stage, phase ^ 1);
          const uint32_t full = bar_full + 8 * stage;
          const uint32_t sb = smem_base + stage * STAGE_BYTES;
          ptx::mbar_arrive_expect_tx(full, STAGE_BYTES);
          ptx::tma_load_2d(sb, &tmA, full, kb * BK, tm * BM);
          ptx::tma_load_2d(sb + A_BYTES, &tmA, full, a.dp + kb * BK, tm * BM);
          ptx::tma_load_2d(sb + 2 * A_BYTES, &tmB, full, kb * BK, tn * BN);
          ptx::tma_load_2d(sb + 2 * A_BYTES + B_BYTES, &tmB, full, a.dp + kb * BK, tn * BN);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (one thread issues every UMMA and every commit)
    if (lane == 0) {
      constexpr uint32_t idesc = ptx::umma_idesc_bf16(BM, BN);
      int stage = 0, as = 0;
      uint32_t phase = 0, aphase = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        ptx::mbar_wait(bar_tempty + 8 * as, aphase ^ 1);
        ptx::tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BN;
        for (int kb = 0; kb < a.num_kb; ++kb) {
          ptx::mbar_wait(bar_full + 8 * stage, phase);
          ptx::tc_fence_after();
          const uint32_t sb = smem_base + stage * STAGE_BYTES;
#pragma unroll
          for (int kk = 0; kk < BK / 16; ++kk) {
            const uint64_t ahi = ptx::umma_desc_k_sw128(sb + kk * 32);
            const uint64_t alo = ptx::umma_desc_k_sw128(sb + A_BYTES + kk * 32);
            const uint64_t bhi = ptx::umma_desc_k_sw128(sb + 2 * A_BYTES + kk * 32);
            const uint64_t blo = ptx::umma_desc_k_sw128(sb + 2 * A_BYTES + B_BYTES + kk * 32);
            ptx::mma_f16_ss(d_tmem, alo, bhi, idesc, (kb | kk) != 0);
            ptx::mma_f16_ss(d_tmem, ahi, blo, idesc, 1);
            ptx::mma_f16_ss(d_tmem, ahi, bhi, idesc, 1);
          }
          ptx::mma_commit(bar_empty + 8 * stage);  // smem stage reusable once these UMMAs retire
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        ptx::mma_commit(bar_tfull + 8 * as);  // accumulator complete
        as ^= 1;
        if (as == 0) aphase ^= 1;
      }
    }
  } else {
    // ===== epilogue warps: TMEM lane quadrant = warp % 4
    const int q = warp & 3;
    float* sbuf = epi_buf + (warp - 2) * EPI_STAGE_FLOATS;
    int as = 0;
    uint32_t aphase = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      const int tm = tile % a.tiles_m, tn = tile / a.tiles_m;
      ptx::mbar_wait(bar_tfull + 8 * as, aphase);
      ptx::tc_fence_after();
      const int row0 = tm * BM + q * 32;
      const int my_row = row0 + lane;
      const float xx_mine = (a.xx != nullptr && my_row < a.m) ? __ldg(a.xx + my_row) : 0.f;
      const int rows_here = min(32, a.m - row0);
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        const int col0 = tn * BN + c * 32;
        if (col0 >= a.n) break;  // warp-uniform
        uint32_t r[32];
        ptx::tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(q * 32) << 16) + as * BN + c * 32, r);
        ptx::tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) sbuf[lane * 33 + j] = __uint_as_float(r[j]);
        __syncwarp();
        const int col = col0 + lane;
        const float yy_c = (a.yy != nullptr && col < a.n) ? __ldg(a.yy + col) : 0.f;
        for (int rr = 0; rr < rows_here; ++rr) {
          const float acc = sbuf[rr * 33 + lane];
          const float xr = __shfl_sync(0xffffffffu, xx_mine, rr);
          if (col < a.n) __stcs(a.out + (size_t)(row0 + rr) * a.ldo + col, dist_epilogue(a.kind, acc, xr, yy_c));
        }
        __syncwarp();
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(bar_tempty + 8 * as);
      as ^= 1;
      if (as == 0) aphase ^= 1;
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// ------------------------------------------------------------------------------------ SIMT cross-check
// Plain fp32 FMA GEMM with the same epilogues; used by the GPU tests as an on-device second opinion.
__global__ void __launch_bounds__(256)
distmat_simt_kernel(const float* __restrict__ x, int m, int64_t ldx, const float* __restrict__ y, int n, int64_t ldy,
                    int d, int kind, float* __restrict__ out, int64_t ldo) {
  __shared__ float sx[16][17], sy[16][17];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int row = blockIdx.y * 16 + ty, col = blockIdx.x * 16 + tx;
  float acc = 0.f, xx = 0.f, yy = 0.f;
  for (int k0 = 0; k0 < d; k0 += 16) {
    const int rx = blockIdx.y * 16 + ty, ry = blockIdx.x * 16 + ty;
    sx[ty][tx] = (rx < m && k0 + tx < d) ? x[(size_t)rx * ldx + k0 + tx] : 0.f;
    sy[ty][tx] = (ry < n && k0 + tx < d) ? y[(size_t)ry * ldy + k0 + tx] : 0.f;
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      const float xv = sx[ty][k], yv = sy[tx][k];
      acc = fmaf(xv, yv, acc);
      xx = fmaf(xv, xv, xx);
      yy = fmaf(yv, yv, yy);
    }
    __syncthreads();
  }
  if (row < m && col < n) out[(size_t)row * ldo + col] = dist_epilogue(kind, acc, xx, yy);
}

// ------------------------------------------------------------------------------------ tensor maps
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn) return fn;
  void* p = nullptr;
  cudaDriverEntryPointQueryResult qres;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess ||
      qres != cudaDriverEntryPointSuccess || p == nullptr)
    return nullptr;
  fn = reinterpret_cast<EncodeTiledFn>(p);
  return fn;
}

// packed operand [rows, 2*dp] bf16, box = {64, box_rows}, 128-byte swizzle, OOB rows/cols read as zero
static int make_operand_map(CUtensorMap* tm, const void* packed, int rows, int dp, int box_rows) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) {
    set_error("cuTensorMapEncodeTiled not available from the driver");
    return REIDB200_ECUDA;
  }
  cuuint64_t gdim[2] = {(cuuint64_t)2 * dp, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)2 * dp * sizeof(__nv_bfloat16)};
  cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(packed), gdim, gstride, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (%d) rows=%d dp=%d", (int)r, rows, dp);
    return REIDB200_ECUDA;
  }
  return REIDB200_OK;
}

int gemm_split_launch(const void* packedA, const float* xx, int m, const void* packedB, const float* yy, int n, int d,
                      int kind, float* out, int64_t ldo, cudaStream_t st) {
  if (m == 0 || n == 0) return REIDB200_OK;
  RB_REQUIRE(kind >= 0 && kind <= 3, "distmat: unknown kind %d", kind);
  RB_REQUIRE(kind == REIDB200_DIST_DOT || (xx && yy), "distmat: row norms required for kind %d", kind);
  static bool attr_set = false;
  if (!attr_set) {
    RB_CUDA(cudaFuncSetAttribute(gemm_split_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GEMM_SMEM));
    attr_set = true;
  }
  const int dp = padded_dim(d);
  CUtensorMap tmA, tmB;
  int rc = make_operand_map(&tmA, packedA, m, dp, BM);
  if (rc) return rc;
  rc = make_operand_map(&tmB, packedB, n, dp, BN);
  if (rc) return rc;
  GemmArgs a;
  a.m = m; a.n = n; a.dp = dp; a.num_kb = dp / BK;
  a.tiles_m = ceil_div(m, BM); a.tiles_n = ceil_div(n, BN);
  a.kind = kind; a.xx = xx; a.yy = yy; a.out = out; a.ldo = ldo;
  const long long total = (long long)a.tiles_m * a.tiles_n;
  const int grid = (int)(total < num_sms() ? total : num_sms());
  { ProfScope _ps("gemm_split_tc_kernel", st); gemm_split_tc_kernel<<<grid, GEMM_THREADS, GEMM_SMEM, st>>>(tmA, tmB, a); }
  RB_LAUNCH_CHECK("gemm_split_tc_kernel");
  return REIDB200_OK;
}

}  // namespace reidb200

using namespace reidb200;

extern "C" size_t reidb200_pack_bytes(int rows, int d) {
  if (rows <= 0 || d <= 0) return 0;
  return align_up(packed_bytes(rows, d), 256);
}

extern "C" int reidb200_pack_rows(const float* x, int rows, int64_t ldx, int d, void* packed, float* sqnorm,
                                  void* stream) {
  if (rows == 0) return REIDB200_OK;
  RB_REQUIRE(x && packed && rows > 0 && d > 0 && ldx >= d, "pack_rows: bad argument");
  RB_REQUIRE((reinterpret_cast<uintptr_t>(packed) & 15) == 0, "pack_rows: packed buffer must be 16-byte aligned");
  return pack_rows_launch(x, rows, ldx, d, packed, sqnorm, (cudaStream_t)stream);
}

extern "C" size_t reidb200_distmat_workspace_bytes(int m, int n, int d) {
  if (m <= 0 || n <= 0 || d <= 0) return 0;
  CarveSize c;
  c.take<char>(packed_bytes(m, d));
  c.take<float>(m);
  c.take<char>(packed_bytes(n, d));
  c.take<float>(n);
  return align_up(c.off, 256);
}

extern "C" int reidb200_distmat(const float* x, int m, int64_t ldx, const float* y, int n, int64_t ldy, int d, int kind,
                                float* out, int64_t ldo, void* workspace, size_t workspace_bytes, void* stream) {
  if (m == 0 || n == 0) return REIDB200_OK;
  RB_REQUIRE(x && y && out && m > 0 && n > 0 && d > 0, "distmat: bad argument m=%d n=%d d=%d", m, n, d);
  RB_REQUIRE(ldx >= d && ldy >= d && ldo >= n, "distmat: leading dimension too small");
  RB_REQUIRE(workspace && workspace_bytes >= reidb200_distmat_workspace_bytes(m, n, d), "distmat: workspace too small");
  RB_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, "distmat: workspace must be 256-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  Carver c(workspace, workspace_bytes);
  void* pa = c.take<char>(packed_bytes(m, d));
  float* xx = c.take<float>(m);
  void* pb = c.take<char>(packed_bytes(n, d));
  float* yy = c.take<float>(n);
  int rc = pack_rows_launch(x, m, ldx, d, pa, xx, st);
  if (rc) return rc;
  if (x == y && m == n && ldx == ldy) {
    pb = pa;
    yy = xx;
  } else {
    rc = pack_rows_launch(y, n, ldy, d, pb, yy, st);
    if (rc) return rc;
  }
  return gemm_split_launch(pa, xx, m, pb, yy, n, d, kind, out, ldo, st);
}

extern "C" int reidb200_distmat_simt(const float* x, int m, int64_t ldx, const float* y, int n, int64_t ldy, int d,
                                     int kind, float* out, int64_t ldo, void* stream) {
  if (m == 0 || n == 0) return REIDB200_OK;
  RB_REQUIRE(x && y && out && m > 0 && n > 0 && d > 0 && kind >= 0 && kind <= 3, "distmat_simt: bad argument");
  dim3 grid(ceil_div(n, 16), ceil_div(m, 16));
  { ProfScope _ps("distmat_simt_kernel", (cudaStream_t)stream); distmat_simt_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(x, m, ldx, y, n, ldy, d, kind, out, ldo); }
  RB_LAUNCH_CHECK("distmat_simt_kernel");
  return REIDB200_OK;
}
