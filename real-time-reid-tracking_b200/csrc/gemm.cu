// (2) Distance matrix as a tcgen05 / TMA tensor-core GEMM with fp32-accurate operands.
//
// Every fp32 operand row is scaled by a power of two (its largest magnitude lands in [2^14, 2^15): exact, and it keeps
// both planes in fp16's normal range) and split into two fp16 planes, s.x = hi + lo with |s.x - hi - lo| <= 2^-22 |s.x|
// (fp16 carries 11 significand bits, so the pair carries 22 of fp32's 24).  Each 64-wide k-block issues three UMMAs
// (kind::f16, fp32 accumulation in TMEM) into one accumulator:
//        acc += Alo.Bhi^T + Ahi.Blo^T + Ahi.Bhi^T            (the dropped lo.lo term is ~2^-22 relative)
// and the epilogue multiplies by the two inverse row scales (exact).  A bf16 split at the same cost would leave
// 2^-18 per element -- 16x more error than this and 64x more than fp32.
// Epilogues restate reid/losses/utils.py:12-35 (euclidean_dist, cosine_dist), reid/evaluate.py:58
// (dot score) and faiss METRIC_L2 (reid/faiss_utils.py:97-111).
//
// Kernel shape: persistent, 1 CTA / SM, 6 warps: warp 0 = TMA producer, warp 1 = MMA issuer (+ TMEM
// owner), warps 2-5 = epilogue (TMEM -> registers -> smem transpose -> coalesced global stores).
// CTA tile 128 x 256, k-block 64 (one SWIZZLE_128B atom), 2 smem stages of 96 KB
// (Ahi, Alo, Bhi, Blo), 2 TMEM accumulator stages of 256 columns.
#include <stdlib.h>
#include <string.h>

#include "gemm.cuh"
#include "ptx.cuh"
#include <math.h>

namespace reidb200 {

// ------------------------------------------------------------------------------------ packing
__global__ void __launch_bounds__(256)
pack_rows_kernel(const float* __restrict__ x, int rows, int64_t ldx, int d, int dp, __half* __restrict__ packed,
                 float* __restrict__ sqnorm, float* __restrict__ iscale) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= rows) return;
  const float* src = x + (size_t)warp * ldx;
  float s = 0.f, mx = 0.f;
  for (int k = lane; k < d; k += 32) {
    const float v = __ldg(src + k);
    s = fmaf(v, v, s);
    mx = fmaxf(mx, fabsf(v));
  }
  s = warp_sum(s);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  // scale = 2^(14 - floor(log2 max)): the row maximum lands in [2^14, 2^15); exponent clamped so the scale stays finite
  int e = 0;
  if (mx > 0.f && mx <= 3.0e38f) e = min(max(14 - ilogbf(mx), -100), 100);
  const float scale = ldexpf(1.0f, e);
  __half* hi = packed + (size_t)warp * 2 * dp;
  __half* lo = hi + dp;
  for (int k = lane; k < dp; k += 32) {
    const float v = k < d ? __fmul_rn(__ldg(src + k), scale) : 0.f;
    const __half h = __float2half_rn(v);
    hi[k] = h;
    lo[k] = __float2half_rn(__fsub_rn(v, __half2float(h)));
  }
  if (lane == 0) {
    if (sqnorm) sqnorm[warp] = s;
    iscale[warp] = ldexpf(1.0f, -e);
  }
}

int pack_rows_launch(const float* x, int rows, int64_t ldx, int d, void* packed, float* sqnorm, float* iscale,
                     cudaStream_t st) {
  if (rows == 0) return REIDB200_OK;
  const int dp = padded_dim(d);
  const int warps_per_block = 8;
  { ProfScope _ps("pack_rows_kernel", st); pack_rows_kernel<<<ceil_div(rows, warps_per_block), warps_per_block * 32, 0, st>>>(
      x, rows, ldx, d, dp, static_cast<__half*>(packed), sqnorm, iscale); }
  RB_LAUNCH_CHECK("pack_rows_kernel");
  return REIDB200_OK;
}

// ------------------------------------------------------------------------------------ tcgen05 kernel
constexpr int BM = 128, BN = 256, BK = 64;
constexpr int STAGES = 2;
constexpr int A_BYTES = BM * BK * 2;                      // 16 KB
constexpr int B_BYTES = BN * BK * 2;                      // 32 KB
constexpr int STAGE_BYTES = 2 * A_BYTES + 2 * B_BYTES;    // 96 KB
constexpr int EPI_WARPS = 8;                              // two per TMEM lane quadrant (SM sub-partition)
constexpr int EPI_STAGE_FLOATS = 32 * 32;                 // per epilogue warp: one 32 x 32 fp32 output box (TMA store source)
constexpr int GEMM_THREADS = 64 + 32 * EPI_WARPS;         // TMA producer warp, MMA warp, epilogue warps
constexpr int TMEM_COLS = 512;
constexpr int EPI_YW_FLOATS = EPI_WARPS * 64;             // per epilogue warp: column norms | column scale factors of a chunk
// dynamic shared memory is declared 1024-byte aligned (SWIZZLE_128B operand tiles); the kernels trap if it is not
constexpr size_t GEMM_SMEM = (size_t)STAGES * STAGE_BYTES + EPI_WARPS * EPI_STAGE_FLOATS * 4 + EPI_YW_FLOATS * 4 + 128;

struct GemmArgs {
  int m, n, dp, num_kb, tiles_m, tiles_n, kind;
  const float* xx;
  const float* yy;
  const float* isa;  // inverse power-of-two scale of every A row / B row (pack_rows_kernel)
  const float* isb;
  float* out;
  int64_t ldo;
  // optional per-(row, 32-column chunk) reduction of the epilogue values (min, or max when cred_max): lets the
  // top-k selection read only the k most promising 128-byte chunks of each row instead of the whole row
  float* cred;
  int64_t ldcr;
  int cred_max;
  int tma_out;  // the output tensor map is valid (16-byte aligned rows): epilogue stores through the TMA engine
  int symmetric;  // A == B (self distances): only tiles with tn >= tm are computed, each also stored transposed
  // TRIANGLE mode of a symmetric product (`lower` != 0; the fused re-ranked evaluation): the tiles are computed with the
  // roles of the two blocks swapped (row block >= column block) and stored ONCE -- element (i, c) exists iff
  // (i >> 8) >= (c >> 8); nothing is stored transposed (half of the output bytes).  A row's view of the missing half is
  // what the chunk-guided selection needs: per (mirror row, 32-column chunk) the minimum in `cred` as before and, in
  // `cred2`, (ordered key of the SECOND smallest value & ~31) | position of the minimum inside the chunk.  A chunk whose
  // second value cannot pass the row's bound contributes exactly its minimum, so topk.cu reads it from these two words;
  // only chunks with two or more passing elements are gathered through the transposed location.  The two cross terms
  // of the split product are issued in the opposite order so that every accumulator -- hence every stored value -- is
  // bit-identical to the one the upper-triangle kernel produces for the same pair.
  int lower;
  uint32_t* cred2;
  // box-major output (the re-ranking workspace): the matrix is stored as 32 x 32 boxes of 4 KB, box (rb, cb) at float
  // offset (rb * blocked + cb) * 1024, row-major inside; blocked = boxes per box row (0 = plain row-major).  Every
  // epilogue store is then one contiguous 4 KB piece of DRAM instead of 32 pieces of 128 B a row stride apart.
  int blocked;
  int debug;      // timing experiments only (REIDB200_GEMM_DEBUG): 1 = no mirror store, 2 = no stores, 4 = no epilogue
  SelArgs sel;    // sel.tau != null: selection fused into the epilogue (gemm.cuh); only sel.st_* of the output is stored
};

// tile index -> (tm, tn) of the CTA-pair kernel.  Symmetric mode enumerates the upper triangle column by column:
// tile = tn (tn + 1) / 2 + tm, tm <= tn.
__device__ __forceinline__ void pair_tile_coords(int tile, int tiles_m2, int symmetric, int& tm, int& tn) {
  if (!symmetric) { tm = tile % tiles_m2; tn = tile / tiles_m2; return; }
  if (symmetric == 2) {
    // selection mode: BAND order -- band d holds the tiles (tm, tm + d), bands d = 0 (the diagonal: every row block's
    // bootstrap tile), 1, 2, ...  Tiles that run at the same time then belong to different row AND column blocks (a block
    // appears in at most two tiles of a band), so no row's list is appended to by many tiles at once while its threshold
    // is still loose (column-by-column order put up to 74 tiles of one column block in flight: their mirrored appends
    // flooded that block's lists before any of them could tighten).
    const float s2 = 2.0f * (float)tiles_m2 + 1.0f;
    int d = (int)((s2 - sqrtf(fmaxf(s2 * s2 - 8.0f * (float)tile, 0.0f))) * 0.5f);
    auto off = [&](int dd) { return dd * tiles_m2 - dd * (dd - 1) / 2; };  // tiles before band dd
    while (d > 0 && off(d) > tile) --d;
    while (off(d + 1) <= tile) ++d;
    tm = tile - off(d);
    tn = tm + d;
    return;
  }
  int b = (int)((sqrtf(8.0f * (float)tile + 1.0f) - 1.0f) * 0.5f);
  while ((b + 1) * (b + 2) / 2 <= tile) ++b;
  while (b * (b + 1) / 2 > tile) --b;
  tn = b;
  tm = tile - b * (b + 1) / 2;
}

// v[j] = value of column j in this lane's row -> reduction (min, or max) over the 32 lanes (rows) of column `lane`
template <bool MAX>
__device__ __forceinline__ float colred_transpose(float (&v)[32], int lane) {
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    const bool up = (lane & off) != 0;
#pragma unroll
    for (int i = 0; i < off; ++i) {
      const float send = up ? v[i] : v[i + off];
      const float keep = up ? v[i + off] : v[i];
      const float got = __shfl_xor_sync(0xffffffffu, send, off);
      v[i] = MAX ? fmaxf(keep, got) : fminf(keep, got);
    }
  }
  return v[0];
}


// value of one output element from the raw accumulator r.  Operands were scaled by powers of two (ia = inverse row
// scale, s = inverse column scale), so every scaling is exact and the fused form rounds like the plain one:
// fma(-2, r ia s, xx + yy) == fma(r ia, -2 s, xx + yy).  sv = -2 s for the two L2 kinds, s otherwise (epi_colscale).
template <int KIND>
__device__ __forceinline__ float epi_colscale(float s) {
  return (KIND == REIDB200_DIST_SQL2 || KIND == REIDB200_DIST_EUCLID) ? __fmul_rn(-2.0f, s) : s;
}
template <int KIND>
__device__ __forceinline__ float epi_value(float r, float ia, float sv, float xx, float yy) {
  const float ra = __fmul_rn(r, ia);
  if (KIND == REIDB200_DIST_SQL2) return fmaf(ra, sv, __fadd_rn(xx, yy));
  if (KIND == REIDB200_DIST_EUCLID) return sqrtf(fmaxf(fmaf(ra, sv, __fadd_rn(xx, yy)), 1e-12f));
  return dist_epilogue(KIND, __fmul_rn(ra, sv), xx, yy);
}

// ---- selection in the epilogue (SelArgs, gemm.cuh)
__device__ __forceinline__ uint32_t sel_key_of(float v, int descending) {
  const uint32_t u = float_to_ordered(v);
  return descending ? ~u : u;
}

// One thread appends the elements of `v[0..32)` whose key passes `tau` to `row`'s list and sketch: the passing ones are
// counted first, ONE atomicAdd reserves their slots (every lane of the warp issues its own in the same instruction, so a
// chunk costs one L2 round trip however many lanes append), then the entries are written and folded into the sketch by
// reductions that return nothing.  idx of element j = idx0 + j.  `valid` = bit j set when element j exists.  Returns the
// row when the append crossed a multiple of the window (the warp then tightens it), else -1.
__device__ __forceinline__ int sel_append32(const SelArgs& s, int row, const float (&v)[32], uint32_t valid, uint32_t tau, int idx0) {
  uint32_t mask = 0;
#pragma unroll
  for (int j = 0; j < 32; ++j) mask |= (sel_key_of(v[j], s.descending) <= tau ? 1u : 0u) << j;
  mask &= valid;
  if (mask == 0 || (s.debug & 1)) return -1;
  const uint32_t cnt = __popc(mask);
  const uint32_t base = atomicAdd(s.resv + row, cnt);
  unsigned long long* lrow = s.list + (size_t)row * s.cap;
  unsigned long long* srow = s.sketch + (size_t)row * s.W;
  uint32_t pos = base;
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    if ((mask >> j) & 1u) {
      const unsigned long long e = ((unsigned long long)sel_key_of(v[j], s.descending) << 32) | (uint32_t)(idx0 + j);
      if (pos < (uint32_t)s.cap) lrow[pos] = e;
      atomicMin(srow + ((uint32_t)(idx0 + j) & (uint32_t)(s.W - 1)), e);  // W is a power of two
      ++pos;
    }
  }
  return (base / (uint32_t)s.window) != ((base + cnt) / (uint32_t)s.window) ? row : -1;
}

// One warp: tau[row] <- min(tau[row], k-th smallest occupied slot of the row's sketch) (see SelArgs).  The W <= 512 slot
// keys are sorted in registers (S per lane, warp_bitonic_sort) and element k - 1 is read off.
template <int S>
__device__ __forceinline__ void sel_tighten_sorted(const SelArgs& s, int row, int lane) {
  const unsigned long long* e = s.sketch + (size_t)row * s.W;
  uint32_t key[S], nov[S];
#pragma unroll
  for (int i = 0; i < S; ++i) {
    const int slot = lane * S + i;
    key[i] = slot < s.W ? (uint32_t)(__ldcg(e + slot) >> 32) : 0xffffffffu;  // an empty slot reads as key 0xFFFFFFFF
  }
  warp_bitonic_sort<S, false>(key, nov, lane);
  const int pos = s.k - 1;
  uint32_t mine = 0xffffffffu;
#pragma unroll
  for (int i = 0; i < S; ++i) mine = (pos % S) == i ? key[i] : mine;
  const uint32_t kth = __shfl_sync(0xffffffffu, mine, pos / S);
  if (lane == 0 && kth != 0xffffffffu) atomicMin(s.tau + row, kth);
}

__device__ __noinline__ void sel_tighten_row(const SelArgs& s, int row, int lane) {
  if (s.W <= 64) sel_tighten_sorted<2>(s, row, lane);
  else if (s.W <= 128) sel_tighten_sorted<4>(s, row, lane);
  else sel_tighten_sorted<16>(s, row, lane);
}

// Epilogue of the accumulator columns owned by one warp: TMEM lane quadrant q = warp % 4 (thread = row), and every
// second 32-column chunk of the 256-column tile (two warps share a quadrant: chunks c0, c0 + 2, ...), so that each
// SM sub-partition has two epilogue warps to hide each other's latencies.
// Fast path (16-byte aligned rows): the warp turns the chunk into output values in registers, stages the 32 x 32 box
// in shared memory (128-byte rows, 16-byte pieces XOR-swizzled by row & 7: four conflict-free wavefronts per 512-byte
// store) and one lane hands it to the TMA store engine, which clips the ragged matrix edges.  Slow path (odd leading
// dimension): transpose through the same shared-memory tile and store row-contiguous scalars.
// Latency hiding: the column norms / scales of the warp's next chunk and its accumulator columns (tcgen05.ld) are
// requested while the current chunk is processed; the first chunk's norms are requested BEFORE the wait on the
// accumulator-full barrier; the wait for the staging buffer (previous TMA store has read it) comes after the math.
// `yw` = 64 floats of shared memory private to this warp (column norms | column scale factors of the current chunk,
// read back as broadcast 128-bit loads).
template <int KIND>
__device__ __forceinline__ void epilogue_slab(const GemmArgs& a, const CUtensorMap* tmO, uint32_t bar_full, uint32_t parity,
                                              uint32_t tmem_addr, int row0, int col_base, int c_first, int lane, float* sb,
                                              float* yw, bool vec_ok, bool sym_t = false) {
  const int my_row = row0 + lane;
  const int rows_here = min(32, a.m - row0);
  const int nch = rows_here <= 0 ? 0 : min(BN / 32, (a.n - col_base + 31) >> 5);  // warp-uniform
  float xx_mine = 0.f, ia_mine = 0.f, yn = 0.f, sn = 0.f;
  if (c_first < nch) {
    if (my_row < a.m) {
      if (KIND != REIDB200_DIST_DOT) xx_mine = __ldg(a.xx + my_row);
      ia_mine = __ldg(a.isa + my_row);
    }
    if (col_base + c_first * 32 + lane < a.n) {
      if (KIND != REIDB200_DIST_DOT) yn = __ldg(a.yy + col_base + c_first * 32 + lane);
      sn = __ldg(a.isb + col_base + c_first * 32 + lane);
    }
  }
  ptx::mbar_wait(bar_full, parity);
  ptx::tc_fence_after();
  if (c_first >= nch || (a.debug & 4)) return;
  if (a.debug & 1) sym_t = false;
  if (a.debug & 2) vec_ok = false;
  const bool want_red = a.cred != nullptr;
  const bool red_max = a.cred_max != 0;
  uint32_t rn[32];
  ptx::tmem_ld_32x32b_x32(tmem_addr + c_first * 32, rn);
#pragma unroll 1
  for (int c = c_first; c < nch; c += 2) {
    const int col0 = col_base + c * 32;
    yw[lane] = yn;
    yw[32 + lane] = epi_colscale<KIND>(sn);
    float o[32];
    ptx::tmem_ld_wait(rn);
#pragma unroll
    for (int j = 0; j < 32; ++j) o[j] = __uint_as_float(rn[j]);
    if (c + 2 < nch) {
      ptx::tmem_ld_32x32b_x32(tmem_addr + (c + 2) * 32, rn);
      yn = 0.f; sn = 0.f;
      if (col0 + 64 + lane < a.n) {
        if (KIND != REIDB200_DIST_DOT) yn = __ldg(a.yy + col0 + 64 + lane);
        sn = __ldg(a.isb + col0 + 64 + lane);
      }
    }
    __syncwarp();
    const bool full = col0 + 32 <= a.n;
    // ---- math: raw accumulators -> output values, in registers (thread = row); row reduction over the chunk
    float red = red_max ? -INFINITY : INFINITY;
#pragma unroll
    for (int j = 0; j < 32; j += 4) {
      const float4 y4 = *reinterpret_cast<const float4*>(yw + j);  // broadcast reads (zero beyond column n)
      const float4 s4 = *reinterpret_cast<const float4*>(yw + 32 + j);
      o[j + 0] = epi_value<KIND>(o[j + 0], ia_mine, s4.x, xx_mine, y4.x);
      o[j + 1] = epi_value<KIND>(o[j + 1], ia_mine, s4.y, xx_mine, y4.y);
      o[j + 2] = epi_value<KIND>(o[j + 2], ia_mine, s4.z, xx_mine, y4.z);
      o[j + 3] = epi_value<KIND>(o[j + 3], ia_mine, s4.w, xx_mine, y4.w);
    }
    if (full) {
      float t[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) t[j] = red_max ? fmaxf(o[j], o[j + 16]) : fminf(o[j], o[j + 16]);
#pragma unroll
      for (int w = 8; w >= 1; w >>= 1) {
#pragma unroll
        for (int j = 0; j < w; ++j) t[j] = red_max ? fmaxf(t[j], t[j + w]) : fminf(t[j], t[j + w]);
      }
      red = t[0];
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (col0 + j < a.n) red = red_max ? fmaxf(red, o[j]) : fminf(red, o[j]);
    }
    if (want_red && my_row < a.m) a.cred[(size_t)my_row * a.ldcr + (col0 >> 5)] = red;
    if (vec_ok) {
      // ---- direct box: wait until the previous TMA store has read the staging buffer, stage, store
      if (lane == 0) ptx::bulk_wait_read<0>();
      __syncwarp();
#pragma unroll
      for (int j = 0; j < 32; j += 4)
        *reinterpret_cast<float4*>(sb + lane * 32 + (((j >> 2) ^ (lane & 7)) << 2)) = make_float4(o[j], o[j + 1], o[j + 2], o[j + 3]);
      ptx::fence_proxy_async();
      __syncwarp();
      if (lane == 0) {
        if (a.blocked) ptx::tma_store_2d(tmO, ptx::smem_u32(sb), 0, ((row0 >> 5) * a.blocked + (col0 >> 5)) * 32);
        else ptx::tma_store_2d(tmO, ptx::smem_u32(sb), col0, row0);
        ptx::bulk_commit();
      }
      if (sym_t) {
        // ---- mirror block out[col0 + j, row0 + lane], straight from the registers: for every column j the warp
        // writes 32 consecutive floats of mirror row col0 + j (one coalesced 128-byte store; no staging buffer, so
        // nothing to wait for).  Its chunk reduction (mirror row = column col0 + lane, over this warp's rows) reads
        // the column back from the staged direct box, which the TMA engine is reading concurrently.
        if (want_red) {
          float v[32];
#pragma unroll
          for (int rr = 0; rr < 32; ++rr) v[rr] = sb[rr * 32 + ((((lane >> 2) ^ (rr & 7)) << 2) | (lane & 3))];
          if (rows_here < 32) {  // ragged last row block (warp-uniform): rows beyond m do not take part
            const float pad = red_max ? -INFINITY : INFINITY;
#pragma unroll
            for (int rr = 0; rr < 32; ++rr) v[rr] = rr < rows_here ? v[rr] : pad;
          }
          if (KIND == REIDB200_DIST_SQL2 && a.cred2) {
            // triangle mode (minimum only): tournament over (smallest, second smallest) pairs, then the place of the
            // smallest (the lowest one among equals; equal values make second == smallest, which sends the chunk to the
            // exact gather in topk.cu)
            float mn[16], sc[16];
#pragma unroll
            for (int rr = 0; rr < 16; ++rr) { mn[rr] = fminf(v[rr], v[rr + 16]); sc[rr] = fmaxf(v[rr], v[rr + 16]); }
#pragma unroll
            for (int w = 8; w >= 1; w >>= 1) {
#pragma unroll
              for (int rr = 0; rr < w; ++rr) {
                const float m1 = mn[rr], m2 = mn[rr + w];
                mn[rr] = fminf(m1, m2);
                sc[rr] = fminf(fmaxf(m1, m2), fminf(sc[rr], sc[rr + w]));
              }
            }
            int arg = 31;
#pragma unroll
            for (int rr = 30; rr >= 0; --rr) arg = v[rr] == mn[0] ? rr : arg;
            if (col0 + lane < a.n) {
              a.cred[(size_t)(col0 + lane) * a.ldcr + (row0 >> 5)] = mn[0];
              a.cred2[(size_t)(col0 + lane) * a.ldcr + (row0 >> 5)] = (float_to_ordered(sc[0]) & ~31u) | (uint32_t)arg;
            }
          } else {
#pragma unroll
            for (int w = 16; w >= 1; w >>= 1) {
#pragma unroll
              for (int rr = 0; rr < w; ++rr) v[rr] = red_max ? fmaxf(v[rr], v[rr + w]) : fminf(v[rr], v[rr + w]);
            }
            const float red_t = v[0];
            if (col0 + lane < a.n) a.cred[(size_t)(col0 + lane) * a.ldcr + (row0 >> 5)] = red_t;
          }
        }
        if (KIND == REIDB200_DIST_SQL2 && a.lower) {
          // triangle mode: nothing is stored transposed
        } else if (a.blocked) {  // box (col0 / 32, row0 / 32), padded to whole boxes: no edge guards
          float* mp = a.out + ((size_t)(col0 >> 5) * a.blocked + (row0 >> 5)) * 1024 + lane;
#pragma unroll
          for (int j = 0; j < 32; ++j) mp[j * 32] = o[j];
        } else if (my_row < a.m) {
          float* mp = a.out + (size_t)col0 * a.ldo + my_row;
          if (full) {
#pragma unroll
            for (int j = 0; j < 32; ++j) mp[(size_t)j * a.ldo] = o[j];
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (col0 + j < a.n) mp[(size_t)j * a.ldo] = o[j];
          }
        }
      }
    } else {
      // ---- slow path: transpose through shared memory (XOR-swizzled columns), row-contiguous scalar stores
#pragma unroll
      for (int j = 0; j < 32; ++j) sb[lane * 32 + (j ^ lane)] = o[j];
      __syncwarp();
      const int col = col0 + lane;
#pragma unroll 8
      for (int rr = 0; rr < 32; ++rr)
        if (rr < rows_here && col < a.n && !(a.debug & 2)) __stcs(a.out + (size_t)(row0 + rr) * a.ldo + col, sb[rr * 32 + (lane ^ rr)]);
      __syncwarp();
    }
  }
}

// Epilogue of one warp's accumulator columns in SELECTION mode: the same math as epilogue_slab, but instead of storing the
// tile the warp filters it against the rows' thresholds (and, for a mirrored tile of a symmetric product, against the
// thresholds of the columns).  Fast path per 32 x 32 chunk: the row minimum of the chunk (already needed) is compared with
// the row's threshold and with the largest threshold among the chunk's 32 columns -- two compares; only a thread that
// passes walks its 32 values and appends.  Rows whose list completed a window are tightened by the warp after the slab.
// Only the part of the output inside sel.st_* is still stored (the query x gallery block the lambda blend reads).
// `yw` = 64 floats of shared memory private to this warp: column norms (then the column thresholds, as uint) | column scales.
template <int KIND>
__device__ __forceinline__ void epilogue_select(const GemmArgs& a, const CUtensorMap* tmO, uint32_t bar_full, uint32_t parity,
                                                uint32_t tmem_addr, int row0, int col_base, int c_first, int lane, float* sb,
                                                float* yw, bool sym_t, bool boot, int gate_r, int gate_c, uint32_t gate_need) {
  const SelArgs& S = a.sel;
  if (!boot && !(S.debug & 4)) {  // the bootstrap tiles of this tile's row block (and column block, if mirrored) have set the thresholds
    if (lane == 0) {
      while (*reinterpret_cast<volatile uint32_t*>(S.gate + gate_r) < gate_need) __nanosleep(64);
      if (gate_c >= 0)
        while (*reinterpret_cast<volatile uint32_t*>(S.gate + gate_c) < gate_need) __nanosleep(64);
      __threadfence();
    }
    __syncwarp();
  }
  const int my_row = row0 + lane;
  const int rows_here = min(32, a.m - row0);
  const int nch = rows_here <= 0 ? 0 : min(BN / 32, (a.n - col_base + 31) >> 5);  // warp-uniform
  float xx_mine = 0.f, ia_mine = 0.f, yn = 0.f, sn = 0.f;
  uint32_t tau_row = 0u, tcn = 0u;
  if (c_first < nch) {
    if (my_row < a.m) {
      if (KIND != REIDB200_DIST_DOT) xx_mine = __ldg(a.xx + my_row);
      ia_mine = __ldg(a.isa + my_row);
      tau_row = __ldcg(S.tau + my_row);  // (a stale value is only a looser filter)
    }
    if (col_base + c_first * 32 + lane < a.n) {
      if (KIND != REIDB200_DIST_DOT) yn = __ldg(a.yy + col_base + c_first * 32 + lane);
      sn = __ldg(a.isb + col_base + c_first * 32 + lane);
      if (sym_t) tcn = __ldcg(S.tau + col_base + c_first * 32 + lane);
    }
  }
  ptx::mbar_wait(bar_full, parity);
  ptx::tc_fence_after();
  if (c_first >= nch) {  // nothing to do for this warp (ragged edge); a bootstrap tile still counts it
    if (boot && lane == 0) atomicAdd(S.gate + gate_r, 1u);
    return;
  }
  const bool st_rows = row0 < S.st_r1 && row0 + 32 > S.st_r0;        // direct block rows inside the stored region
  const bool st_rows_t = sym_t && row0 < S.st_c1 && row0 + 32 > S.st_c0;  // mirrored block: these rows are its COLUMNS
  int pending = -1;
  uint32_t rn[32];
  ptx::tmem_ld_32x32b_x32(tmem_addr + c_first * 32, rn);
#pragma unroll 1
  for (int c = c_first; c < nch; c += 2) {
    const int col0 = col_base + c * 32;
    yw[lane] = yn;
    yw[32 + lane] = epi_colscale<KIND>(sn);
    const uint32_t tc_cur = tcn;
    const uint32_t tc_max = __reduce_max_sync(0xffffffffu, tcn);
    float o[32];
    ptx::tmem_ld_wait(rn);
#pragma unroll
    for (int j = 0; j < 32; ++j) o[j] = __uint_as_float(rn[j]);
    // the row's threshold is re-read for every chunk (requested here, used by the next chunk): several tiles work on the
    // same row block at the same time, and each must see what the others' appends have tightened meanwhile
    uint32_t tau_next = tau_row;
    if (my_row < a.m) tau_next = __ldcg(S.tau + my_row);
    if (c + 2 < nch) {
      ptx::tmem_ld_32x32b_x32(tmem_addr + (c + 2) * 32, rn);
      yn = 0.f; sn = 0.f; tcn = 0u;
      if (col0 + 64 + lane < a.n) {
        if (KIND != REIDB200_DIST_DOT) yn = __ldg(a.yy + col0 + 64 + lane);
        sn = __ldg(a.isb + col0 + 64 + lane);
        if (sym_t) tcn = __ldcg(S.tau + col0 + 64 + lane);
      }
    }
    __syncwarp();
    const bool full = col0 + 32 <= a.n;
#pragma unroll
    for (int j = 0; j < 32; j += 4) {
      const float4 y4 = *reinterpret_cast<const float4*>(yw + j);
      const float4 s4 = *reinterpret_cast<const float4*>(yw + 32 + j);
      o[j + 0] = epi_value<KIND>(o[j + 0], ia_mine, s4.x, xx_mine, y4.x);
      o[j + 1] = epi_value<KIND>(o[j + 1], ia_mine, s4.y, xx_mine, y4.y);
      o[j + 2] = epi_value<KIND>(o[j + 2], ia_mine, s4.z, xx_mine, y4.z);
      o[j + 3] = epi_value<KIND>(o[j + 3], ia_mine, s4.w, xx_mine, y4.w);
    }
    // best value of this thread's row over the chunk (columns beyond n do not take part)
    float red = S.descending ? -INFINITY : INFINITY;
    if (full) {
      float t[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) t[j] = S.descending ? fmaxf(o[j], o[j + 16]) : fminf(o[j], o[j + 16]);
#pragma unroll
      for (int w = 8; w >= 1; w >>= 1) {
#pragma unroll
        for (int j = 0; j < w; ++j) t[j] = S.descending ? fmaxf(t[j], t[j + w]) : fminf(t[j], t[j + w]);
      }
      red = t[0];
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (col0 + j < a.n) red = S.descending ? fmaxf(red, o[j]) : fminf(red, o[j]);
    }
    const uint32_t kred = sel_key_of(red, S.descending);
    const bool row_ok = my_row < a.m;
    if (boot && (S.debug & 8)) {
    } else if (boot) {
      // bootstrap tile (threshold still +inf): every column is a candidate.  No counters: column c of the tile goes to slot
      // c of the row's list (the warp sets resv to the tile's valid column count afterwards; nobody else appends to these
      // rows before the gate opens) and into the sketch by a reduction that returns nothing.
      if (row_ok) {
        unsigned long long* lrow = S.list + (size_t)my_row * S.cap + (col0 - col_base);
        unsigned long long* srow = S.sketch + (size_t)my_row * S.W;
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          if (full || col0 + j < a.n) {
            const unsigned long long e = ((unsigned long long)sel_key_of(o[j], S.descending) << 32) | (uint32_t)(col0 + j);
            lrow[j] = e;
            atomicMin(srow + ((uint32_t)(col0 + j) % (uint32_t)S.W), e);
          }
        }
      }
    } else if (!(S.debug & 16)) {
      const uint32_t valid = full ? 0xffffffffu : ((1u << (a.n - col0)) - 1u);
      // direct: this thread's row against its own threshold (entered by the lanes whose row minimum passes)
      if (row_ok && kred <= tau_row) {
        const int p = sel_append32(S, my_row, o, valid, tau_row, col0);
        if (p >= 0) pending = p;
      }
      // mirror (symmetric product): the chunk is transposed through the staging box so that lane j holds column j, i.e. the
      // 32 elements (row0 .. row0 + 31) of the mirrored row col0 + j, and the same code runs against that row's threshold
      if (sym_t && __any_sync(0xffffffffu, row_ok && kred <= tc_max)) {
        if (lane == 0) ptx::bulk_wait_read<0>();
        __syncwarp();
#pragma unroll
        for (int j = 0; j < 32; j += 4)
          *reinterpret_cast<float4*>(sb + lane * 32 + (((j >> 2) ^ (lane & 7)) << 2)) = make_float4(o[j], o[j + 1], o[j + 2], o[j + 3]);
        __syncwarp();
        float t[32];
#pragma unroll
        for (int rr = 0; rr < 32; ++rr) t[rr] = sb[rr * 32 + ((((lane >> 2) ^ (rr & 7)) << 2) | (lane & 3))];
        __syncwarp();
        const uint32_t rvalid = rows_here >= 32 ? 0xffffffffu : ((1u << rows_here) - 1u);
        if (col0 + lane < a.n) {
          const int p = sel_append32(S, col0 + lane, t, rvalid, tc_cur, row0);
          if (p >= 0) pending = p;
        }
      }
    }
    // ---- the part of the output that is still stored
    const bool st_direct = st_rows && col0 < S.st_c1 && col0 + 32 > S.st_c0;
    const bool st_mirror = st_rows_t && col0 < S.st_r1 && col0 + 32 > S.st_r0;
    if (st_direct) {  // warp-uniform
      if (lane == 0) ptx::bulk_wait_read<0>();
      __syncwarp();
#pragma unroll
      for (int j = 0; j < 32; j += 4)
        *reinterpret_cast<float4*>(sb + lane * 32 + (((j >> 2) ^ (lane & 7)) << 2)) = make_float4(o[j], o[j + 1], o[j + 2], o[j + 3]);
      ptx::fence_proxy_async();
      __syncwarp();
      if (lane == 0) {
        ptx::tma_store_2d(tmO, ptx::smem_u32(sb), col0, row0);
        ptx::bulk_commit();
      }
    }
    if (st_mirror && row_ok) {  // out[col0 + j, row0 + lane]: one coalesced 128-byte store per mirror row
      float* mp = a.out + (size_t)col0 * a.ldo + my_row;
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (col0 + j < a.n) mp[(size_t)j * a.ldo] = o[j];
    }
    tau_row = tau_next;
    __syncwarp();  // the threshold / norm staging of this chunk has been read by every lane
    // ---- tighten the rows whose list completed a window during this chunk (the warp works on one row at a time); done
    // per chunk so that tiles working on the same rows at the same time see each other's progress early
    uint32_t pend = __ballot_sync(0xffffffffu, pending >= 0 && !(S.debug & 2));
    while (pend) {
      const int src = __ffs(pend) - 1;
      pend &= pend - 1;
      sel_tighten_row(S, __shfl_sync(0xffffffffu, pending, src), lane);
    }
    pending = -1;
  }
  if (boot) {
    // bootstrap tile: every row of this warp leaves with its list length set and a threshold taken from its sketch (its
    // own 128 columns here at least, the sibling warp adds the other 128), then the warp opens the gate
    if (my_row < a.m) S.resv[my_row] = (uint32_t)min(BN, a.n - col_base);
    for (int r = 0; r < rows_here && !(S.debug & 2); ++r) sel_tighten_row(S, row0 + r, lane);
    __threadfence();
    __syncwarp();
    if (lane == 0) atomicAdd(S.gate + gate_r, 1u);
  }
}

template <int KIND, bool SELECT>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_split_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                     const __grid_constant__ CUtensorMap tmO, GemmArgs a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  if ((ptx::smem_u32(smem) & 1023u) != 0) __trap();
  float* epi_buf = reinterpret_cast<float*>(smem + (size_t)STAGES * STAGE_BYTES);
  float* yw_buf = epi_buf + EPI_WARPS * EPI_STAGE_FLOATS;
  uint64_t* bars = reinterpret_cast<uint64_t*>(yw_buf + EPI_YW_FLOATS);
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
      ptx::mbar_init(bar_tempty + 8 * s, EPI_WARPS);  // one arrival per epilogue warp
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
      constexpr uint32_t idesc = ptx::umma_idesc_f16(BM, BN);
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
    const bool vec_ok = a.tma_out != 0;
    const int c_first = (warp - 2) >> 2;  // the two warps of a quadrant interleave the 32-column chunks
    int as = 0;
    uint32_t aphase = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      const int tm = tile % a.tiles_m, tn = tile / a.tiles_m;
      if (SELECT)
        epilogue_select<KIND>(a, &tmO, bar_tfull + 8 * as, aphase, tmem_base + ((uint32_t)(q * 32) << 16) + as * BN,
                              tm * BM + q * 32, tn * BN, c_first, lane, sbuf, yw_buf + (warp - 2) * 64, false, tn == 0, tm, -1,
                              (uint32_t)EPI_WARPS);
      else
        epilogue_slab<KIND>(a, &tmO, bar_tfull + 8 * as, aphase, tmem_base + ((uint32_t)(q * 32) << 16) + as * BN,
                            tm * BM + q * 32, tn * BN, c_first, lane, sbuf, yw_buf + (warp - 2) * 64, vec_ok);
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(bar_tempty + 8 * as);
      as ^= 1;
      if (as == 0) aphase ^= 1;
    }
    if (lane == 0) ptx::bulk_wait<0>();  // every TMA store of this warp has been performed
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, TMEM_COLS);
  }
}


// ------------------------------------------------------------------------------------ tcgen05 kernel, CTA pairs
// cta_group::2: a cluster of two CTAs owns a 256 x 256 output tile.  Each CTA stages its own 128 A rows and HALF of
// the B tile (128 of the 256 rows), so a k-block costs 64 KB of L2->SMEM traffic per SM instead of 96 KB for the same
// number of UMMA cycles; three stages fit.  The even CTA issues every UMMA (M = 256) and multicasts the commits.
constexpr int STAGES2 = 3;
constexpr int STAGE2_BYTES = 4 * A_BYTES;  // Ahi, Alo, Bhi(half), Blo(half): 64 KB
constexpr size_t GEMM2_SMEM = (size_t)STAGES2 * STAGE2_BYTES + EPI_WARPS * EPI_STAGE_FLOATS * 4 + EPI_YW_FLOATS * 4 + 128;
static_assert(GEMM_SMEM <= 232448 && GEMM2_SMEM <= 232448, "shared memory budget (227 KB per CTA)");

template <int KIND, bool SELECT>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(GEMM_THREADS, 1)
gemm_split_tc2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                      const __grid_constant__ CUtensorMap tmO, GemmArgs a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  if ((ptx::smem_u32(smem) & 1023u) != 0) __trap();
  float* epi_buf = reinterpret_cast<float*>(smem + (size_t)STAGES2 * STAGE2_BYTES);
  float* yw_buf = epi_buf + EPI_WARPS * EPI_STAGE_FLOATS;
  uint64_t* bars = reinterpret_cast<uint64_t*>(yw_buf + EPI_YW_FLOATS);
  // bars: [0,S) full, [S,2S) empty, then tmem_full[2], tmem_empty[2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES2 + 4);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;
  const uint32_t rank = ptx::cluster_ctarank();
  const bool leader = rank == 0;
  const uint32_t bar_full = ptx::smem_u32(bars);
  const uint32_t bar_empty = ptx::smem_u32(bars + STAGES2);
  const uint32_t bar_tfull = ptx::smem_u32(bars + 2 * STAGES2);
  const uint32_t bar_tempty = ptx::smem_u32(bars + 2 * STAGES2 + 2);
  const uint32_t smem_base = ptx::smem_u32(smem);

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES2; ++s) {
      ptx::mbar_init(bar_full + 8 * s, 1);   // the leader's arrive.expect_tx; bytes of both CTAs land here
      ptx::mbar_init(bar_empty + 8 * s, 1);  // one multicast commit
    }
    for (int s = 0; s < 2; ++s) {
      ptx::mbar_init(bar_tfull + 8 * s, 1);
      ptx::mbar_init(bar_tempty + 8 * s, 2 * EPI_WARPS);  // epilogue warps of both CTAs arrive on the leader's barrier
    }
    ptx::fence_barrier_init();
    ptx::prefetch_tensormap(&tmA);
    ptx::prefetch_tensormap(&tmB);
  }
  if (warp == 1) {
    ptx::tmem_alloc_2cta(ptx::smem_u32(tmem_slot), TMEM_COLS);
    ptx::tmem_relinquish_2cta();
  }
  ptx::tc_fence_before();
  ptx::cluster_sync();  // barriers of both CTAs initialised before any remote arrive / multicast
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int tiles_m2 = (a.tiles_m + 1) / 2;
  const int total_tiles = a.symmetric ? tiles_m2 * (tiles_m2 + 1) / 2 : tiles_m2 * a.tiles_n;
  const int cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;

  if (warp == 0) {
    // ===== TMA producer (one thread per CTA): own A rows, own half of the B rows
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = cluster_id; tile < total_tiles; tile += num_clusters) {
        int tm, tn;
        pair_tile_coords(tile, tiles_m2, a.symmetric, tm, tn);
        if (KIND == REIDB200_DIST_SQL2 && a.lower) { const int t = tm; tm = tn; tn = t; }
        const int arow = tm * 2 * BM + (int)rank * BM;
        const int brow = tn * BN + (int)rank * (BN / 2);
        for (int kb = 0; kb < a.num_kb; ++kb) {
          ptx::mbar_wait(bar_empty + 8 * stage, phase ^ 1);
          const uint32_t full = bar_full + 8 * stage;
          const uint32_t sb = smem_base + stage * STAGE2_BYTES;
          if (leader) ptx::mbar_arrive_expect_tx(full, 2 * STAGE2_BYTES);
          ptx::tma_load_2d_2cta(sb, &tmA, full, kb * BK, arow);
          ptx::tma_load_2d_2cta(sb + A_BYTES, &tmA, full, a.dp + kb * BK, arow);
          ptx::tma_load_2d_2cta(sb + 2 * A_BYTES, &tmB, full, kb * BK, brow);
          ptx::tma_load_2d_2cta(sb + 3 * A_BYTES, &tmB, full, a.dp + kb * BK, brow);
          if (++stage == STAGES2) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: one thread of the even CTA
    if (leader && lane == 0) {
      constexpr uint32_t idesc = ptx::umma_idesc_f16(2 * BM, BN);
      int stage = 0, as = 0;
      uint32_t phase = 0, aphase = 0;
      for (int tile = cluster_id; tile < total_tiles; tile += num_clusters) {
        // triangle mode: an off-diagonal tile has the roles of its two blocks swapped and issues the two cross terms in
        // the opposite order (same products, same order of accumulation as the upper-triangle tile); a diagonal tile is
        // computed in full exactly as before
        bool swapped = false;
        if (KIND == REIDB200_DIST_SQL2 && !SELECT && a.lower) {
          int tm, tn;
          pair_tile_coords(tile, tiles_m2, a.symmetric, tm, tn);
          swapped = tm != tn;
        }
        ptx::mbar_wait(bar_tempty + 8 * as, aphase ^ 1);
        ptx::tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BN;
        for (int kb = 0; kb < a.num_kb; ++kb) {
          ptx::mbar_wait(bar_full + 8 * stage, phase);
          ptx::tc_fence_after();
          const uint32_t sb = smem_base + stage * STAGE2_BYTES;
#pragma unroll
          for (int kk = 0; kk < BK / 16; ++kk) {
            const uint64_t ahi = ptx::umma_desc_k_sw128(sb + kk * 32);
            const uint64_t alo = ptx::umma_desc_k_sw128(sb + A_BYTES + kk * 32);
            const uint64_t bhi = ptx::umma_desc_k_sw128(sb + 2 * A_BYTES + kk * 32);
            const uint64_t blo = ptx::umma_desc_k_sw128(sb + 3 * A_BYTES + kk * 32);
            if (swapped) {
              ptx::mma_f16_ss_2cta(d_tmem, ahi, blo, idesc, (kb | kk) != 0);
              ptx::mma_f16_ss_2cta(d_tmem, alo, bhi, idesc, 1);
            } else {
              ptx::mma_f16_ss_2cta(d_tmem, alo, bhi, idesc, (kb | kk) != 0);
              ptx::mma_f16_ss_2cta(d_tmem, ahi, blo, idesc, 1);
            }
            ptx::mma_f16_ss_2cta(d_tmem, ahi, bhi, idesc, 1);
          }
          ptx::mma_commit_2cta(bar_empty + 8 * stage, 3);  // frees this stage in BOTH CTAs
          if (++stage == STAGES2) { stage = 0; phase ^= 1; }
        }
        ptx::mma_commit_2cta(bar_tfull + 8 * as, 3);  // accumulator complete, both CTAs
        as ^= 1;
        if (as == 0) aphase ^= 1;
      }
    }
  } else {
    // ===== epilogue warps (both CTAs): own 128 rows of the 256-row tile
    const int q = warp & 3;
    float* sbuf = epi_buf + (warp - 2) * EPI_STAGE_FLOATS;
    const bool vec_ok = a.tma_out != 0;
    const int c_first = (warp - 2) >> 2;  // the two warps of a quadrant interleave the 32-column chunks
    int as = 0;
    uint32_t aphase = 0;
    for (int tile = cluster_id; tile < total_tiles; tile += num_clusters) {
      int tm, tn;
      pair_tile_coords(tile, tiles_m2, a.symmetric, tm, tn);
      if (!SELECT && KIND == REIDB200_DIST_SQL2 && a.lower) { const int t = tm; tm = tn; tn = t; }
      if (SELECT)
        epilogue_select<KIND>(a, &tmO, bar_tfull + 8 * as, aphase, tmem_base + ((uint32_t)(q * 32) << 16) + as * BN,
                              tm * 2 * BM + (int)rank * BM + q * 32, tn * BN, c_first, lane, sbuf, yw_buf + (warp - 2) * 64,
                              a.symmetric && tn != tm, a.symmetric ? tn == tm : tn == 0, tm, (a.symmetric && tn != tm) ? tn : -1,
                              2u * EPI_WARPS);
      else
        epilogue_slab<KIND>(a, &tmO, bar_tfull + 8 * as, aphase, tmem_base + ((uint32_t)(q * 32) << 16) + as * BN,
                            tm * 2 * BM + (int)rank * BM + q * 32, tn * BN, c_first, lane, sbuf, yw_buf + (warp - 2) * 64,
                            vec_ok, a.symmetric && tn != tm);
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive_leader(bar_tempty + 8 * as);
      as ^= 1;
      if (as == 0) aphase ^= 1;
    }
    if (lane == 0) ptx::bulk_wait<0>();  // every TMA store of this warp has been performed
  }

  ptx::tc_fence_before();
  ptx::cluster_sync();  // nobody leaves while the peer may still read this CTA's smem / signal its barriers
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc_2cta(tmem_base, TMEM_COLS);
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

// packed operand [rows, 2*dp] fp16, box = {64, box_rows}, 128-byte swizzle, OOB rows/cols read as zero
int make_operand_map(CUtensorMap* tm, const void* packed, int rows, int dp, int box_rows) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) {
    set_error("cuTensorMapEncodeTiled not available from the driver");
    return REIDB200_ECUDA;
  }
  cuuint64_t gdim[2] = {(cuuint64_t)2 * dp, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)2 * dp * sizeof(__half)};
  cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(packed), gdim, gstride, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (%d) rows=%d dp=%d", (int)r, rows, dp);
    return REIDB200_ECUDA;
  }
  return REIDB200_OK;
}

// output matrix f32[m, n] with row stride ldo (elements): box = 32 x 32, SWIZZLE_128B (128-byte box rows)
static int make_output_map(CUtensorMap* tm, float* out, int m, int n, int64_t ldo) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) {
    set_error("cuTensorMapEncodeTiled not available from the driver");
    return REIDB200_ECUDA;
  }
  cuuint64_t gdim[2] = {(cuuint64_t)n, (cuuint64_t)m};
  cuuint64_t gstride[1] = {(cuuint64_t)ldo * sizeof(float)};
  cuuint32_t box[2] = {32, 32};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, out, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled (output) failed (%d) m=%d n=%d ldo=%lld", (int)r, m, n, (long long)ldo);
    return REIDB200_ECUDA;
  }
  return REIDB200_OK;
}

static int gemm_launch_impl(const void* packedA, const float* xx, const float* isa, int m, const void* packedB, const float* yy,
                            const float* isb, int n, int d, int kind, float* out, int64_t ldo, cudaStream_t st, float* chunk_red,
                            int64_t ld_chunk_red, int chunk_red_max, int blocked, const SelArgs* sel, uint32_t* chunk_red2 = nullptr);

int gemm_split_launch(const void* packedA, const float* xx, const float* isa, int m, const void* packedB, const float* yy,
                      const float* isb, int n, int d, int kind, float* out, int64_t ldo, cudaStream_t st, float* chunk_red,
                      int64_t ld_chunk_red, int chunk_red_max, int blocked) {
  return gemm_launch_impl(packedA, xx, isa, m, packedB, yy, isb, n, d, kind, out, ldo, st, chunk_red, ld_chunk_red, chunk_red_max,
                          blocked, nullptr);
}

int gemm_split_tri_launch(const void* packed, const float* sq, const float* isc, int n, int d, float* out, int64_t ldo,
                          cudaStream_t st, float* chunk_red, uint32_t* chunk_red2, int64_t ld_chunk_red) {
  RB_REQUIRE(chunk_red && chunk_red2 && out, "distmat(triangle): chunk reductions and an output matrix are required");
  return gemm_launch_impl(packed, sq, isc, n, packed, sq, isc, n, d, REIDB200_DIST_SQL2, out, ldo, st, chunk_red, ld_chunk_red, 0, 0,
                          nullptr, chunk_red2);
}

int gemm_split_select_launch(const void* packedA, const float* xx, const float* isa, int m, const void* packedB, const float* yy,
                             const float* isb, int n, int d, int kind, float* out, int64_t ldo, const SelArgs& sel, cudaStream_t st) {
  RB_REQUIRE(sel.tau && sel.resv && sel.list && sel.sketch && sel.gate && sel.cap > 0 && sel.k > 0 && sel.W >= sel.k && sel.W <= 512 &&
             (sel.W & 31) == 0 && sel.window > 0, "distmat(select): bad selection state");
  return gemm_launch_impl(packedA, xx, isa, m, packedB, yy, isb, n, d, kind, out, ldo, st, nullptr, 0, 0, 0, &sel);
}

static int gemm_launch_impl(const void* packedA, const float* xx, const float* isa, int m, const void* packedB, const float* yy,
                            const float* isb, int n, int d, int kind, float* out, int64_t ldo, cudaStream_t st, float* chunk_red,
                            int64_t ld_chunk_red, int chunk_red_max, int blocked, const SelArgs* sel, uint32_t* chunk_red2) {
  if (m == 0 || n == 0) return REIDB200_OK;
  RB_REQUIRE(isa && isb, "distmat: operand scales missing");
  RB_REQUIRE(kind >= 0 && kind <= 3, "distmat: unknown kind %d", kind);
  RB_REQUIRE(kind == REIDB200_DIST_DOT || (xx && yy), "distmat: row norms required for kind %d", kind);
  const int dp = padded_dim(d);
  const bool pair = m > BM && !getenv("REIDB200_GEMM_1CTA");  // CTA pairs (cta_group::2) unless A fits one 128-row tile
  CUtensorMap tmA, tmB;
  int rc = make_operand_map(&tmA, packedA, m, dp, BM);
  if (rc) return rc;
  rc = make_operand_map(&tmB, packedB, n, dp, pair ? BN / 2 : BN);
  if (rc) return rc;
  GemmArgs a;
  a.m = m; a.n = n; a.dp = dp; a.num_kb = dp / BK;
  a.tiles_m = ceil_div(m, BM); a.tiles_n = ceil_div(n, BN);
  a.kind = kind; a.xx = xx; a.yy = yy; a.isa = isa; a.isb = isb; a.out = out; a.ldo = ldo;
  a.cred = chunk_red; a.ldcr = ld_chunk_red; a.cred_max = chunk_red_max;
  memset(&a.sel, 0, sizeof(a.sel));
  const bool want_store = !sel || (sel->st_r1 > sel->st_r0 && sel->st_c1 > sel->st_c0);
  if (sel) {
    a.sel = *sel;
    if (!want_store) { a.sel.st_r0 = a.sel.st_r1 = a.sel.st_c0 = a.sel.st_c1 = 0; }
    RB_REQUIRE(!want_store || out, "distmat(select): a stored block needs an output matrix");
  }
  CUtensorMap tmO;
  memset(&tmO, 0, sizeof(tmO));
  a.tma_out = out != nullptr && (ldo & 3) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0 &&
              (yy == nullptr || (reinterpret_cast<uintptr_t>(yy) & 15) == 0) && (reinterpret_cast<uintptr_t>(isb) & 15) == 0;
  a.blocked = blocked ? ceil_div(n, 32) : 0;
  if (blocked) {
    RB_REQUIRE(a.tma_out, "distmat: the box-major layout needs 16-byte aligned buffers");
    rc = make_output_map(&tmO, out, ceil_div(m, 32) * a.blocked * 32, 32, 32);  // a tall matrix of 128-byte rows
    if (rc) return rc;
  } else if (a.tma_out) {
    rc = make_output_map(&tmO, out, m, n, ldo);
    if (rc) return rc;
  }
  RB_REQUIRE(!sel || !want_store || a.tma_out, "distmat(select): the stored block needs 16-byte aligned rows");
  // self distances: compute the upper triangle of 256 x 256 tiles only and store every off-diagonal tile twice
  a.debug = getenv("REIDB200_GEMM_DEBUG") ? atoi(getenv("REIDB200_GEMM_DEBUG")) : 0;
  a.symmetric = pair && (a.tma_out || sel) && packedA == packedB && m == n && xx == yy && isa == isb && !getenv("REIDB200_GEMM_NOSYM");
  if (a.symmetric && sel) a.symmetric = 2;  // selection: diagonal tiles first (pair_tile_coords)
  a.lower = 0;
  a.cred2 = nullptr;
  if (chunk_red2) {  // triangle mode: the caller's selection relies on it, so it is never silently dropped
    RB_REQUIRE(a.symmetric == 1 && !blocked && !chunk_red_max, "distmat(triangle): needs the symmetric CTA-pair kernel (n > %d, aligned rows)", BM);
    a.lower = 1;
    a.cred2 = chunk_red2;
  }
  void (*k1)(CUtensorMap, CUtensorMap, CUtensorMap, GemmArgs) = nullptr;
  void (*k2)(CUtensorMap, CUtensorMap, CUtensorMap, GemmArgs) = nullptr;
  if (sel) {  // selection in the epilogue: SQL2 (ascending) and DOT (descending) only, separate instantiations so that the
              // storing kernels keep their register allocation
    RB_REQUIRE(kind == REIDB200_DIST_SQL2 || kind == REIDB200_DIST_DOT, "distmat(select): metric must be SQL2 or DOT");
    if (kind == REIDB200_DIST_DOT) { k1 = gemm_split_tc_kernel<REIDB200_DIST_DOT, true>; k2 = gemm_split_tc2_kernel<REIDB200_DIST_DOT, true>; }
    else { k1 = gemm_split_tc_kernel<REIDB200_DIST_SQL2, true>; k2 = gemm_split_tc2_kernel<REIDB200_DIST_SQL2, true>; }
  } else {
    switch (kind) {
      case REIDB200_DIST_EUCLID: k1 = gemm_split_tc_kernel<REIDB200_DIST_EUCLID, false>; k2 = gemm_split_tc2_kernel<REIDB200_DIST_EUCLID, false>; break;
      case REIDB200_DIST_COSINE_HALF: k1 = gemm_split_tc_kernel<REIDB200_DIST_COSINE_HALF, false>; k2 = gemm_split_tc2_kernel<REIDB200_DIST_COSINE_HALF, false>; break;
      case REIDB200_DIST_DOT: k1 = gemm_split_tc_kernel<REIDB200_DIST_DOT, false>; k2 = gemm_split_tc2_kernel<REIDB200_DIST_DOT, false>; break;
      default: k1 = gemm_split_tc_kernel<REIDB200_DIST_SQL2, false>; k2 = gemm_split_tc2_kernel<REIDB200_DIST_SQL2, false>; break;
    }
  }
  static DeviceOnce attr_done[2][2][4];
  if (attr_done[sel ? 1 : 0][pair ? 1 : 0][kind].pending()) {
    RB_CUDA(cudaFuncSetAttribute(pair ? (const void*)k2 : (const void*)k1, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)(pair ? GEMM2_SMEM : GEMM_SMEM)));
  }
  if (pair) {
    const long long tm2 = (a.tiles_m + 1) / 2;
    const long long total2 = a.symmetric ? tm2 * (tm2 + 1) / 2 : tm2 * a.tiles_n;
    const int max_clusters = num_sms() / 2;
    const int clusters = (int)(total2 < max_clusters ? total2 : max_clusters);
    { ProfScope _ps("gemm_split_tc_kernel", st); k2<<<2 * clusters, GEMM_THREADS, GEMM2_SMEM, st>>>(tmA, tmB, tmO, a); }
    RB_LAUNCH_CHECK("gemm_split_tc2_kernel");
    return REIDB200_OK;
  }
  const long long total = (long long)a.tiles_m * a.tiles_n;
  const int grid = (int)(total < num_sms() ? total : num_sms());
  { ProfScope _ps("gemm_split_tc_kernel", st); k1<<<grid, GEMM_THREADS, GEMM_SMEM, st>>>(tmA, tmB, tmO, a); }
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
                                  float* iscale, void* stream) {
  if (rows == 0) return REIDB200_OK;
  RB_REQUIRE(x && packed && iscale && rows > 0 && d > 0 && ldx >= d, "pack_rows: bad argument");
  RB_REQUIRE((reinterpret_cast<uintptr_t>(packed) & 15) == 0, "pack_rows: packed buffer must be 16-byte aligned");
  return pack_rows_launch(x, rows, ldx, d, packed, sqnorm, iscale, (cudaStream_t)stream);
}

extern "C" size_t reidb200_distmat_workspace_bytes(int m, int n, int d) {
  if (m <= 0 || n <= 0 || d <= 0) return 0;
  CarveSize c;
  c.take<char>(packed_bytes(m, d));
  c.take<float>(2 * align_up((size_t)m, 4));
  c.take<char>(packed_bytes(n, d));
  c.take<float>(2 * align_up((size_t)n, 4));
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
  float* xx = c.take<float>(2 * align_up((size_t)m, 4));
  void* pb = c.take<char>(packed_bytes(n, d));
  float* yy = c.take<float>(2 * align_up((size_t)n, 4));
  float* isa = xx + align_up((size_t)m, 4);  // 16-byte aligned halves: the epilogue loads float4s of yy and isb
  float* isb = yy + align_up((size_t)n, 4);
  int rc = pack_rows_launch(x, m, ldx, d, pa, xx, isa, st);
  if (rc) return rc;
  if (x == y && m == n && ldx == ldy) {
    pb = pa;
    yy = xx;
    isb = isa;
  } else {
    rc = pack_rows_launch(y, n, ldy, d, pb, yy, isb, st);
    if (rc) return rc;
  }
  return gemm_split_launch(pa, xx, isa, m, pb, yy, isb, n, d, kind, out, ldo, st);
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
