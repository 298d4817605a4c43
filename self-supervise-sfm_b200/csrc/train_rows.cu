// HBM-bound row kernels of the TRAINING step (forward with saved intermediates + backward of one Block,
// sailrecon/layers/block.py:86-112 under torch.utils.checkpoint as in models/aggregator.py:654-657,726-729,758-761):
//   sr_transpose_bf16      [R, C] -> [C, R_pad]  (operands of the dgrad / wgrad GEMMs, which run on sr_gemm_bf16)
//   sr_colsum              out[c] += sum_r a[r, c] * b[r, c]            (bias and LayerScale-gamma gradients)
//   sr_scale_cast_bf16     out = bf16(x * gamma)                         (gradient entering a LayerScale'd Linear)
//   sr_ls_residual_f32     out = resid + gamma * z                       (block.py:110-111 with z kept for d gamma)
//   sr_gelu_fwd / _bwd     g = gelu_erf(u);  du = dg * gelu'(u)          (mlp.py:36)
//   sr_layernorm_bwd       dx += LN'(dy), dw += ..., db += ...           (block.py:87-91 norm1 / norm2)
//   sr_qknorm_rope_fwd/bwd per-head LayerNorm(64) + 2D RoPE on the q, k thirds of a [M, 3C] bf16 buffer
//                          (attention.py:78-82) and its gradient, in place
//   sr_adam_step_f32       torch.optim.Adam update (train/train_imc.py:475,381) on one flat fp32 parameter bucket
#include "common.cuh"
#include "sr_internal.h"

namespace sr {
extern int g_launch_count;

__device__ __forceinline__ float bf16_lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t w) { return __uint_as_float(w & 0xffff0000u); }

// ------------------------------------------------------------------------------------------------ transpose
// 64 x 64 tiles through shared memory; reads and writes are 128-byte rows of bf16.
__global__ void __launch_bounds__(256)
transpose_bf16_kernel(const __nv_bfloat16* __restrict__ src, int64_t lds, __nv_bfloat16* __restrict__ dst, int64_t ldd,
                      int R, int C, int R_pad) {
  __shared__ __nv_bfloat16 tile[64][66];
  const int r0 = blockIdx.x * 64, c0 = blockIdx.y * 64;
  const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;          // 64 x 4
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const int r = r0 + ty + 4 * i, c = c0 + tx;
    tile[ty + 4 * i][tx] = (r < R && c < C) ? src[static_cast<int64_t>(r) * lds + c] : __float2bfloat16(0.f);
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const int c = c0 + ty + 4 * i, r = r0 + tx;
    if (c < C && r < R_pad) dst[static_cast<int64_t>(c) * ldd + r] = tile[tx][ty + 4 * i];
  }
}

// Vector form (16-byte global accesses both ways): 64 x 64 tile, thread t loads chunks (row t/8 [+32], 8 columns), the
// write side gives every group of 8 lanes one output row (128 contiguous bytes, rows r0 + 8k .. + 7 of source column c).
// The 8 values a thread gathers for its chunk are 8 rows of ONE source column, so the column sums of src (the bias
// gradient of the Linear whose dY this is) come for free: 3 shuffles + one atomic per (column, tile).
__global__ void __launch_bounds__(256)
transpose_bf16_vec_kernel(const __nv_bfloat16* __restrict__ src, int64_t lds, __nv_bfloat16* __restrict__ dst, int64_t ldd,
                          int R, int C, int R_pad, float* __restrict__ colsum) {
  constexpr int kTilesPerCta = 4;                                  // 256 source rows per CTA: one atomic per column and CTA
  __shared__ __align__(16) uint32_t tile[2][64 * 33];              // 64 rows x 64 bf16, row stride 33 words, double-buffered
  const int c0 = blockIdx.y * 64;
  const int t = threadIdx.x;
  const int warp = t >> 5, lane = t & 31, k = lane & 7;
  float sums[2] = {0.f, 0.f};
#pragma unroll 1
  for (int it = 0; it < kTilesPerCta; ++it) {
    const int r0 = (blockIdx.x * kTilesPerCta + it) * 64;
    if (r0 >= R_pad) break;
    uint32_t* tl = tile[it & 1];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int r = (t >> 3) + 32 * h, c8 = t & 7;
      uint4 v = make_uint4(0, 0, 0, 0);
      if (r0 + r < R && c0 + c8 * 8 < C) v = *reinterpret_cast<const uint4*>(src + static_cast<int64_t>(r0 + r) * lds + c0 + c8 * 8);
      uint32_t* w = tl + r * 33 + c8 * 4;
      w[0] = v.x; w[1] = v.y; w[2] = v.z; w[3] = v.w;
    }
    __syncthreads();                                               // (the other buffer is free: its readers passed the previous barrier)
    const __nv_bfloat16* tb = reinterpret_cast<const __nv_bfloat16*>(tl);
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int c = warp * 8 + i * 4 + (lane >> 3);
      uint32_t o[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const __nv_bfloat16 a = tb[(8 * k + 2 * j) * 66 + c], b = tb[(8 * k + 2 * j + 1) * 66 + c];
        o[j] = static_cast<uint32_t>(__bfloat16_as_ushort(a)) | (static_cast<uint32_t>(__bfloat16_as_ushort(b)) << 16);
        sums[i] += __bfloat162float(a) + __bfloat162float(b);
      }
      if (c0 + c < C) *reinterpret_cast<uint4*>(dst + static_cast<int64_t>(c0 + c) * ldd + r0 + 8 * k) = make_uint4(o[0], o[1], o[2], o[3]);
    }
  }
  if (colsum != nullptr) {
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int c = warp * 8 + i * 4 + (lane >> 3);
      float sum = sums[i];
      sum += __shfl_xor_sync(0xffffffffu, sum, 1);
      sum += __shfl_xor_sync(0xffffffffu, sum, 2);
      sum += __shfl_xor_sync(0xffffffffu, sum, 4);
      if (k == 0 && c0 + c < C) atomicAdd(colsum + c0 + c, sum);
    }
  }
}

// ------------------------------------------------------------------------------------------------ column sums
// One thread per column, 128 columns x 128 rows per block, atomics into out (zeroed by the launcher unless accumulate).
template <bool A_BF16>
__global__ void __launch_bounds__(128)
colsum_kernel(const void* __restrict__ a_, int64_t lda, const __nv_bfloat16* __restrict__ b, int64_t ldb, float* __restrict__ out,
              int R, int C) {
  const int c = blockIdx.x * 128 + threadIdx.x;
  if (c >= C) return;
  const int r0 = blockIdx.y * 128, r1 = min(R, r0 + 128);
  float acc = 0.f;
  for (int r = r0; r < r1; ++r) {
    const float av = A_BF16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(a_)[static_cast<int64_t>(r) * lda + c])
                            : reinterpret_cast<const float*>(a_)[static_cast<int64_t>(r) * lda + c];
    acc += b ? av * __bfloat162float(b[static_cast<int64_t>(r) * ldb + c]) : av;
  }
  atomicAdd(out + c, acc);
}

// Bias gradient db = colsum(dY) on bf16 dY (round 2: no longer fused into a transpose, the wgrad GEMM reads dY in place): 256
// columns x 256 rows per block; a thread owns 8 adjacent columns (16-byte loads: a warp reads 512 contiguous bytes per row),
// the 8 warps take rows r0 + w, r0 + w + 8, ... (4 rows in flight), partials meet in shared memory, one atomic per column.
__global__ void __launch_bounds__(256)
colsum_bf16_kernel(const __nv_bfloat16* __restrict__ a, int64_t lda, float* __restrict__ out, int R, int C) {
  __shared__ float part[8][256 + 8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c0 = blockIdx.x * 256 + lane * 8;
  const int r0 = blockIdx.y * 256, r1 = min(R, r0 + 256);
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (c0 < C) {
    for (int r = r0 + warp; r < r1; r += 32) {
      uint4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int rr = r + u * 8;
        v[u] = rr < r1 ? __ldg(reinterpret_cast<const uint4*>(a + static_cast<int64_t>(rr) * lda + c0)) : make_uint4(0u, 0u, 0u, 0u);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        acc[0] += bf16_lo(v[u].x); acc[1] += bf16_hi(v[u].x);
        acc[2] += bf16_lo(v[u].y); acc[3] += bf16_hi(v[u].y);
        acc[4] += bf16_lo(v[u].z); acc[5] += bf16_hi(v[u].z);
        acc[6] += bf16_lo(v[u].w); acc[7] += bf16_hi(v[u].w);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) part[warp][lane * 8 + i] = acc[i];
  __syncthreads();
  const int c = blockIdx.x * 256 + threadIdx.x;
  if (c < C) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += part[w][threadIdx.x];
    atomicAdd(out + c, s);
  }
}

// LayerScale gamma gradient: out[c] += sum_r a[r, c] * b[r, c] with a fp32 (the incoming gradient) and b bf16 (the saved Linear
// output); same 256 x 256 block shape as colsum_bf16_kernel, 2 rows in flight per warp (48 B per lane and row).
__global__ void __launch_bounds__(256)
colsum_f32xbf16_kernel(const float* __restrict__ a, int64_t lda, const __nv_bfloat16* __restrict__ b, int64_t ldb,
                       float* __restrict__ out, int R, int C) {
  __shared__ float part[8][256 + 8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c0 = blockIdx.x * 256 + lane * 8;
  const int r0 = blockIdx.y * 256, r1 = min(R, r0 + 256);
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (c0 < C) {
    for (int r = r0 + warp; r < r1; r += 16) {
      float4 x0[2], x1[2];
      uint4 v[2];
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int rr = r + u * 8;
        const bool ok = rr < r1;
        const float4* pa = reinterpret_cast<const float4*>(a + static_cast<int64_t>(rr) * lda + c0);
        x0[u] = ok ? __ldg(pa) : make_float4(0.f, 0.f, 0.f, 0.f);
        x1[u] = ok ? __ldg(pa + 1) : make_float4(0.f, 0.f, 0.f, 0.f);
        v[u] = ok ? __ldg(reinterpret_cast<const uint4*>(b + static_cast<int64_t>(rr) * ldb + c0)) : make_uint4(0u, 0u, 0u, 0u);
      }
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        acc[0] = fmaf(x0[u].x, bf16_lo(v[u].x), acc[0]); acc[1] = fmaf(x0[u].y, bf16_hi(v[u].x), acc[1]);
        acc[2] = fmaf(x0[u].z, bf16_lo(v[u].y), acc[2]); acc[3] = fmaf(x0[u].w, bf16_hi(v[u].y), acc[3]);
        acc[4] = fmaf(x1[u].x, bf16_lo(v[u].z), acc[4]); acc[5] = fmaf(x1[u].y, bf16_hi(v[u].z), acc[5]);
        acc[6] = fmaf(x1[u].z, bf16_lo(v[u].w), acc[6]); acc[7] = fmaf(x1[u].w, bf16_hi(v[u].w), acc[7]);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) part[warp][lane * 8 + i] = acc[i];
  __syncthreads();
  const int c = blockIdx.x * 256 + threadIdx.x;
  if (c < C) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += part[w][threadIdx.x];
    atomicAdd(out + c, s);
  }
}

// ------------------------------------------------------------------------------------------------ elementwise
__global__ void __launch_bounds__(256)
scale_cast_kernel(const float* __restrict__ x, int64_t ldx, const float* __restrict__ gamma, __nv_bfloat16* __restrict__ out,
                  int64_t ldo, int64_t rows, int C) {
  const int64_t e = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) * 4;
  if (e >= rows * C) return;
  const int64_t r = e / C;
  const int c = static_cast<int>(e % C);
  const float4 v = *reinterpret_cast<const float4*>(x + r * ldx + c);
  const float4 g = gamma ? __ldg(reinterpret_cast<const float4*>(gamma + c)) : make_float4(1.f, 1.f, 1.f, 1.f);
  uint2 w;
  w.x = pack_bf16x2(v.x * g.x, v.y * g.y);
  w.y = pack_bf16x2(v.z * g.z, v.w * g.w);
  *reinterpret_cast<uint2*>(out + r * ldo + c) = w;
}

__global__ void __launch_bounds__(256)
ls_residual_kernel(const float* __restrict__ resid, const __nv_bfloat16* __restrict__ z, const float* __restrict__ gamma,
                   float* __restrict__ out, int64_t rows, int C) {
  const int64_t e = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) * 4;
  if (e >= rows * C) return;
  const int c = static_cast<int>(e % C);
  const float4 r = *reinterpret_cast<const float4*>(resid + e);
  const uint2 zz = *reinterpret_cast<const uint2*>(z + e);
  const float4 g = gamma ? __ldg(reinterpret_cast<const float4*>(gamma + c)) : make_float4(1.f, 1.f, 1.f, 1.f);
  float4 y;
  y.x = r.x + g.x * bf16_lo(zz.x);
  y.y = r.y + g.y * bf16_hi(zz.x);
  y.z = r.z + g.z * bf16_lo(zz.y);
  y.w = r.w + g.w * bf16_hi(zz.y);
  *reinterpret_cast<float4*>(out + e) = y;
}

// erf by Abramowitz-Stegun 7.1.26 (|error| <= 1.5e-7, the form the GEMM's fused GELU epilogue uses): one rcp, one ex2 and
// 8 FMAs instead of the ~27 branchy instructions of erff(); the same exponential exp(-x^2 / 2) is the Gaussian density
// the derivative needs.  erf_abs = erf(|x| / sqrt 2), e = exp(-x^2 / 2).
__device__ __forceinline__ void erf_gauss(float x, float& erf_abs, float& e) {
  const float z = fabsf(x) * 0.70710678118654752440f;
  float t;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(0.3275911f, z, 1.0f)));
  float poly = fmaf(1.061405429f, t, -1.453152027f);
  poly = fmaf(poly, t, 1.421413741f);
  poly = fmaf(poly, t, -0.284496736f);
  poly = fmaf(poly, t, 0.254829592f);
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(z * z * -1.4426950408889634f));
  erf_abs = fmaf(-poly * t, e, 1.0f);
}

// mode 0: out = bf16(gelu_erf(u));  mode 1: out = bf16(dg * gelu'(u)),  gelu'(u) = Phi(u) + u * phi(u)
template <int kMode>
__global__ void __launch_bounds__(256)
gelu_kernel(const __nv_bfloat16* __restrict__ u, const __nv_bfloat16* __restrict__ dg, __nv_bfloat16* __restrict__ out, int64_t n) {
  const int64_t e = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) * 8;
  if (e >= n) return;
  const uint4 uu = *reinterpret_cast<const uint4*>(u + e);
  uint4 gg = make_uint4(0, 0, 0, 0);
  if (kMode == 1) gg = *reinterpret_cast<const uint4*>(dg + e);
  const uint32_t uw[4] = {uu.x, uu.y, uu.z, uu.w}, gw[4] = {gg.x, gg.y, gg.z, gg.w};
  uint32_t ow[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float r[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const float x = h ? bf16_hi(uw[i]) : bf16_lo(uw[i]);
      float erf_abs, e;
      erf_gauss(x, erf_abs, e);
      if (kMode == 0) {
        const float hx = 0.5f * x;
        r[h] = fmaf(fabsf(hx), erf_abs, hx);                 // 0.5 x (1 + sign(x) erf_abs)
      } else {
        const float cdf = 0.5f + copysignf(0.5f * erf_abs, x);
        r[h] = (h ? bf16_hi(gw[i]) : bf16_lo(gw[i])) * fmaf(x, 0.39894228040143267794f * e, cdf);
      }
    }
    ow[i] = pack_bf16x2(r[0], r[1]);
  }
  *reinterpret_cast<uint4*>(out + e) = make_uint4(ow[0], ow[1], ow[2], ow[3]);
}

// LayerScale backward in ONE pass (round 2, late): reads the incoming gradient dy (fp32) and the saved Linear output z (bf16)
// once and produces everything block.py:110-111 needs from them —
//   dz[r, c]    = bf16(dy * gamma[c])          (the gradient entering the Linear under the LayerScale)
//   dgamma[c]  += sum_r dy * z                 (layer_scale.py:22)
//   dbias[c]   += sum_r float(dz)              (the Linear's bias gradient: column sums of the ROUNDED dz, as colsum(dz) gives)
// 8 B per element instead of the 14 B of colsum(dy, z) + scale_cast + colsum(dz).  Same 256 x 256 block shape as
// colsum_f32xbf16_kernel; two atomics per column and block.
__global__ void __launch_bounds__(256)
ls_bwd_kernel(const float* __restrict__ dy, int64_t lddy, const __nv_bfloat16* __restrict__ z, int64_t ldz,
              const float* __restrict__ gamma, __nv_bfloat16* __restrict__ dz, int64_t lddz, float* __restrict__ dgamma,
              float* __restrict__ dbias, int R, int C) {
  __shared__ float part[2][8][256 + 8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c0 = blockIdx.x * 256 + lane * 8;
  const int r0 = blockIdx.y * 256, r1 = min(R, r0 + 256);
  float ag[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, ab[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (c0 < C) {
    const float4 g0 = __ldg(reinterpret_cast<const float4*>(gamma + c0)), g1 = __ldg(reinterpret_cast<const float4*>(gamma + c0) + 1);
    const float gm[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
    for (int r = r0 + warp; r < r1; r += 16) {
      float4 x0[2], x1[2];
      uint4 v[2];
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int rr = r + u * 8;
        const bool ok = rr < r1;
        const float4* pa = reinterpret_cast<const float4*>(dy + static_cast<int64_t>(rr) * lddy + c0);
        x0[u] = ok ? __ldg(pa) : make_float4(0.f, 0.f, 0.f, 0.f);
        x1[u] = ok ? __ldg(pa + 1) : make_float4(0.f, 0.f, 0.f, 0.f);
        v[u] = ok ? __ldg(reinterpret_cast<const uint4*>(z + static_cast<int64_t>(rr) * ldz + c0)) : make_uint4(0u, 0u, 0u, 0u);
      }
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int rr = r + u * 8;
        if (rr >= r1) continue;
        const float d[8] = {x0[u].x, x0[u].y, x0[u].z, x0[u].w, x1[u].x, x1[u].y, x1[u].z, x1[u].w};
        const float zz[8] = {bf16_lo(v[u].x), bf16_hi(v[u].x), bf16_lo(v[u].y), bf16_hi(v[u].y),
                             bf16_lo(v[u].z), bf16_hi(v[u].z), bf16_lo(v[u].w), bf16_hi(v[u].w)};
        uint32_t o[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          ag[2 * i] = fmaf(d[2 * i], zz[2 * i], ag[2 * i]);
          ag[2 * i + 1] = fmaf(d[2 * i + 1], zz[2 * i + 1], ag[2 * i + 1]);
          o[i] = pack_bf16x2(d[2 * i] * gm[2 * i], d[2 * i + 1] * gm[2 * i + 1]);
          ab[2 * i] += bf16_lo(o[i]);
          ab[2 * i + 1] += bf16_hi(o[i]);
        }
        *reinterpret_cast<uint4*>(dz + static_cast<int64_t>(rr) * lddz + c0) = make_uint4(o[0], o[1], o[2], o[3]);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) { part[0][warp][lane * 8 + i] = ag[i]; part[1][warp][lane * 8 + i] = ab[i]; }
  __syncthreads();
  const int c = blockIdx.x * 256 + threadIdx.x;
  if (c < C) {
    float sg = 0.f, sb = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) { sg += part[0][w][threadIdx.x]; sb += part[1][w][threadIdx.x]; }
    atomicAdd(dgamma + c, sg);
    if (dbias) atomicAdd(dbias + c, sb);
  }
}

// GELU backward with the fc1 bias gradient in the same pass: du = bf16(dg * gelu'(u)) (in place over dg allowed) and
// dbias[c] += sum_r float(du) — saves the separate colsum(du) read of the [T, 4096] gradient.
__global__ void __launch_bounds__(256)
gelu_bwd_colsum_kernel(const __nv_bfloat16* __restrict__ u, int64_t ldu, const __nv_bfloat16* dg, int64_t lddg, __nv_bfloat16* du,
                       int64_t lddu, float* __restrict__ dbias, int R, int C) {
  __shared__ float part[8][256 + 8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c0 = blockIdx.x * 256 + lane * 8;
  const int r0 = blockIdx.y * 256, r1 = min(R, r0 + 256);
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (c0 < C) {
    for (int r = r0 + warp; r < r1; r += 16) {
      uint4 uu[2], gg[2];
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        const int rr = r + k * 8;
        const bool ok = rr < r1;
        uu[k] = ok ? __ldg(reinterpret_cast<const uint4*>(u + static_cast<int64_t>(rr) * ldu + c0)) : make_uint4(0u, 0u, 0u, 0u);
        gg[k] = ok ? *reinterpret_cast<const uint4*>(dg + static_cast<int64_t>(rr) * lddg + c0) : make_uint4(0u, 0u, 0u, 0u);
      }
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        const int rr = r + k * 8;
        if (rr >= r1) continue;
        const uint32_t uw[4] = {uu[k].x, uu[k].y, uu[k].z, uu[k].w}, gw[4] = {gg[k].x, gg[k].y, gg[k].z, gg[k].w};
        uint32_t ow[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          float rv[2];
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            const float x = h ? bf16_hi(uw[i]) : bf16_lo(uw[i]);
            float erf_abs, e;
            erf_gauss(x, erf_abs, e);
            const float cdf = 0.5f + copysignf(0.5f * erf_abs, x);
            rv[h] = (h ? bf16_hi(gw[i]) : bf16_lo(gw[i])) * fmaf(x, 0.39894228040143267794f * e, cdf);
          }
          ow[i] = pack_bf16x2(rv[0], rv[1]);
          acc[2 * i] += bf16_lo(ow[i]);
          acc[2 * i + 1] += bf16_hi(ow[i]);
        }
        *reinterpret_cast<uint4*>(du + static_cast<int64_t>(rr) * lddu + c0) = make_uint4(ow[0], ow[1], ow[2], ow[3]);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) part[warp][lane * 8 + i] = acc[i];
  __syncthreads();
  const int c = blockIdx.x * 256 + threadIdx.x;
  if (c < C) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += part[w][threadIdx.x];
    atomicAdd(dbias + c, s);
  }
}


// ------------------------------------------------------------------------------------------------ LayerNorm backward
// One warp per row, lane holds columns {128 i + 4 lane .. +3}, i < NCH (C = 128 NCH).  dx = dres + rstd * (g - mean(g) -
// xhat * mean(g * xhat)) with g = dy * w.  The per-column dw / db partials live in shared memory (each lane owns its own
// 4 NCH addresses of its warp's slice, so no synchronisation inside the row loop) instead of 2 x 4 NCH registers: 4 warps
// per CTA at <= 128 registers -> 4 CTAs = 16 warps per SM (the register-resident version ran 8); one atomic per column
// and CTA at the end.
template <int NCH>
__global__ void __launch_bounds__(128, 4)
layernorm_bwd_kernel(const float* __restrict__ x, int64_t ldx, const float* __restrict__ w, const __nv_bfloat16* __restrict__ dy,
                     int64_t lddy, const float* dres, int64_t lddres, float* dx, int64_t lddx, float* __restrict__ dw,
                     float* __restrict__ db, int rows, int rows_per_block, float eps) {
  constexpr int C = NCH * 128;
  __shared__ __align__(16) float red_w[4][C];
  __shared__ __align__(16) float red_b[4][C];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int i = 0; i < NCH; ++i) {
    *reinterpret_cast<float4*>(&red_w[warp][i * 128 + lane * 4]) = make_float4(0.f, 0.f, 0.f, 0.f);
    *reinterpret_cast<float4*>(&red_b[warp][i * 128 + lane * 4]) = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  const int r_begin = blockIdx.x * rows_per_block, r_end = min(rows, r_begin + rows_per_block);
  for (int r = r_begin + warp; r < r_end; r += 4) {
    float xv[NCH * 4], gv[NCH * 4];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < NCH; ++i) {
      const float4 v = *reinterpret_cast<const float4*>(x + static_cast<int64_t>(r) * ldx + i * 128 + lane * 4);
      xv[4 * i] = v.x; xv[4 * i + 1] = v.y; xv[4 * i + 2] = v.z; xv[4 * i + 3] = v.w;
      s += v.x + v.y + v.z + v.w;
    }
    const float mean = warp_sum(s) * (1.0f / C);
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < NCH * 4; ++i) { xv[i] -= mean; q += xv[i] * xv[i]; }
    const float rstd = rsqrtf(warp_sum(q) * (1.0f / C) + eps);
    float sg = 0.f, sgx = 0.f;
#pragma unroll
    for (int i = 0; i < NCH; ++i) {
      const uint2 d = *reinterpret_cast<const uint2*>(dy + static_cast<int64_t>(r) * lddy + i * 128 + lane * 4);
      const float4 wv = w ? __ldg(reinterpret_cast<const float4*>(w + i * 128 + lane * 4)) : make_float4(1.f, 1.f, 1.f, 1.f);
      const float dyv[4] = {bf16_lo(d.x), bf16_hi(d.x), bf16_lo(d.y), bf16_hi(d.y)};
      const float wq[4] = {wv.x, wv.y, wv.z, wv.w};
      float4 aw = *reinterpret_cast<float4*>(&red_w[warp][i * 128 + lane * 4]);
      float4 ab = *reinterpret_cast<float4*>(&red_b[warp][i * 128 + lane * 4]);
      float* awp = &aw.x;
      float* abp = &ab.x;
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const float xh = xv[4 * i + u] * rstd;
        xv[4 * i + u] = xh;
        awp[u] += dyv[u] * xh;
        abp[u] += dyv[u];
        const float g = dyv[u] * wq[u];
        gv[4 * i + u] = g;
        sg += g;
        sgx += g * xh;
      }
      *reinterpret_cast<float4*>(&red_w[warp][i * 128 + lane * 4]) = aw;
      *reinterpret_cast<float4*>(&red_b[warp][i * 128 + lane * 4]) = ab;
    }
    const float mg = warp_sum(sg) * (1.0f / C), mgx = warp_sum(sgx) * (1.0f / C);
#pragma unroll
    for (int i = 0; i < NCH; ++i) {
      float* o = dx + static_cast<int64_t>(r) * lddx + i * 128 + lane * 4;
      float4 v = *reinterpret_cast<const float4*>(dres + static_cast<int64_t>(r) * lddres + i * 128 + lane * 4);   // may alias o
      v.x += rstd * (gv[4 * i] - mg - xv[4 * i] * mgx);
      v.y += rstd * (gv[4 * i + 1] - mg - xv[4 * i + 1] * mgx);
      v.z += rstd * (gv[4 * i + 2] - mg - xv[4 * i + 2] * mgx);
      v.w += rstd * (gv[4 * i + 3] - mg - xv[4 * i + 3] * mgx);
      *reinterpret_cast<float4*>(o) = v;
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += 128) {
    if (dw != nullptr) atomicAdd(dw + c, red_w[0][c] + red_w[1][c] + red_w[2][c] + red_w[3][c]);
    if (db != nullptr) atomicAdd(db + c, red_b[0][c] + red_b[1][c] + red_b[2][c] + red_b[3][c]);
  }
}

// ------------------------------------------------------------------------------------------------ QK-norm + RoPE
// One warp = FOUR adjacent heads of a row (512 contiguous bytes): lane = 8 g + q holds features 8 q .. 8 q + 7 of head g as
// one 16-byte load.  LayerNorm(64) reduces over the 8 lanes of a head (3 shuffles); the rotate-half partner of feature j
// inside a half is j ^ 16 -> lane ^ 2.  Two rows per iteration, the next two rows' loads are issued before the current
// ones are computed.  blockIdx.x = group of 4 head slots in [0, 3H / 4): q heads, k heads, v heads (H % 4 == 0).
struct QkRopeParams {
  const float* qn_w; const float* qn_b; const float* kn_w; const float* kn_b;
  float eps;
  int qk_norm, rope;
  const float* rope_cos; const float* rope_sin;     // [max_pos, 16]
  const int32_t* pos;                                // [M, 2] (y, x)
  int H;
  int64_t M;
  int rows_per_warp;
};

__device__ __forceinline__ float sum8(float v) {      // over the 8 lanes of one head
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  v += __shfl_xor_sync(0xffffffffu, v, 2);
  v += __shfl_xor_sync(0xffffffffu, v, 4);
  return v;
}
__device__ __forceinline__ void unpack8(const uint4& v, float* f) {
  f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
  f[4] = bf16_lo(v.z); f[5] = bf16_hi(v.z); f[6] = bf16_lo(v.w); f[7] = bf16_hi(v.w);
}

template <bool BWD>
__global__ void __launch_bounds__(256, 2)
qknorm_rope_kernel(const __nv_bfloat16* __restrict__ raw, int64_t ldr, __nv_bfloat16* io, int64_t ldio,
                   float* __restrict__ dqn_w, float* __restrict__ dqn_b, float* __restrict__ dkn_w, float* __restrict__ dkn_b,
                   const QkRopeParams p) {
  constexpr int U = 2;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int which = (blockIdx.x * 4) / p.H;           // 0 q, 1 k, 2 v
  const int col = blockIdx.x * 256 + lane * 8;
  const int q = lane & 7;
  const int64_t r_begin = (static_cast<int64_t>(blockIdx.y) * 8 + warp) * p.rows_per_warp;
  const int64_t r_end = min(p.M, r_begin + p.rows_per_warp);
  if (which == 2) {
    if (!BWD) {
      for (int64_t r = r_begin; r < r_end; r += 4) {
        uint4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (r + u < r_end) v[u] = *reinterpret_cast<const uint4*>(raw + (r + u) * ldr + col);
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (r + u < r_end) *reinterpret_cast<uint4*>(io + (r + u) * ldio + col) = v[u];
      }
    }
    return;
  }
  const float* nw = which == 0 ? p.qn_w : p.kn_w;
  const float* nb = which == 0 ? p.qn_b : p.kn_b;
  float w[8], bb[8], aw[8], ab[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    w[e] = p.qk_norm ? nw[8 * q + e] : 1.f;
    bb[e] = p.qk_norm ? nb[8 * q + e] : 0.f;
    aw[e] = ab[e] = 0.f;
  }
  const int f = 8 * (q & 1);                          // table column of the lane's first feature (j & 15)
  const bool is_a = (q & 2) == 0;                     // first member of the (i, i + 16) pair
  const bool x_half = q >= 4;
  const uint4 zero4 = make_uint4(0, 0, 0, 0);
  uint4 nx[U], nd[U];                                 // prefetched rows
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const bool ok = r_begin + u < r_end;
    nx[u] = ok ? *reinterpret_cast<const uint4*>(raw + (r_begin + u) * ldr + col) : zero4;
    nd[u] = (BWD && ok) ? *reinterpret_cast<const uint4*>(io + (r_begin + u) * ldio + col) : zero4;
  }
  for (int64_t r = r_begin; r < r_end; r += U) {
    uint4 cx[U], cd[U];
    float cs[U][8], sn[U][8];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      cx[u] = nx[u];
      cd[u] = nd[u];
      const bool ok = r + u < r_end;
#pragma unroll
      for (int e = 0; e < 8; ++e) { cs[u][e] = 1.f; sn[u][e] = 0.f; }
      if (p.rope && ok) {
        const int2 yx = __ldg(reinterpret_cast<const int2*>(p.pos) + (r + u));
        const int64_t t = static_cast<int64_t>(x_half ? yx.y : yx.x) * 16 + f;
        const float4 c0 = __ldg(reinterpret_cast<const float4*>(p.rope_cos + t)), c1 = __ldg(reinterpret_cast<const float4*>(p.rope_cos + t + 4));
        const float4 s0 = __ldg(reinterpret_cast<const float4*>(p.rope_sin + t)), s1 = __ldg(reinterpret_cast<const float4*>(p.rope_sin + t + 4));
        cs[u][0] = c0.x; cs[u][1] = c0.y; cs[u][2] = c0.z; cs[u][3] = c0.w; cs[u][4] = c1.x; cs[u][5] = c1.y; cs[u][6] = c1.z; cs[u][7] = c1.w;
        sn[u][0] = s0.x; sn[u][1] = s0.y; sn[u][2] = s0.z; sn[u][3] = s0.w; sn[u][4] = s1.x; sn[u][5] = s1.y; sn[u][6] = s1.z; sn[u][7] = s1.w;
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {                     // next iteration's rows: in flight under the arithmetic below
      const bool ok = r + U + u < r_end;
      nx[u] = ok ? *reinterpret_cast<const uint4*>(raw + (r + U + u) * ldr + col) : zero4;
      nd[u] = (BWD && ok) ? *reinterpret_cast<const uint4*>(io + (r + U + u) * ldio + col) : zero4;
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      float x[8];
      unpack8(cx[u], x);
      float rstd = 1.f;
      if (p.qk_norm) {
        float sm = 0.f;
#pragma unroll
        for (int e = 0; e < 8; ++e) sm += x[e];
        const float mean = sum8(sm) * (1.0f / 64.0f);
        float sq = 0.f;
#pragma unroll
        for (int e = 0; e < 8; ++e) { x[e] -= mean; sq += x[e] * x[e]; }
        rstd = rsqrtf(sum8(sq) * (1.0f / 64.0f) + p.eps);
#pragma unroll
        for (int e = 0; e < 8; ++e) x[e] *= rstd;      // xhat
      }
      float o[8];
      if (!BWD) {
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float y = p.qk_norm ? x[e] * w[e] + bb[e] : x[e];
          const float oth = __shfl_xor_sync(0xffffffffu, y, 2);
          // pair (a, b) = (i, i + 16): a' = a c - b s ; b' = b c + a s
          o[e] = is_a ? y * cs[u][e] - oth * sn[u][e] : y * cs[u][e] + oth * sn[u][e];
        }
      } else {
        float d[8];
        unpack8(cd[u], d);
        float sg = 0.f, sgx = 0.f;
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float oth = __shfl_xor_sync(0xffffffffu, d[e], 2);
          // da = da' c + db' s ; db = db' c - da' s
          d[e] = is_a ? d[e] * cs[u][e] + oth * sn[u][e] : d[e] * cs[u][e] - oth * sn[u][e];
          if (p.qk_norm) {
            aw[e] += d[e] * x[e];
            ab[e] += d[e];
            d[e] *= w[e];                              // g
            sg += d[e];
            sgx += d[e] * x[e];
          }
        }
        if (p.qk_norm) {
          const float mg = sum8(sg) * (1.0f / 64.0f), mgx = sum8(sgx) * (1.0f / 64.0f);
#pragma unroll
          for (int e = 0; e < 8; ++e) d[e] = rstd * (d[e] - mg - x[e] * mgx);
        }
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = d[e];
      }
      if (r + u < r_end)
        *reinterpret_cast<uint4*>(io + (r + u) * ldio + col) =
            make_uint4(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]), pack_bf16x2(o[4], o[5]), pack_bf16x2(o[6], o[7]));
    }
  }
  if (BWD && p.qk_norm) {
    // the 64 + 64 parameter gradients live in 16 sectors of L2: atomics from every warp serialise there (measured: 2/3 of
    // the kernel's time), so reduce over the warp's 4 heads (shuffles) and the CTA's 8 warps (shared memory) first
    __shared__ float red[8][2][64];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      aw[e] += __shfl_xor_sync(0xffffffffu, aw[e], 8);
      aw[e] += __shfl_xor_sync(0xffffffffu, aw[e], 16);
      ab[e] += __shfl_xor_sync(0xffffffffu, ab[e], 8);
      ab[e] += __shfl_xor_sync(0xffffffffu, ab[e], 16);
    }
    if (lane < 8) {
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        red[warp][0][8 * q + e] = aw[e];
        red[warp][1][8 * q + e] = ab[e];
      }
    }
    __syncthreads();
    if (threadIdx.x < 128) {
      const int wb = threadIdx.x >> 6, c = threadIdx.x & 63;
      float t = 0.f;
#pragma unroll
      for (int k = 0; k < 8; ++k) t += red[k][wb][c];
      float* dst = wb == 0 ? (which == 0 ? dqn_w : dkn_w) : (which == 0 ? dqn_b : dkn_b);
      if (dst) atomicAdd(dst + c, t);
    }
  }
}

// dst[idx[r], :] += src[r, :] for unique idx (the gradient of the scene-token gather, aggregator.py:617-625)
__global__ void __launch_bounds__(256)
scatter_add_rows_kernel(float* __restrict__ dst, int64_t ldd, const int32_t* __restrict__ idx, const float* __restrict__ src,
                        int64_t lds, int64_t n, int C) {
  const int64_t e = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) * 4;
  if (e >= n * C) return;
  const int64_t r = e / C;
  const int c = static_cast<int>(e % C);
  const float4 v = *reinterpret_cast<const float4*>(src + r * lds + c);
  float* o = dst + static_cast<int64_t>(idx[r]) * ldd + c;
  float4 t = *reinterpret_cast<float4*>(o);
  t.x += v.x; t.y += v.y; t.z += v.z; t.w += v.w;
  *reinterpret_cast<float4*>(o) = t;
}


// ------------------------------------------------------------------------------------------------ Adam
// torch.optim.Adam (betas, eps, no weight decay, no amsgrad; train_imc.py:475) on a flat bucket: 16 B read + 12 B written
// per parameter, float4 lanes, grid-stride over a grid of 148 x 8 CTAs.  g is multiplied by grad_scale first (1 / world
// after a sum all-reduce, or the GradScaler's 1 / scale, train_imc.py:380-381).
__global__ void __launch_bounds__(256)
adam_step_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v, int64_t n4,
                 float step_size, float beta1, float beta2, float omb1, float omb2, float eps, float inv_sqrt_bc2, float grad_scale) {
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n4; i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    float4 pp = reinterpret_cast<float4*>(p)[i];
    const float4 gg = reinterpret_cast<const float4*>(g)[i];
    float4 mm = reinterpret_cast<float4*>(m)[i];
    float4 vv = reinterpret_cast<float4*>(v)[i];
    auto upd = [&](float& pe, float ge, float& me, float& ve) {
      ge *= grad_scale;
      me = beta1 * me + omb1 * ge;                     // omb = 1 - beta evaluated in double on the host, as torch does
      ve = beta2 * ve + omb2 * ge * ge;
      pe -= step_size * me / (sqrtf(ve) * inv_sqrt_bc2 + eps);
    };
    upd(pp.x, gg.x, mm.x, vv.x);
    upd(pp.y, gg.y, mm.y, vv.y);
    upd(pp.z, gg.z, mm.z, vv.z);
    upd(pp.w, gg.w, mm.w, vv.w);
    reinterpret_cast<float4*>(p)[i] = pp;
    reinterpret_cast<float4*>(m)[i] = mm;
    reinterpret_cast<float4*>(v)[i] = vv;
  }
}

}  // namespace sr

using namespace sr;

#define SR_LAUNCHED()          \
  do {                         \
    ++g_launch_count;          \
    SR_CHECK_CUDA(cudaGetLastError()); \
  } while (0)

extern "C" int sr_transpose_bf16(const void* src, int64_t lds, void* dst, int64_t ldd, int64_t R, int64_t C, int64_t R_pad,
                                 float* colsum, void* stream) {
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(src && dst && R > 0 && C > 0 && R_pad >= R && ldd >= R_pad && lds >= C, "sr_transpose_bf16: bad shape");
  dim3 grid(static_cast<unsigned>((R_pad + 63) / 64), static_cast<unsigned>((C + 63) / 64));
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const bool vec = R_pad % 64 == 0 && C % 8 == 0 && lds % 8 == 0 && ldd % 8 == 0 &&
                   ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0;
  SR_REQUIRE(vec || colsum == nullptr, "sr_transpose_bf16: the fused column sum needs the 16-byte aligned layout "
                                       "(R_pad % 64 == 0, C, lds, ldd multiples of 8)");
  if (vec) {
    dim3 vgrid(static_cast<unsigned>((R_pad / 64 + 3) / 4), grid.y);
    transpose_bf16_vec_kernel<<<vgrid, 256, 0, st>>>(reinterpret_cast<const __nv_bfloat16*>(src), lds,
                                                     reinterpret_cast<__nv_bfloat16*>(dst), ldd, static_cast<int>(R),
                                                     static_cast<int>(C), static_cast<int>(R_pad), colsum);
  }
  else
    transpose_bf16_kernel<<<grid, 256, 0, st>>>(reinterpret_cast<const __nv_bfloat16*>(src), lds,
                                                reinterpret_cast<__nv_bfloat16*>(dst), ldd, static_cast<int>(R), static_cast<int>(C),
                                                static_cast<int>(R_pad));
  SR_LAUNCHED();
  return SR_OK;
}

extern "C" int sr_colsum(const void* a, int a_bf16, int64_t lda, const void* b, int64_t ldb, float* out, int64_t R, int64_t C,
                         int accumulate, void* stream) {
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(a && out && R > 0 && C > 0, "sr_colsum: bad arguments");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (!accumulate) SR_CHECK_CUDA(cudaMemsetAsync(out, 0, static_cast<size_t>(C) * sizeof(float), st));
  dim3 grid(static_cast<unsigned>((C + 127) / 128), static_cast<unsigned>((R + 127) / 128));
  if (a_bf16 && b == nullptr && C % 8 == 0 && lda % 8 == 0 && (reinterpret_cast<uintptr_t>(a) & 15) == 0) {
    dim3 g2(static_cast<unsigned>((C + 255) / 256), static_cast<unsigned>((R + 255) / 256));
    colsum_bf16_kernel<<<g2, 256, 0, st>>>(reinterpret_cast<const __nv_bfloat16*>(a), lda, out, static_cast<int>(R), static_cast<int>(C));
  } else if (!a_bf16 && b != nullptr && C % 8 == 0 && lda % 4 == 0 && ldb % 8 == 0 &&
             ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b)) & 15) == 0) {
    dim3 g2(static_cast<unsigned>((C + 255) / 256), static_cast<unsigned>((R + 255) / 256));
    colsum_f32xbf16_kernel<<<g2, 256, 0, st>>>(reinterpret_cast<const float*>(a), lda, reinterpret_cast<const __nv_bfloat16*>(b), ldb, out,
                                               static_cast<int>(R), static_cast<int>(C));
  } else if (a_bf16)
    colsum_kernel<true><<<grid, 128, 0, st>>>(a, lda, reinterpret_cast<const __nv_bfloat16*>(b), ldb, out, static_cast<int>(R),
                                              static_cast<int>(C));
  else
    colsum_kernel<false><<<grid, 128, 0, st>>>(a, lda, reinterpret_cast<const __nv_bfloat16*>(b), ldb, out, static_cast<int>(R),
                                               static_cast<int>(C));
  SR_LAUNCHED();
  return SR_OK;
}

extern "C" int sr_scale_cast_bf16(const float* x, int64_t ldx, const float* gamma, void* out, int64_t ldo, int64_t rows, int64_t C,
                                  void* stream) {
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(x && out && rows > 0 && C > 0 && C % 4 == 0 && ldx % 4 == 0 && ldo % 4 == 0, "sr_scale_cast_bf16: bad shape");
  const int64_t n4 = rows * C / 4;
  scale_cast_kernel<<<static_cast<unsigned>((n4 + 255) / 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      x, ldx, gamma, reinterpret_cast<__nv_bfloat16*>(out), ldo, rows, static_cast<int>(C));
  SR_LAUNCHED();
  return SR_OK;
}

extern "C" int sr_ls_residual_f32(const float* resid, const void* z, const float* gamma, float* out, int64_t rows, int64_t C,
                                  void* stream) {
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(resid && z && out && rows > 0 && C > 0 && C % 4 == 0, "sr_ls_residual_f32: bad shape (contiguous [rows, C], C % 4 == 0)");
  const int64_t n4 = rows * C / 4;
  ls_residual_kernel<<<static_cast<unsigned>((n4 + 255) / 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      resid, reinterpret_cast<const __nv_bfloat16*>(z), gamma, out, rows, static_cast<int>(C));
  SR_LAUNCHED();
  return SR_OK;
}

extern "C" int sr_gelu_fwd_bf16(const void* u, void* out, int64_t n, void* stream) {
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(u && out && n > 0 && n % 8 == 0, "sr_gelu_fwd_bf16: n must be a positive multiple of 8");
  gelu_kernel<0><<<static_cast<unsigned>((n / 8 + 255) / 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const __nv_bfloat16*>(u), nullptr, reinterpret_cast<__nv_bfloat16*>(out), n);
  SR_LAUNCHED();
  return SR_OK;
}

extern "C" int sr_gelu_bwd_bf16(const void* u, const void* dg, void* du, int64_t n, void* stream) {
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(u && dg && du && n > 0 && n % 8 == 0, "sr_gelu_bwd_bf16: n must be a positive multiple of 8");
  gelu_kernel<1><<<static_cast<unsigned>((n / 8 + 255) / 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const __nv_bfloat16*>(u), reinterpret_cast<const __nv_bfloat16*>(dg), reinterpret_cast<__nv_bfloat16*>(du), n);
  SR_LAUNCHED();
  return SR_OK;
}

extern "C" int sr_ls_bwd_bf16(const float* dy, int64_t lddy, const void* z, int64_t ldz, const float* gamma, void* dz, int64_t lddz,
                              float* dgamma, float* dbias, int64_t rows, int64_t C, void* stream) {
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(dy && z && gamma && dz && dgamma && rows > 0 && C > 0, "sr_ls_bwd_bf16: bad arguments");
  SR_REQUIRE(C % 8 == 0 && lddy % 4 == 0 && ldz % 8 == 0 && lddz % 8 == 0 && lddy >= C && ldz >= C && lddz >= C &&
                 ((reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(z) | reinterpret_cast<uintptr_t>(dz) |
                   reinterpret_cast<uintptr_t>(gamma)) & 15) == 0,
             "sr_ls_bwd_bf16: needs C % 8 == 0 and 16-byte aligned rows");
  SR_REQUIRE(rows < (1ll << 31) && (rows + 255) / 256 <= 65535, "sr_ls_bwd_bf16: too many rows");
  dim3 grid(static_cast<unsigned>((C + 255) / 256), static_cast<unsigned>((rows + 255) / 256));
  ls_bwd_kernel<<<grid, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      dy, lddy, reinterpret_cast<const __nv_bfloat16*>(z), ldz, gamma, reinterpret_cast<__nv_bfloat16*>(dz), lddz, dgamma, dbias,
      static_cast<int>(rows), static_cast<int>(C));
  SR_LAUNCHED();
  return SR_OK;
}

extern "C" int sr_gelu_bwd_colsum_bf16(const void* u, int64_t ldu, const void* dg, int64_t lddg, void* du, int64_t lddu, float* dbias,
                                       int64_t rows, int64_t C, void* stream) {
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(u && dg && du && dbias && rows > 0 && C > 0, "sr_gelu_bwd_colsum_bf16: bad arguments");
  SR_REQUIRE(C % 8 == 0 && ldu % 8 == 0 && lddg % 8 == 0 && lddu % 8 == 0 && ldu >= C && lddg >= C && lddu >= C &&
                 ((reinterpret_cast<uintptr_t>(u) | reinterpret_cast<uintptr_t>(dg) | reinterpret_cast<uintptr_t>(du)) & 15) == 0,
             "sr_gelu_bwd_colsum_bf16: needs C % 8 == 0 and 16-byte aligned rows");
  SR_REQUIRE(rows < (1ll << 31) && (rows + 255) / 256 <= 65535, "sr_gelu_bwd_colsum_bf16: too many rows");
  dim3 grid(static_cast<unsigned>((C + 255) / 256), static_cast<unsigned>((rows + 255) / 256));
  gelu_bwd_colsum_kernel<<<grid, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const __nv_bfloat16*>(u), ldu, reinterpret_cast<const __nv_bfloat16*>(dg), lddg,
      reinterpret_cast<__nv_bfloat16*>(du), lddu, dbias, static_cast<int>(rows), static_cast<int>(C));
  SR_LAUNCHED();
  return SR_OK;
}

extern "C" int sr_layernorm_bwd(const float* x, int64_t ldx, const float* weight, const void* dy, int64_t lddy, const float* dres,
                                int64_t lddres, float* dx, int64_t lddx, float* dweight, float* dbias, int64_t rows, int64_t C,
                                float eps, void* stream) {
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(x && dy && dx && rows > 0, "sr_layernorm_bwd: null pointer / empty");
  SR_REQUIRE(C == 1024, "sr_layernorm_bwd: C must be 1024 (got %lld)", static_cast<long long>(C));
  if (dres == nullptr) { dres = dx; lddres = lddx; }
  SR_REQUIRE(ldx % 4 == 0 && lddy % 4 == 0 && lddx % 4 == 0 && lddres % 4 == 0, "sr_layernorm_bwd: row strides must be multiples of 4");
  const int rows_per_block = 32;
  layernorm_bwd_kernel<8><<<static_cast<unsigned>((rows + rows_per_block - 1) / rows_per_block), 128, 0,
                            reinterpret_cast<cudaStream_t>(stream)>>>(x, ldx, weight, reinterpret_cast<const __nv_bfloat16*>(dy), lddy,
                                                                      dres, lddres, dx, lddx, dweight, dbias, static_cast<int>(rows),
                                                                      rows_per_block, eps);
  SR_LAUNCHED();
  return SR_OK;
}

static int qk_launch(bool bwd, const void* raw, int64_t ldr, void* io, int64_t ldio, float* dqw, float* dqb, float* dkw, float* dkb,
                     const sr_epilogue_t* e, int64_t M, void* stream) {
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(raw && io && e && M > 0, "sr_qknorm_rope: null pointer / empty");
  SR_REQUIRE(e->head_dim == 64 && e->qkv_dim > 0 && e->qkv_dim % 64 == 0, "sr_qknorm_rope: head_dim must be 64");
  SR_REQUIRE(ldr % 8 == 0 && ldio % 8 == 0 && ((reinterpret_cast<uintptr_t>(raw) | reinterpret_cast<uintptr_t>(io)) & 15) == 0 &&
                 (e->qkv_dim / 64) % 4 == 0,
             "sr_qknorm_rope: buffers must be 16-byte aligned, row strides multiples of 8, heads a multiple of 4");
  if (e->qk_norm) SR_REQUIRE(e->qn_w && e->qn_b && e->kn_w && e->kn_b, "sr_qknorm_rope: q/k norm parameters missing");
  if (e->rope) SR_REQUIRE(e->rope_cos && e->rope_sin && e->pos, "sr_qknorm_rope: RoPE tables / positions missing");
  QkRopeParams p;
  p.qn_w = e->qn_w; p.qn_b = e->qn_b; p.kn_w = e->kn_w; p.kn_b = e->kn_b;
  p.eps = e->qk_eps;
  p.qk_norm = e->qk_norm; p.rope = e->rope;
  p.rope_cos = e->rope_cos; p.rope_sin = e->rope_sin; p.pos = e->pos;
  p.H = static_cast<int>(e->qkv_dim / 64);
  p.M = M;
  p.rows_per_warp = 32;
  dim3 grid(static_cast<unsigned>(3 * p.H / 4), static_cast<unsigned>((M + 8 * p.rows_per_warp - 1) / (8 * p.rows_per_warp)));
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (bwd)
    qknorm_rope_kernel<true><<<grid, 256, 0, st>>>(reinterpret_cast<const __nv_bfloat16*>(raw), ldr, reinterpret_cast<__nv_bfloat16*>(io),
                                                   ldio, dqw, dqb, dkw, dkb, p);
  else
    qknorm_rope_kernel<false><<<grid, 256, 0, st>>>(reinterpret_cast<const __nv_bfloat16*>(raw), ldr, reinterpret_cast<__nv_bfloat16*>(io),
                                                    ldio, nullptr, nullptr, nullptr, nullptr, p);
  SR_LAUNCHED();
  return SR_OK;
}

extern "C" int sr_qknorm_rope_fwd(const void* raw, int64_t ldr, void* out, int64_t ldo, const sr_epilogue_t* qkv, int64_t M,
                                  void* stream) {
  return qk_launch(false, raw, ldr, out, ldo, nullptr, nullptr, nullptr, nullptr, qkv, M, stream);
}

extern "C" int sr_qknorm_rope_bwd(const void* raw, int64_t ldr, void* dqkv, int64_t ldd, const sr_epilogue_t* qkv, float* dqn_w,
                                  float* dqn_b, float* dkn_w, float* dkn_b, int64_t M, void* stream) {
  return qk_launch(true, raw, ldr, dqkv, ldd, dqn_w, dqn_b, dkn_w, dkn_b, qkv, M, stream);
}

extern "C" int sr_scatter_add_rows_f32(float* dst, int64_t ldd, const int32_t* idx, const float* src, int64_t lds, int64_t n,
                                       int64_t C, void* stream) {
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(dst && idx && src && n > 0 && C > 0 && C % 4 == 0 && ldd % 4 == 0 && lds % 4 == 0, "sr_scatter_add_rows_f32: bad shape");
  const int64_t n4 = n * C / 4;
  scatter_add_rows_kernel<<<static_cast<unsigned>((n4 + 255) / 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      dst, ldd, idx, src, lds, n, static_cast<int>(C));
  SR_LAUNCHED();
  return SR_OK;
}

extern "C" int sr_adam_step_f32(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int64_t n, double lr, double beta1,
                                double beta2, double eps, int64_t step, double grad_scale, void* stream) {
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(param && grad && exp_avg && exp_avg_sq && n > 0 && n % 4 == 0 && step >= 1, "sr_adam_step_f32: bad arguments (n % 4 == 0, step >= 1)");
  SR_REQUIRE(((reinterpret_cast<uintptr_t>(param) | reinterpret_cast<uintptr_t>(grad) | reinterpret_cast<uintptr_t>(exp_avg) |
               reinterpret_cast<uintptr_t>(exp_avg_sq)) & 15) == 0, "sr_adam_step_f32: buffers must be 16-byte aligned");
  const double bc1 = 1.0 - pow(beta1, static_cast<double>(step));
  const double bc2 = 1.0 - pow(beta2, static_cast<double>(step));
  const int64_t n4 = n / 4;
  const int64_t want = (n4 + 255) / 256;
  const unsigned grid = static_cast<unsigned>(want < 148 * 8 ? want : 148 * 8);
  adam_step_kernel<<<grid, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(param, grad, exp_avg, exp_avg_sq, n4,
                                                                             static_cast<float>(lr / bc1), static_cast<float>(beta1),
                                                                             static_cast<float>(beta2), static_cast<float>(1.0 - beta1),
                                                                             static_cast<float>(1.0 - beta2), static_cast<float>(eps),
                                                                             static_cast<float>(1.0 / sqrt(bc2)),
                                                                             static_cast<float>(grad_scale));
  SR_LAUNCHED();
  return SR_OK;
}
