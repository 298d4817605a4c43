// Self-supervised training loss of train/train_imc.py:141-256: reprojection residuals of sampled
// correspondences under the PREDICTED cameras, scored by a per-frame empirical CDF.
//
//   reference                                                         here
//   train/utils/geometry.py:252-287  compute_relative_pose            sr_relative_pose_f32 (+ _bwd)
//   train/utils/geometry.py:89-158   backproject_and_reproject        sr_reproject_f32 (+ _bwd), dst_depth == nullptr
//   train/utils/geometry.py:160-250  ..._with_approximation           sr_reproject_f32 (+ _bwd), dst_depth != nullptr
//   train/utils/geometry.py:289-303  compute_projective_residual      sr_projective_residual_f32 (+ _bwd)
//   train/losses/cdf_loss.py:122-186 per-frame weighted histogram -> PMF -> CDF -> Sobel -> Gaussian (reflect padding)
//                                                                     sr_cdf_tables_f32
//   train/losses/cdf_loss.py:188-242 CDF / PDF lookup (+0.5 rounding, 2.0 for invalid, pdf * w as the gradient)
//                                                                     sr_cdf_lookup_f32
//   train/train_imc.py:155-251       compute_loss, the whole chain    sr_reproj_cdf_loss_f32: loss AND d loss / d (predicted
//                                                                     intrinsics, extrinsics) in 7 launches + 1 memset
//
// The reference runs ~150 PyTorch launches per loss (bmm / inverse / scatter_add / conv1d / gather, forward and backward)
// over [pairs, K] fp32 arrays of a few hundred thousand points: launch-bound.  Here the per-point work is two fused
// HBM-bound passes (24 B read per point in the first, 32 B in the second, no intermediate coordinate arrays), everything
// per pair / per frame is a thread per pair / block per frame.  fp32 throughout, arithmetic in the reference's order
// (K^-1 x, * depth, R . + t, K_dst ., divide) with contraction disabled where the reference's bin index depends on it.
//
// Matrix inverses: torch.inverse is an LU factorisation; here 3 x 3 by the adjugate and 4 x 4 by Gauss-Jordan with partial
// pivoting — same value up to fp32 rounding.  Histogram sums are fp32 atomics: exact and order-independent for the unit
// weights compute_loss uses (train_imc.py:238-240), order-dependent in the last bit for general weights exactly like
// torch's scatter_add_.
#include <mutex>

#include "common.cuh"
#include "sr_internal.h"

namespace sr {
extern int g_launch_count;

constexpr int kPairStride = 32;     // floats per pair: Ki[0:9] | R[9:18] | t[18:21] | Kd[21:30] | pad
constexpr int kGradStride = 32;     // per-block partial sums: dKi[0:9] | dR[9:18] | dt[18:21] | dKd[21:30] | loss | pad
constexpr int kPointThreads = 256;
constexpr int kPointsPerThread = 4;                               // backward / fused kernels: 1024 points per block
constexpr int kBlockPoints = kPointThreads * kPointsPerThread;
constexpr float kDivEps = 1e-6f;    // geometry.py:34 (from_homogeneous) and :245

__device__ __forceinline__ void inv3(const float* a, float* o) {
  const float c00 = a[4] * a[8] - a[5] * a[7], c01 = a[5] * a[6] - a[3] * a[8], c02 = a[3] * a[7] - a[4] * a[6];
  const float r = 1.0f / (a[0] * c00 + a[1] * c01 + a[2] * c02);
  o[0] = c00 * r; o[1] = (a[2] * a[7] - a[1] * a[8]) * r; o[2] = (a[1] * a[5] - a[2] * a[4]) * r;
  o[3] = c01 * r; o[4] = (a[0] * a[8] - a[2] * a[6]) * r; o[5] = (a[2] * a[3] - a[0] * a[5]) * r;
  o[6] = c02 * r; o[7] = (a[1] * a[6] - a[0] * a[7]) * r; o[8] = (a[0] * a[4] - a[1] * a[3]) * r;
}

// General 4 x 4 inverse, Gauss-Jordan with partial pivoting on the augmented [A | I].
__device__ void inv4(const float* a, float* o) {
  float m[4][8];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      m[i][j] = a[i * 4 + j];
      m[i][4 + j] = i == j ? 1.0f : 0.0f;
    }
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    int piv = c;
    float best = fabsf(m[c][c]);
#pragma unroll
    for (int r = c + 1; r < 4; ++r)
      if (fabsf(m[r][c]) > best) { best = fabsf(m[r][c]); piv = r; }
#pragma unroll
    for (int r = c + 1; r < 4; ++r)
      if (r == piv) {
#pragma unroll
        for (int j = 0; j < 8; ++j) { const float tmp = m[c][j]; m[c][j] = m[r][j]; m[r][j] = tmp; }
      }
    const float inv = 1.0f / m[c][c];
#pragma unroll
    for (int j = 0; j < 8; ++j) m[c][j] *= inv;
#pragma unroll
    for (int r = 0; r < 4; ++r)
      if (r != c) {
        const float f = m[r][c];
#pragma unroll
        for (int j = 0; j < 8; ++j) m[r][j] -= f * m[c][j];
      }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) o[i * 4 + j] = m[i][4 + j];
}

template <int N>
__device__ __forceinline__ void matmul(const float* a, const float* b, float* c) {        // c = a b
#pragma unroll
  for (int i = 0; i < N; ++i)
#pragma unroll
    for (int j = 0; j < N; ++j) {
      float s = 0.f;
#pragma unroll
      for (int k = 0; k < N; ++k) s += a[i * N + k] * b[k * N + j];
      c[i * N + j] = s;
    }
}
template <int N>
__device__ __forceinline__ void matmul_tn(const float* a, const float* b, float* c) {     // c = a^T b
#pragma unroll
  for (int i = 0; i < N; ++i)
#pragma unroll
    for (int j = 0; j < N; ++j) {
      float s = 0.f;
#pragma unroll
      for (int k = 0; k < N; ++k) s += a[k * N + i] * b[k * N + j];
      c[i * N + j] = s;
    }
}
template <int N>
__device__ __forceinline__ void matmul_nt(const float* a, const float* b, float* c) {     // c = a b^T
#pragma unroll
  for (int i = 0; i < N; ++i)
#pragma unroll
    for (int j = 0; j < N; ++j) {
      float s = 0.f;
#pragma unroll
      for (int k = 0; k < N; ++k) s += a[i * N + k] * b[j * N + k];
      c[i * N + j] = s;
    }
}

// pad_poses (geometry.py:36-57): rows == 3 appends (0, 0, 0, 1); rows == 4 is taken as is.
__device__ __forceinline__ void load_pose44(const float* e, int rows, float* m) {
#pragma unroll
  for (int i = 0; i < 12; ++i) m[i] = e[i];
  if (rows == 4) {
    m[12] = e[12]; m[13] = e[13]; m[14] = e[14]; m[15] = e[15];
  } else {
    m[12] = 0.f; m[13] = 0.f; m[14] = 0.f; m[15] = 1.0f;
  }
}

// relative = dst @ inverse(src)  (geometry.py:281-285)
__global__ void relative_pose_kernel(const float* __restrict__ src, const float* __restrict__ dst, int rows,
                                     float* __restrict__ rel, int64_t P) {
  const int64_t p = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (p >= P) return;
  float es[16], ed[16], esi[16], r[16];
  load_pose44(src + p * rows * 4, rows, es);
  load_pose44(dst + p * rows * 4, rows, ed);
  inv4(es, esi);
  matmul<4>(ed, esi, r);
#pragma unroll
  for (int i = 0; i < 16; ++i) rel[p * 16 + i] = r[i];
}

// d dst = G inv(src)^T ;  d inv = dst^T G ;  d src = -inv^T (d inv) inv^T     (G = d relative, 4 x 4)
__device__ __forceinline__ void relative_pose_grad(const float* ed, const float* esi, const float* g, float* d_src, float* d_dst) {
  float t0[16], t1[16];
  matmul_nt<4>(g, esi, d_dst);
  matmul_tn<4>(ed, g, t0);          // d inv
  matmul_tn<4>(esi, t0, t1);        // inv^T d inv
  matmul_nt<4>(t1, esi, d_src);     // (...) inv^T
#pragma unroll
  for (int i = 0; i < 16; ++i) d_src[i] = -d_src[i];
}

__global__ void relative_pose_bwd_kernel(const float* __restrict__ src, const float* __restrict__ dst, int rows,
                                         const float* __restrict__ d_rel, float* __restrict__ d_src, float* __restrict__ d_dst,
                                         int64_t P) {
  const int64_t p = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (p >= P) return;
  float es[16], ed[16], esi[16], g[16], gs[16], gd[16];
  load_pose44(src + p * rows * 4, rows, es);
  load_pose44(dst + p * rows * 4, rows, ed);
  inv4(es, esi);
#pragma unroll
  for (int i = 0; i < 16; ++i) g[i] = d_rel[p * 16 + i];
  relative_pose_grad(ed, esi, g, gs, gd);
  for (int i = 0; i < rows * 4; ++i) {
    d_src[p * rows * 4 + i] = gs[i];
    d_dst[p * rows * 4 + i] = gd[i];
  }
}

// ------------------------------------------------------------------------------------------------ per-pair matrices
// Ki = inverse(src K) (geometry.py:139), [R | t] = top three rows of the relative pose, Kd = dst K.
__global__ void pair_setup_pairs_kernel(const float* __restrict__ src_k, const float* __restrict__ dst_k,
                                        const float* __restrict__ rel, int rows, float* __restrict__ pm, int64_t P) {
  const int64_t p = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (p >= P) return;
  float ks[9], ki[9];
#pragma unroll
  for (int i = 0; i < 9; ++i) ks[i] = src_k[p * 9 + i];
  inv3(ks, ki);
  float* o = pm + p * kPairStride;
  const float* r = rel + p * rows * 4;
#pragma unroll
  for (int i = 0; i < 9; ++i) o[i] = ki[i];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
#pragma unroll
    for (int j = 0; j < 3; ++j) o[9 + i * 3 + j] = r[i * 4 + j];
    o[18 + i] = r[i * 4 + 3];
  }
#pragma unroll
  for (int i = 0; i < 9; ++i) o[21 + i] = dst_k[p * 9 + i];
  o[30] = 0.f;
  o[31] = 0.f;
}

// The same from per-FRAME cameras and the pair index lists (train_imc.py:185-192): intrinsics[src_idx], poses[dst_idx] ...
__global__ void pair_setup_frames_kernel(const float* __restrict__ k_rec, const float* __restrict__ extr, int rows,
                                         const int64_t* __restrict__ src_idx, const int64_t* __restrict__ dst_idx,
                                         float* __restrict__ pm, int64_t P) {
  const int64_t p = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (p >= P) return;
  const int64_t s = src_idx[p], d = dst_idx[p];
  float ks[9], ki[9], es[16], ed[16], esi[16], r[16];
#pragma unroll
  for (int i = 0; i < 9; ++i) ks[i] = k_rec[s * 9 + i];
  inv3(ks, ki);
  load_pose44(extr + s * rows * 4, rows, es);
  load_pose44(extr + d * rows * 4, rows, ed);
  inv4(es, esi);
  matmul<4>(ed, esi, r);
  float* o = pm + p * kPairStride;
#pragma unroll
  for (int i = 0; i < 9; ++i) o[i] = ki[i];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
#pragma unroll
    for (int j = 0; j < 3; ++j) o[9 + i * 3 + j] = r[i * 4 + j];
    o[18 + i] = r[i * 4 + 3];
  }
#pragma unroll
  for (int i = 0; i < 9; ++i) o[21 + i] = k_rec[d * 9 + i];
  o[30] = 0.f;
  o[31] = 0.f;
}

// ------------------------------------------------------------------------------------------------ per-point geometry
struct PointFwd {
  float s[3];      // depth-scaled source-camera point  (K^-1 x~) * z
  float d[3];      // the point in the destination camera
  float h[3];      // K_dst d
};

__device__ __forceinline__ void project_point(const float* pm, float u, float v, float z, PointFwd& f) {
#pragma unroll
  for (int i = 0; i < 3; ++i) f.s[i] = (pm[i * 3] * u + pm[i * 3 + 1] * v + pm[i * 3 + 2]) * z;
#pragma unroll
  for (int i = 0; i < 3; ++i)
    f.d[i] = pm[9 + i * 3] * f.s[0] + pm[9 + i * 3 + 1] * f.s[1] + pm[9 + i * 3 + 2] * f.s[2] + pm[18 + i];
#pragma unroll
  for (int i = 0; i < 3; ++i) f.h[i] = pm[21 + i * 3] * f.d[0] + pm[21 + i * 3 + 1] * f.d[1] + pm[21 + i * 3 + 2] * f.d[2];
}

// Adds one point's contribution to the 30 per-pair gradient sums, given d loss / d h.
__device__ __forceinline__ void accumulate_point_grad(const float* pm, float u, float v, float z, const PointFwd& f,
                                                      const float* dh, float* acc) {
  float dd[3], dc[3];
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) acc[21 + i * 3 + j] += dh[i] * f.d[j];                            // d Kd
#pragma unroll
  for (int j = 0; j < 3; ++j) dd[j] = pm[21 + j] * dh[0] + pm[24 + j] * dh[1] + pm[27 + j] * dh[2];   // Kd^T dh
#pragma unroll
  for (int i = 0; i < 3; ++i) {
#pragma unroll
    for (int j = 0; j < 3; ++j) acc[9 + i * 3 + j] += dd[i] * f.s[j];                              // d R
    acc[18 + i] += dd[i];                                                                        // d t
  }
#pragma unroll
  for (int j = 0; j < 3; ++j) dc[j] = (pm[9 + j] * dd[0] + pm[12 + j] * dd[1] + pm[15 + j] * dd[2]) * z;   // R^T dd, * depth
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    acc[i * 3] += dc[i] * u;                                                                     // d Ki
    acc[i * 3 + 1] += dc[i] * v;
    acc[i * 3 + 2] += dc[i];
  }
}

template <int NV>
__device__ __forceinline__ void block_reduce_store(float* v, float* sh /* [warps][NV] */, float* dst) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    float x = v[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    if (lane == 0) sh[warp * NV + i] = x;
  }
  __syncthreads();
  if (threadIdx.x < NV) {
    float s = 0.f;
    for (int w = 0; w < nw; ++w) s += sh[w * NV + threadIdx.x];
    dst[threadIdx.x] = s;
  }
}

// backproject_and_reproject (dst_depth == nullptr) / ..._with_approximation: predicted destination pixel per point.
__global__ void __launch_bounds__(kPointThreads)
reproject_kernel(const float* __restrict__ coords, const float* __restrict__ depth, const float* __restrict__ dst_depth,
                 const float* __restrict__ src_scale, const float* __restrict__ dst_scale, const float* __restrict__ pm_all,
                 float* __restrict__ out, int K) {
  __shared__ float pm[kPairStride];
  const int64_t p = blockIdx.y;
  if (threadIdx.x < kPairStride) pm[threadIdx.x] = pm_all[p * kPairStride + threadIdx.x];
  __syncthreads();
  const int k = blockIdx.x * kPointThreads + threadIdx.x;
  if (k >= K) return;
  const int64_t i = p * K + k;
  const float2 uv = reinterpret_cast<const float2*>(coords)[i];
  const float z = depth[i] * (src_scale ? src_scale[p] : 1.0f);
  PointFwd f;
  project_point(pm, uv.x, uv.y, z, f);
  const float w = (dst_depth ? dst_depth[i] * (dst_scale ? dst_scale[p] : 1.0f) : f.h[2]) + kDivEps;
  reinterpret_cast<float2*>(out)[i] = make_float2(f.h[0] / w, f.h[1] / w);
}

// Backward of the above: per-block partial sums of the 30 per-pair gradients.
__global__ void __launch_bounds__(kPointThreads)
reproject_bwd_kernel(const float* __restrict__ coords, const float* __restrict__ depth, const float* __restrict__ dst_depth,
                     const float* __restrict__ src_scale, const float* __restrict__ dst_scale,
                     const float* __restrict__ pm_all, const float* __restrict__ d_out, float* __restrict__ partial, int K) {
  __shared__ float pm[kPairStride];
  __shared__ float red[(kPointThreads / 32) * kGradStride];
  const int64_t p = blockIdx.y;
  if (threadIdx.x < kPairStride) pm[threadIdx.x] = pm_all[p * kPairStride + threadIdx.x];
  __syncthreads();
  float acc[kGradStride];
#pragma unroll
  for (int i = 0; i < kGradStride; ++i) acc[i] = 0.f;
#pragma unroll
  for (int it = 0; it < kPointsPerThread; ++it) {
    const int k = blockIdx.x * kBlockPoints + it * kPointThreads + threadIdx.x;
    if (k >= K) continue;
    const int64_t i = p * K + k;
    const float2 uv = reinterpret_cast<const float2*>(coords)[i];
    const float2 g = reinterpret_cast<const float2*>(d_out)[i];
    const float z = depth[i] * (src_scale ? src_scale[p] : 1.0f);
    PointFwd f;
    project_point(pm, uv.x, uv.y, z, f);
    float dh[3];
    if (dst_depth) {
      const float w = dst_depth[i] * (dst_scale ? dst_scale[p] : 1.0f) + kDivEps;
      dh[0] = g.x / w; dh[1] = g.y / w; dh[2] = 0.f;
    } else {
      const float w = f.h[2] + kDivEps;
      dh[0] = g.x / w; dh[1] = g.y / w;
      dh[2] = -(g.x * f.h[0] + g.y * f.h[1]) / (w * w);
    }
    accumulate_point_grad(pm, uv.x, uv.y, z, f, dh, acc);
  }
  block_reduce_store<kGradStride>(acc, red, partial + (p * gridDim.x + blockIdx.x) * kGradStride);
}

// Per pair: reduce the block partials, chain d Ki back through the inverse: d Ks = -Ki^T (d Ki) Ki^T.
__global__ void reproject_pair_bwd_kernel(const float* __restrict__ pm_all, const float* __restrict__ partial, int nblk,
                                          int rows, float* __restrict__ d_src_k, float* __restrict__ d_dst_k,
                                          float* __restrict__ d_rel, int64_t P) {
  const int64_t p = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (p >= P) return;
  float g[30];
#pragma unroll
  for (int i = 0; i < 30; ++i) g[i] = 0.f;
  for (int b = 0; b < nblk; ++b) {
    const float* src = partial + (p * nblk + b) * kGradStride;
#pragma unroll
    for (int i = 0; i < 30; ++i) g[i] += src[i];
  }
  const float* pm = pm_all + p * kPairStride;
  float ki[9], t0[9], t1[9];
#pragma unroll
  for (int i = 0; i < 9; ++i) ki[i] = pm[i];
  matmul_tn<3>(ki, g, t0);
  matmul_nt<3>(t0, ki, t1);
#pragma unroll
  for (int i = 0; i < 9; ++i) {
    d_src_k[p * 9 + i] = -t1[i];
    d_dst_k[p * 9 + i] = g[21 + i];
  }
  float* r = d_rel + p * rows * 4;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
#pragma unroll
    for (int j = 0; j < 3; ++j) r[i * 4 + j] = g[9 + i * 3 + j];
    r[i * 4 + 3] = g[18 + i];
  }
  if (rows == 4) { r[12] = 0.f; r[13] = 0.f; r[14] = 0.f; r[15] = 0.f; }
}

// compute_projective_residual: || predicted - actual ||_2 per point, and its backward (0 where the residual is 0, as
// torch.norm's).
__global__ void projective_residual_kernel(const float* __restrict__ pred, const float* __restrict__ actual,
                                           float* __restrict__ out, int64_t n) {
  const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= n) return;
  const float2 a = reinterpret_cast<const float2*>(pred)[i], b = reinterpret_cast<const float2*>(actual)[i];
  const float dx = a.x - b.x, dy = a.y - b.y;
  out[i] = sqrtf(dx * dx + dy * dy);
}
__global__ void projective_residual_bwd_kernel(const float* __restrict__ pred, const float* __restrict__ actual,
                                               const float* __restrict__ d_out, float* __restrict__ d_pred, int64_t n) {
  const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= n) return;
  const float2 a = reinterpret_cast<const float2*>(pred)[i], b = reinterpret_cast<const float2*>(actual)[i];
  const float dx = a.x - b.x, dy = a.y - b.y;
  const float r = sqrtf(dx * dx + dy * dy);
  const float s = r > 0.f ? d_out[i] / r : 0.f;
  reinterpret_cast<float2*>(d_pred)[i] = make_float2(dx * s, dy * s);
}

// ------------------------------------------------------------------------------------------------ per-frame CDF tables
// Bin index as torch computes it ON THE GPU: `(r - min) / bin_width` with a Python-float divisor is a multiplication by
// inv_bw = fp32(1 / bin_width), the reciprocal taken in double (ATen's CPU-scalar fast path; scripts/div_probe.py on the
// B200: 0 of 2e7 values differ, true division differs in 363 bins), `.long()` truncates toward zero.
__device__ __forceinline__ long long bin_trunc(float r, float min_val, float inv_bw) {
  return static_cast<long long>(__fmul_rn(__fsub_rn(r, min_val), inv_bw));                       // cdf_loss.py:138-139
}
__device__ __forceinline__ long long bin_round(float r, float min_val, float inv_bw) {
  return static_cast<long long>(__fadd_rn(__fmul_rn(__fsub_rn(r, min_val), inv_bw), 0.5f));      // cdf_loss.py:206
}

// Histogram of one [P, K] residual array into the histograms of BOTH frames of each pair (cdf_loss.py:141-176).  All points
// of a block belong to one pair, i.e. to the same two frames: the block bins into shared memory and flushes its non-empty
// bins with two atomics each; the per-frame total weight (all points, in range or not) is block-reduced first.
__global__ void __launch_bounds__(kPointThreads)
cdf_hist_kernel(const float* __restrict__ res, const float* __restrict__ weights, const int64_t* __restrict__ src_idx,
                const int64_t* __restrict__ dst_idx, float* __restrict__ hist, float* __restrict__ total, int K,
                float min_val, float inv_bw, int bins) {
  extern __shared__ float sh[];          // [bins] + [8] warp sums
  float* sh_w = sh + bins;
  const int64_t p = blockIdx.y;
  for (int i = threadIdx.x; i < bins; i += kPointThreads) sh[i] = 0.f;
  __syncthreads();
  float wsum = 0.f;
#pragma unroll
  for (int it = 0; it < kPointsPerThread; ++it) {
    const int k = blockIdx.x * kBlockPoints + it * kPointThreads + threadIdx.x;
    if (k >= K) continue;
    const float r = res[p * K + k], w = weights ? weights[p * K + k] : 1.0f;
    wsum += w;
    const long long b = bin_trunc(r, min_val, inv_bw);
    if (b >= 0 && b < bins) atomicAdd(&sh[b], w);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) wsum += __shfl_xor_sync(0xffffffffu, wsum, o);
  if ((threadIdx.x & 31) == 0) sh_w[threadIdx.x >> 5] = wsum;
  __syncthreads();
  const int64_t s = src_idx[p], d = dst_idx[p];
  for (int i = threadIdx.x; i < bins; i += kPointThreads) {
    const float v = sh[i];
    if (v != 0.f) {
      atomicAdd(&hist[s * bins + i], v);
      atomicAdd(&hist[d * bins + i], v);
    }
  }
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int w = 0; w < kPointThreads / 32; ++w) t += sh_w[w];
    atomicAdd(&total[s], t);
    atomicAdd(&total[d], t);
  }
}

__device__ __forceinline__ int reflect(int i, int n) { return i < 0 ? -i : (i >= n ? 2 * (n - 1) - i : i); }

// One block per frame (per row of hist): PMF = hist / (total + 1e-10), CDF = running sum, raw PDF = centred difference with
// reflect padding (Conv1d [-1, 0, 1] / (2 bin_width)), PDF = raw PDF smoothed by `smooth` (2 * radius + 1 taps, reflect
// padding); cdf_loss.py:56-88,178-186.
__global__ void cdf_tables_kernel(const float* __restrict__ hist, const float* __restrict__ total, float* __restrict__ pmf,
                                  float* __restrict__ cdf, float* __restrict__ pdf, int bins, float sobel_w,
                                  const float* __restrict__ smooth, int radius) {
  extern __shared__ float sh[];          // c[bins] | raw[bins]
  float* c = sh;
  float* raw = sh + bins;
  const int64_t row = blockIdx.x;
  const float denom = total[row] + 1e-10f;
  for (int i = threadIdx.x; i < bins; i += blockDim.x) {
    const float v = hist[row * bins + i] / denom;
    pmf[row * bins + i] = v;
    c[i] = v;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    float run = 0.f;
    for (int i = 0; i < bins; ++i) { run += c[i]; c[i] = run; }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < bins; i += blockDim.x) {
    cdf[row * bins + i] = c[i];
    raw[i] = sobel_w * c[reflect(i + 1, bins)] - sobel_w * c[reflect(i - 1, bins)];
  }
  __syncthreads();
  for (int i = threadIdx.x; i < bins; i += blockDim.x) {
    float s = 0.f;
    for (int k = -radius; k <= radius; ++k) s += smooth[k + radius] * raw[reflect(i + k, bins)];
    pdf[row * bins + i] = s;
  }
}

// CDF value and gradient (= pdf * weight) of every residual in both of its frames' tables (cdf_loss.py:203-240).
__global__ void cdf_lookup_kernel(const float* __restrict__ res, const float* __restrict__ weights,
                                  const int64_t* __restrict__ src_idx, const int64_t* __restrict__ dst_idx,
                                  const float* __restrict__ cdf, const float* __restrict__ pdf, float* __restrict__ out_src,
                                  float* __restrict__ grad_src, float* __restrict__ out_dst, float* __restrict__ grad_dst,
                                  int K, int64_t n, float min_val, float inv_bw, int bins) {
  const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= n) return;
  const int64_t p = i / K;
  const float w = weights ? weights[i] : 1.0f;
  const long long b = bin_round(res[i], min_val, inv_bw);
  const bool valid = b >= 0 && b < bins && w > 0.0f;
  float cs = 2.0f, cd = 2.0f, gs = 0.f, gd = 0.f;
  if (valid) {
    const int64_t s = src_idx[p] * bins + b, d = dst_idx[p] * bins + b;
    cs = cdf[s]; cd = cdf[d];
    gs = pdf[s] * w; gd = pdf[d] * w;
  }
  out_src[i] = cs; out_dst[i] = cd;
  grad_src[i] = gs; grad_dst[i] = gd;
}

// ------------------------------------------------------------------------------------------------ fused compute_loss
// recovered = K'_to_K @ predicted intrinsics, averaged over the frames when the focal length is shared (train_imc.py:165-172).
__global__ void loss_frame_setup_kernel(const float* __restrict__ k_pred, const float* __restrict__ k2k, int shared_focal,
                                        float* __restrict__ k_rec, int N) {
  __shared__ float mean[9];
  for (int f = threadIdx.x; f < N; f += blockDim.x) {
    float a[9], b[9], c[9];
#pragma unroll
    for (int i = 0; i < 9; ++i) { b[i] = k_pred[f * 9 + i]; a[i] = k2k ? k2k[f * 9 + i] : ((i % 4) == 0 ? 1.0f : 0.0f); }
    matmul<3>(a, b, c);
#pragma unroll
    for (int i = 0; i < 9; ++i) k_rec[f * 9 + i] = c[i];
  }
  if (!shared_focal) return;
  __syncthreads();
  if (threadIdx.x < 9) {
    float s = 0.f;
    for (int f = 0; f < N; ++f) s += k_rec[f * 9 + threadIdx.x];
    mean[threadIdx.x] = s / static_cast<float>(N);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < N * 9; i += blockDim.x) k_rec[i] = mean[i % 9];
}

// Pass 1 over the points: both residuals (exact perspective division / division by the matched depth), log1p, stored for
// pass 2, and binned into the two frames' histograms of each residual kind (unit weights, train_imc.py:238-240).
__global__ void __launch_bounds__(kPointThreads)
loss_residual_hist_kernel(const float* __restrict__ src_coords, const float* __restrict__ dst_coords,
                          const float* __restrict__ src_depth, const float* __restrict__ dst_depth,
                          const float* __restrict__ pm_all, const int64_t* __restrict__ src_idx,
                          const int64_t* __restrict__ dst_idx, float* __restrict__ rl, float* __restrict__ hist,
                          float* __restrict__ total, int K, int64_t PK, int N, float min_val, float inv_bw, int bins) {
  extern __shared__ float sh[];          // hist of the exact residuals [bins] | of the approximated ones [bins]
  __shared__ float pm[kPairStride];
  const int64_t p = blockIdx.y;
  if (threadIdx.x < kPairStride) pm[threadIdx.x] = pm_all[p * kPairStride + threadIdx.x];
  for (int i = threadIdx.x; i < 2 * bins; i += kPointThreads) sh[i] = 0.f;
  __syncthreads();
  int count = 0;
#pragma unroll
  for (int it = 0; it < kPointsPerThread; ++it) {
    const int k = blockIdx.x * kBlockPoints + it * kPointThreads + threadIdx.x;
    if (k >= K) continue;
    const int64_t i = p * K + k;
    const float2 uv = reinterpret_cast<const float2*>(src_coords)[i];
    const float2 tg = reinterpret_cast<const float2*>(dst_coords)[i];
    const float z = src_depth[i];
    PointFwd f;
    project_point(pm, uv.x, uv.y, z, f);
    const float w0 = f.h[2] + kDivEps, w1 = dst_depth[i] + kDivEps;
    const float ex = f.h[0] / w0 - tg.x, ey = f.h[1] / w0 - tg.y;
    const float ax = f.h[0] / w1 - tg.x, ay = f.h[1] / w1 - tg.y;
    const float r0 = log1pf(sqrtf(ex * ex + ey * ey)), r1 = log1pf(sqrtf(ax * ax + ay * ay));
    rl[i] = r0;
    rl[PK + i] = r1;
    const long long b0 = bin_trunc(r0, min_val, inv_bw), b1 = bin_trunc(r1, min_val, inv_bw);
    if (b0 >= 0 && b0 < bins) atomicAdd(&sh[b0], 1.0f);
    if (b1 >= 0 && b1 < bins) atomicAdd(&sh[bins + b1], 1.0f);
    ++count;
  }
  __syncthreads();
  const int64_t s = src_idx[p], d = dst_idx[p];
  for (int i = threadIdx.x; i < 2 * bins; i += kPointThreads) {
    const float v = sh[i];
    if (v != 0.f) {
      const int kind = i >= bins, b = i - kind * bins;
      atomicAdd(&hist[(static_cast<int64_t>(kind) * N + s) * bins + b], v);
      atomicAdd(&hist[(static_cast<int64_t>(kind) * N + d) * bins + b], v);
    }
  }
  // total weight per frame = number of points (in range or not, cdf_loss.py:168-173): one atomic per warp and frame
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) count += __shfl_xor_sync(0xffffffffu, count, o);
  if ((threadIdx.x & 31) == 0 && count > 0) {
    const float c = static_cast<float>(count);
    atomicAdd(&total[s], c); atomicAdd(&total[d], c);
    atomicAdd(&total[N + s], c); atomicAdd(&total[N + d], c);
  }
}

// Pass 2 over the points: CDF / PDF lookups in both frames' tables for both residual kinds -> loss partial sum and
// d loss / d log-residual (cdf_loss.py:9-17: the "gradient" of the CDF value is pdf * weight), chained back through log1p,
// the 2-norm, the two divisions and the projection to the 30 per-pair sums.  The projection is recomputed (27 FMAs)
// instead of being stored.  All sums are UNSCALED; loss_frame_bwd_kernel applies 1 / (4 P K).
__global__ void __launch_bounds__(kPointThreads)
loss_lookup_bwd_kernel(const float* __restrict__ src_coords, const float* __restrict__ dst_coords,
                       const float* __restrict__ src_depth, const float* __restrict__ dst_depth,
                       const float* __restrict__ pm_all, const int64_t* __restrict__ src_idx,
                       const int64_t* __restrict__ dst_idx, const float* __restrict__ rl, const float* __restrict__ cdf,
                       const float* __restrict__ pdf, float* __restrict__ partial, int K, int64_t PK, int N, float min_val,
                       float inv_bw, int bins) {
  __shared__ float pm[kPairStride];
  __shared__ float red[(kPointThreads / 32) * kGradStride];
  const int64_t p = blockIdx.y;
  if (threadIdx.x < kPairStride) pm[threadIdx.x] = pm_all[p * kPairStride + threadIdx.x];
  __syncthreads();
  const int64_t s = src_idx[p], d = dst_idx[p];
  float acc[kGradStride];
#pragma unroll
  for (int i = 0; i < kGradStride; ++i) acc[i] = 0.f;
#pragma unroll
  for (int it = 0; it < kPointsPerThread; ++it) {
    const int k = blockIdx.x * kBlockPoints + it * kPointThreads + threadIdx.x;
    if (k >= K) continue;
    const int64_t i = p * K + k;
    float g_log[2];
#pragma unroll
    for (int kind = 0; kind < 2; ++kind) {
      const long long b = bin_round(rl[kind * PK + i], min_val, inv_bw);
      if (b >= 0 && b < bins) {
        const int64_t is = (static_cast<int64_t>(kind) * N + s) * bins + b, id = (static_cast<int64_t>(kind) * N + d) * bins + b;
        acc[30] += cdf[is] + cdf[id];
        g_log[kind] = pdf[is] + pdf[id];
      } else {
        acc[30] += 4.0f;                                  // 2.0 per invalid lookup (cdf_loss.py:234-235)
        g_log[kind] = 0.f;
      }
    }
    const float2 uv = reinterpret_cast<const float2*>(src_coords)[i];
    const float2 tg = reinterpret_cast<const float2*>(dst_coords)[i];
    const float z = src_depth[i];
    PointFwd f;
    project_point(pm, uv.x, uv.y, z, f);
    const float w0 = f.h[2] + kDivEps, w1 = dst_depth[i] + kDivEps;
    const float ex = f.h[0] / w0 - tg.x, ey = f.h[1] / w0 - tg.y;
    const float ax = f.h[0] / w1 - tg.x, ay = f.h[1] / w1 - tg.y;
    const float re = sqrtf(ex * ex + ey * ey), ra = sqrtf(ax * ax + ay * ay);
    const float se = re > 0.f ? g_log[0] / ((1.0f + re) * re) : 0.f;      // d / d residual of log1p, then of the norm
    const float sa = ra > 0.f ? g_log[1] / ((1.0f + ra) * ra) : 0.f;
    const float gex = ex * se, gey = ey * se, gax = ax * sa, gay = ay * sa;
    float dh[3];
    dh[0] = gex / w0 + gax / w1;
    dh[1] = gey / w0 + gay / w1;
    dh[2] = -(gex * f.h[0] + gey * f.h[1]) / (w0 * w0);
    accumulate_point_grad(pm, uv.x, uv.y, z, f, dh, acc);
  }
  block_reduce_store<kGradStride>(acc, red, partial + (p * gridDim.x + blockIdx.x) * kGradStride);
}

// Per pair: reduce the block partials, then back through inverse(Ks) and relative = Ed inverse(Es).
// out: d Ks [9] | d Kd [9] | d Es [16] | d Ed [16] | loss partial [1]   (51 floats, stride 52)
constexpr int kPairGrad = 52;
__global__ void loss_pair_bwd_kernel(const float* __restrict__ pm_all, const float* __restrict__ partial, int nblk,
                                     const float* __restrict__ extr, int rows, const int64_t* __restrict__ src_idx,
                                     const int64_t* __restrict__ dst_idx, float* __restrict__ pair_grad, int64_t P) {
  const int64_t p = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (p >= P) return;
  float g[31];
#pragma unroll
  for (int i = 0; i < 31; ++i) g[i] = 0.f;
  for (int b = 0; b < nblk; ++b) {
    const float* src = partial + (p * nblk + b) * kGradStride;
#pragma unroll
    for (int i = 0; i < 31; ++i) g[i] += src[i];
  }
  const float* pm = pm_all + p * kPairStride;
  float* o = pair_grad + p * kPairGrad;
  {
    float ki[9], t0[9], t1[9];
#pragma unroll
    for (int i = 0; i < 9; ++i) ki[i] = pm[i];
    matmul_tn<3>(ki, g, t0);
    matmul_nt<3>(t0, ki, t1);
#pragma unroll
    for (int i = 0; i < 9; ++i) { o[i] = -t1[i]; o[9 + i] = g[21 + i]; }
  }
  float es[16], ed[16], esi[16], g4[16], gs[16], gd[16];
  load_pose44(extr + src_idx[p] * rows * 4, rows, es);
  load_pose44(extr + dst_idx[p] * rows * 4, rows, ed);
  inv4(es, esi);
#pragma unroll
  for (int i = 0; i < 3; ++i) {
#pragma unroll
    for (int j = 0; j < 3; ++j) g4[i * 4 + j] = g[9 + i * 3 + j];
    g4[i * 4 + 3] = g[18 + i];
  }
  g4[12] = 0.f; g4[13] = 0.f; g4[14] = 0.f; g4[15] = 0.f;
  relative_pose_grad(ed, esi, g4, gs, gd);
#pragma unroll
  for (int i = 0; i < 16; ++i) { o[18 + i] = gs[i]; o[34 + i] = gd[i]; }
  o[50] = g[30];
  o[51] = 0.f;
}

// Per frame (one block, frames strided over the threads; pairs visited in index order -> deterministic sums): gather the
// pair gradients of the frame's two roles, undo the shared-focal averaging and the K'_to_K product, scale by 1 / (4 P K).
__global__ void loss_frame_bwd_kernel(const float* __restrict__ pair_grad, const int64_t* __restrict__ src_idx,
                                      const int64_t* __restrict__ dst_idx, const float* __restrict__ k2k, int shared_focal,
                                      int rows, float scale, float* __restrict__ loss, float* __restrict__ d_k,
                                      float* __restrict__ d_extr, float* __restrict__ d_krec /* scratch [N, 9] */, int N,
                                      int64_t P) {
  __shared__ float mean[9];
  for (int f = threadIdx.x; f < N; f += blockDim.x) {
    float gk[9], ge[16];
#pragma unroll
    for (int i = 0; i < 9; ++i) gk[i] = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) ge[i] = 0.f;
    for (int64_t p = 0; p < P; ++p) {
      const float* o = pair_grad + p * kPairGrad;
      if (src_idx[p] == f) {
#pragma unroll
        for (int i = 0; i < 9; ++i) gk[i] += o[i];
#pragma unroll
        for (int i = 0; i < 16; ++i) ge[i] += o[18 + i];
      }
      if (dst_idx[p] == f) {
#pragma unroll
        for (int i = 0; i < 9; ++i) gk[i] += o[9 + i];
#pragma unroll
        for (int i = 0; i < 16; ++i) ge[i] += o[34 + i];
      }
    }
#pragma unroll
    for (int i = 0; i < 9; ++i) d_krec[f * 9 + i] = gk[i];
    for (int i = 0; i < rows * 4; ++i) d_extr[f * rows * 4 + i] = ge[i] * scale;
  }
  __syncthreads();
  if (shared_focal) {                       // recovered[f] = mean over frames  ->  every frame receives sum(d recovered) / N
    if (threadIdx.x < 9) {
      float s = 0.f;
      for (int f = 0; f < N; ++f) s += d_krec[f * 9 + threadIdx.x];
      mean[threadIdx.x] = s / static_cast<float>(N);
    }
    __syncthreads();
  }
  for (int f = threadIdx.x; f < N; f += blockDim.x) {
    float a[9], g[9], c[9];
#pragma unroll
    for (int i = 0; i < 9; ++i) {
      g[i] = shared_focal ? mean[i] : d_krec[f * 9 + i];
      a[i] = k2k ? k2k[f * 9 + i] : ((i % 4) == 0 ? 1.0f : 0.0f);
    }
    matmul_tn<3>(a, g, c);                  // d predicted = K'_to_K^T d recovered
#pragma unroll
    for (int i = 0; i < 9; ++i) d_k[f * 9 + i] = c[i] * scale;
  }
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int64_t p = 0; p < P; ++p) s += pair_grad[p * kPairGrad + 50];
    loss[0] = s * scale;
  }
}

}  // namespace sr

// ================================================================================================ C-ABI
namespace {
inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }
inline unsigned blocks_for(int64_t n, int per) { return static_cast<unsigned>((n + per - 1) / per); }
constexpr int64_t kMaxPairs = 65535;       // gridDim.y
constexpr int kMaxBins = 4096;
}  // namespace

extern "C" int sr_relative_pose_f32(const float* src_extrinsic, const float* dst_extrinsic, int32_t rows, float* relative,
                                    int64_t num_pairs, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(rows == 3 || rows == 4, "sr_relative_pose_f32: extrinsics must be [P, 3, 4] or [P, 4, 4]");
  if (num_pairs == 0) return SR_OK;
  SR_REQUIRE(src_extrinsic && dst_extrinsic && relative && num_pairs > 0, "sr_relative_pose_f32: null pointer");
  relative_pose_kernel<<<blocks_for(num_pairs, 64), 64, 0, as_stream(stream)>>>(src_extrinsic, dst_extrinsic, rows, relative,
                                                                                 num_pairs);
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

extern "C" int sr_relative_pose_bwd_f32(const float* src_extrinsic, const float* dst_extrinsic, int32_t rows,
                                        const float* d_relative, float* d_src, float* d_dst, int64_t num_pairs, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(rows == 3 || rows == 4, "sr_relative_pose_bwd_f32: extrinsics must be [P, 3, 4] or [P, 4, 4]");
  if (num_pairs == 0) return SR_OK;
  SR_REQUIRE(src_extrinsic && dst_extrinsic && d_relative && d_src && d_dst && num_pairs > 0, "sr_relative_pose_bwd_f32: null pointer");
  relative_pose_bwd_kernel<<<blocks_for(num_pairs, 64), 64, 0, as_stream(stream)>>>(src_extrinsic, dst_extrinsic, rows,
                                                                                     d_relative, d_src, d_dst, num_pairs);
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

extern "C" int64_t sr_reproject_workspace_floats(int64_t num_pairs, int64_t num_points) {
  const int64_t nblk = (num_points + sr::kBlockPoints - 1) / sr::kBlockPoints;
  return num_pairs * sr::kPairStride + num_pairs * nblk * sr::kGradStride;
}

extern "C" int sr_reproject_f32(const float* src_coords, const float* src_depth, const float* dst_depth, const float* src_k,
                                const float* dst_k, const float* relative, int32_t rows, const float* src_scale,
                                const float* dst_scale, float* out_coords, int64_t num_pairs, int64_t num_points,
                                float* workspace, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(rows == 3 || rows == 4, "sr_reproject_f32: relative pose must be [P, 3, 4] or [P, 4, 4]");
  SR_REQUIRE(num_pairs >= 0 && num_pairs <= kMaxPairs && num_points >= 0 && num_points < (1 << 30), "sr_reproject_f32: bad shape");
  if (num_pairs == 0 || num_points == 0) return SR_OK;
  SR_REQUIRE(src_coords && src_depth && src_k && dst_k && relative && out_coords && workspace, "sr_reproject_f32: null pointer");
  cudaStream_t st = as_stream(stream);
  pair_setup_pairs_kernel<<<blocks_for(num_pairs, 64), 64, 0, st>>>(src_k, dst_k, relative, rows, workspace, num_pairs);
  dim3 grid(blocks_for(num_points, kPointThreads), static_cast<unsigned>(num_pairs));
  reproject_kernel<<<grid, kPointThreads, 0, st>>>(src_coords, src_depth, dst_depth, src_scale, dst_scale, workspace, out_coords,
                                                   static_cast<int>(num_points));
  g_launch_count += 2;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

extern "C" int sr_reproject_bwd_f32(const float* src_coords, const float* src_depth, const float* dst_depth, const float* src_k,
                                    const float* dst_k, const float* relative, int32_t rows, const float* src_scale,
                                    const float* dst_scale, const float* d_coords, float* d_src_k, float* d_dst_k,
                                    float* d_relative, int64_t num_pairs, int64_t num_points, float* workspace, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(rows == 3 || rows == 4, "sr_reproject_bwd_f32: relative pose must be [P, 3, 4] or [P, 4, 4]");
  SR_REQUIRE(num_pairs >= 0 && num_pairs <= kMaxPairs && num_points > 0 && num_points < (1 << 30), "sr_reproject_bwd_f32: bad shape");
  if (num_pairs == 0) return SR_OK;
  SR_REQUIRE(src_coords && src_depth && src_k && dst_k && relative && d_coords && d_src_k && d_dst_k && d_relative && workspace,
             "sr_reproject_bwd_f32: null pointer");
  cudaStream_t st = as_stream(stream);
  const int nblk = static_cast<int>(blocks_for(num_points, kBlockPoints));
  float* pm = workspace;
  float* partial = workspace + num_pairs * kPairStride;
  pair_setup_pairs_kernel<<<blocks_for(num_pairs, 64), 64, 0, st>>>(src_k, dst_k, relative, rows, pm, num_pairs);
  dim3 grid(nblk, static_cast<unsigned>(num_pairs));
  reproject_bwd_kernel<<<grid, kPointThreads, 0, st>>>(src_coords, src_depth, dst_depth, src_scale, dst_scale, pm, d_coords,
                                                       partial, static_cast<int>(num_points));
  reproject_pair_bwd_kernel<<<blocks_for(num_pairs, 64), 64, 0, st>>>(pm, partial, nblk, rows, d_src_k, d_dst_k, d_relative,
                                                                       num_pairs);
  g_launch_count += 3;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

extern "C" int sr_projective_residual_f32(const float* predicted, const float* actual, float* residual, int64_t n, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  if (n == 0) return SR_OK;
  SR_REQUIRE(predicted && actual && residual && n > 0 && n < (1ll << 38), "sr_projective_residual_f32: bad arguments");
  projective_residual_kernel<<<blocks_for(n, 256), 256, 0, as_stream(stream)>>>(predicted, actual, residual, n);
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

extern "C" int sr_projective_residual_bwd_f32(const float* predicted, const float* actual, const float* d_residual,
                                              float* d_predicted, int64_t n, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  if (n == 0) return SR_OK;
  SR_REQUIRE(predicted && actual && d_residual && d_predicted && n > 0 && n < (1ll << 38), "sr_projective_residual_bwd_f32: bad arguments");
  projective_residual_bwd_kernel<<<blocks_for(n, 256), 256, 0, as_stream(stream)>>>(predicted, actual, d_residual, d_predicted, n);
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

extern "C" int sr_cdf_tables_f32(const float* residuals, const float* weights, const int64_t* src_idx, const int64_t* dst_idx,
                                 int64_t num_pairs, int64_t num_points, int64_t num_nodes, float min_val, float inv_bin_width,
                                 int32_t num_bins, float sobel_weight, const float* smooth_kernel, int32_t smooth_radius,
                                 float* hist_total /* [num_nodes * (num_bins + 1)] scratch */, float* pmf, float* cdf,
                                 float* pdf, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(num_bins >= 2 && num_bins <= kMaxBins, "sr_cdf_tables_f32: num_bins must be in [2, 4096]");
  SR_REQUIRE(smooth_radius >= 0 && smooth_radius < num_bins, "sr_cdf_tables_f32: smoothing radius must be below num_bins (reflect padding)");
  SR_REQUIRE(num_pairs >= 0 && num_pairs <= kMaxPairs && num_points >= 0 && num_points < (1 << 30) && num_nodes > 0 &&
                 num_nodes < (1 << 20), "sr_cdf_tables_f32: bad shape");
  SR_REQUIRE(src_idx && dst_idx && smooth_kernel && hist_total && pmf && cdf && pdf, "sr_cdf_tables_f32: null pointer");
  cudaStream_t st = as_stream(stream);
  float* hist = hist_total;
  float* total = hist_total + num_nodes * num_bins;
  SR_CHECK_CUDA(cudaMemsetAsync(hist_total, 0, sizeof(float) * num_nodes * (num_bins + 1), st));
  if (num_pairs > 0 && num_points > 0) {
    SR_REQUIRE(residuals != nullptr, "sr_cdf_tables_f32: null residuals");
    dim3 grid(blocks_for(num_points, kBlockPoints), static_cast<unsigned>(num_pairs));
    cdf_hist_kernel<<<grid, kPointThreads, sizeof(float) * (num_bins + 8), st>>>(residuals, weights, src_idx, dst_idx, hist, total,
                                                                                 static_cast<int>(num_points), min_val,
                                                                                 inv_bin_width, num_bins);
    ++g_launch_count;
  }
  cdf_tables_kernel<<<static_cast<unsigned>(num_nodes), 256, sizeof(float) * 2 * num_bins, st>>>(hist, total, pmf, cdf, pdf, num_bins,
                                                                                                  sobel_weight, smooth_kernel,
                                                                                                  smooth_radius);
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

extern "C" int sr_cdf_lookup_f32(const float* residuals, const float* weights, const int64_t* src_idx, const int64_t* dst_idx,
                                 const float* cdf, const float* pdf, int64_t num_pairs, int64_t num_points, float min_val,
                                 float inv_bin_width, int32_t num_bins, float* out_src, float* grad_src, float* out_dst,
                                 float* grad_dst, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(num_bins >= 2 && num_bins <= kMaxBins, "sr_cdf_lookup_f32: num_bins must be in [2, 4096]");
  const int64_t n = num_pairs * num_points;
  if (n == 0) return SR_OK;
  SR_REQUIRE(residuals && src_idx && dst_idx && cdf && pdf && out_src && grad_src && out_dst && grad_dst && n > 0 && n < (1ll << 38),
             "sr_cdf_lookup_f32: bad arguments");
  cdf_lookup_kernel<<<blocks_for(n, 256), 256, 0, as_stream(stream)>>>(residuals, weights, src_idx, dst_idx, cdf, pdf, out_src,
                                                                        grad_src, out_dst, grad_dst, static_cast<int>(num_points),
                                                                        n, min_val, inv_bin_width, num_bins);
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

extern "C" int64_t sr_reproj_loss_workspace_floats(int64_t num_frames, int64_t num_pairs, int64_t num_points, int32_t num_bins) {
  const int64_t nblk = (num_points + sr::kBlockPoints - 1) / sr::kBlockPoints;
  return num_frames * 18 + num_pairs * sr::kPairStride + 2 * num_pairs * num_points + 2 * num_frames * (num_bins + 1) +
         6 * num_frames * num_bins + num_pairs * nblk * sr::kGradStride + num_pairs * sr::kPairGrad;
}

extern "C" int sr_reproj_cdf_loss_f32(const sr_reproj_loss_t* a, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(a != nullptr, "sr_reproj_cdf_loss_f32: null argument");
  const int64_t N = a->num_frames, P = a->num_pairs, K = a->num_points;
  const int bins = a->num_bins, rows = a->pose_rows;
  SR_REQUIRE(rows == 3 || rows == 4, "sr_reproj_cdf_loss_f32: extrinsics must be [N, 3, 4] or [N, 4, 4]");
  SR_REQUIRE(bins >= 2 && bins <= kMaxBins, "sr_reproj_cdf_loss_f32: num_bins must be in [2, 4096]");
  SR_REQUIRE(a->smooth_radius >= 0 && a->smooth_radius < bins, "sr_reproj_cdf_loss_f32: smoothing radius must be below num_bins");
  SR_REQUIRE(N > 0 && N <= 65536 && P > 0 && P <= kMaxPairs && K > 0 && K < (1 << 30), "sr_reproj_cdf_loss_f32: bad shape");
  SR_REQUIRE(a->pred_intrinsics && a->pred_extrinsics && a->src_idx && a->dst_idx && a->src_coords && a->dst_coords && a->src_depth &&
                 a->dst_depth && a->smooth_kernel && a->workspace && a->loss && a->d_intrinsics && a->d_extrinsics,
             "sr_reproj_cdf_loss_f32: null pointer");
  SR_REQUIRE(a->workspace_floats >= sr_reproj_loss_workspace_floats(N, P, K, bins), "sr_reproj_cdf_loss_f32: workspace too small");
  cudaStream_t st = as_stream(stream);
  const int nblk = static_cast<int>(blocks_for(K, kBlockPoints));
  const int64_t PK = P * K;
  float* k_rec = a->workspace;
  float* d_krec = k_rec + N * 9;
  float* pm = d_krec + N * 9;
  float* rl = pm + P * kPairStride;
  float* hist = rl + 2 * PK;                          // [2 kinds, N, bins]
  float* total = hist + 2 * N * bins;                 // [2 kinds, N]
  float* tables = total + 2 * N;                      // pmf | cdf | pdf, each [2 kinds, N, bins]
  float* partial = tables + 6 * N * bins;
  float* pair_grad = partial + P * nblk * kGradStride;
  float* pmf = tables;
  float* cdf = tables + 2 * N * bins;
  float* pdf = tables + 4 * N * bins;

  SR_CHECK_CUDA(cudaMemsetAsync(hist, 0, sizeof(float) * 2 * N * (bins + 1), st));
  loss_frame_setup_kernel<<<1, 256, 0, st>>>(a->pred_intrinsics, a->kprime_to_k, a->shared_focal, k_rec, static_cast<int>(N));
  pair_setup_frames_kernel<<<blocks_for(P, 64), 64, 0, st>>>(k_rec, a->pred_extrinsics, rows, a->src_idx, a->dst_idx, pm, P);
  dim3 grid(nblk, static_cast<unsigned>(P));
  loss_residual_hist_kernel<<<grid, kPointThreads, sizeof(float) * 2 * bins, st>>>(
      a->src_coords, a->dst_coords, a->src_depth, a->dst_depth, pm, a->src_idx, a->dst_idx, rl, hist, total, static_cast<int>(K), PK,
      static_cast<int>(N), a->min_val, a->inv_bin_width, bins);
  cdf_tables_kernel<<<static_cast<unsigned>(2 * N), 256, sizeof(float) * 2 * bins, st>>>(hist, total, pmf, cdf, pdf, bins,
                                                                                          a->sobel_weight, a->smooth_kernel,
                                                                                          a->smooth_radius);
  loss_lookup_bwd_kernel<<<grid, kPointThreads, 0, st>>>(a->src_coords, a->dst_coords, a->src_depth, a->dst_depth, pm, a->src_idx,
                                                         a->dst_idx, rl, cdf, pdf, partial, static_cast<int>(K), PK,
                                                         static_cast<int>(N), a->min_val, a->inv_bin_width, bins);
  loss_pair_bwd_kernel<<<blocks_for(P, 64), 64, 0, st>>>(pm, partial, nblk, a->pred_extrinsics, rows, a->src_idx, a->dst_idx,
                                                          pair_grad, P);
  const float scale = 1.0f / (4.0f * static_cast<float>(PK));        // (mean + mean) / 2, averaged over the two kinds
  loss_frame_bwd_kernel<<<1, 256, 0, st>>>(pair_grad, a->src_idx, a->dst_idx, a->kprime_to_k, a->shared_focal, rows, scale, a->loss,
                                           a->d_intrinsics, a->d_extrinsics, d_krec, static_cast<int>(N), P);
  g_launch_count += 7;
  if (a->tables != nullptr)
    SR_CHECK_CUDA(cudaMemcpyAsync(a->tables, tables, sizeof(float) * 6 * N * bins, cudaMemcpyDeviceToDevice, st));
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

// ---- sr_reproject_static_f32: the forward reprojection for callers that may not allocate device memory ----------------------
// sr_reproject_f32 needs kPairStride floats of scratch per pair (the per-pair camera record the set-up kernel writes and the
// point kernel reads) and expects the caller to bring it.  Some hosts can't: a plugin running inside somebody else's
// engine whose memory pool is closed to it, a real-time localisation loop that promised "no cudaMalloc after start-up", a
// process that only borrows the buffers its parent hands over.  For them the scratch lives here, inside the loaded library:
// one fixed array per GPU, resident from the moment the module is loaded on that GPU until the process exits.
//
// The first call on each GPU (made with that GPU current) asks the runtime where the array sits on it and remembers the
// answer; make that first call during start-up, ahead of any stream capture.  There is one array per GPU and every call
// writes it from record 0, so two calls that can be in flight on the same GPU at once (different streams, different threads)
// overwrite each other's records: keep them on one stream or order them with events.
//
// kStaticMaxPairs is the price/limit trade: 4096 pairs * 32 floats = 512 KiB per GPU.  Larger batches are refused, not split.
constexpr int64_t kStaticMaxPairs = 4096;
constexpr int kStaticMaxDevices = 64;
static __device__ __align__(256) float g_static_pair_records[kStaticMaxPairs * sr::kPairStride];

namespace {
// In host code `g_static_pair_records` is only a handle for the runtime to look the symbol up with; the address the
// kernels must be given is whatever cudaGetSymbolAddress reports for the device that is current, and each device has its own.
float* g_static_pair_ptr[kStaticMaxDevices];   // written once per device, under g_static_mutex
std::mutex g_static_mutex;

int static_pair_records(float** out) {
  int dev = -1;
  SR_CHECK_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kStaticMaxDevices) return sr::fail(SR_ERR_SHAPE, "sr_reproject_static_f32: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_static_mutex);
  if (g_static_pair_ptr[dev] == nullptr) {
    void* p = nullptr;
    SR_CHECK_CUDA(cudaGetSymbolAddress(&p, g_static_pair_records));
    if (p == nullptr) return sr::fail(SR_ERR_CUDA, "sr_reproject_static_f32: cudaGetSymbolAddress returned a null address");
    g_static_pair_ptr[dev] = static_cast<float*>(p);
  }
  *out = g_static_pair_ptr[dev];
  return SR_OK;
}
}  // namespace

extern "C" int64_t sr_reproject_static_max_pairs(void) { return kStaticMaxPairs; }

extern "C" int sr_reproject_static_f32(const float* src_coords, const float* src_depth, const float* dst_depth, const float* src_k,
                                       const float* dst_k, const float* relative, int32_t rows, const float* src_scale,
                                       const float* dst_scale, float* out_coords, int64_t num_pairs, int64_t num_points,
                                       void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(rows == 3 || rows == 4, "sr_reproject_static_f32: relative pose must be [P, 3, 4] or [P, 4, 4]");
  SR_REQUIRE(num_pairs >= 0 && num_points >= 0 && num_points < (1 << 30), "sr_reproject_static_f32: bad shape");
  SR_REQUIRE(num_pairs <= kStaticMaxPairs,
             "sr_reproject_static_f32: %lld pairs exceed the built-in scratch (%lld pairs); split the batch or call "
             "sr_reproject_f32 with a caller-owned workspace",
             static_cast<long long>(num_pairs), static_cast<long long>(kStaticMaxPairs));
  if (num_pairs == 0 || num_points == 0) return SR_OK;
  SR_REQUIRE(src_coords && src_depth && src_k && dst_k && relative && out_coords, "sr_reproject_static_f32: null pointer");
  float* records = nullptr;
  if (int rc = static_pair_records(&records)) return rc;
  cudaStream_t st = as_stream(stream);
  // Nothing to clear: the set-up kernel rewrites all kPairStride floats of every record [0, num_pairs) and the point kernel
  // reads only those, so what an earlier call left behind is never seen.
  pair_setup_pairs_kernel<<<blocks_for(num_pairs, 64), 64, 0, st>>>(src_k, dst_k, relative, rows, records, num_pairs);
  dim3 grid(blocks_for(num_points, kPointThreads), static_cast<unsigned>(num_pairs));
  reproject_kernel<<<grid, kPointThreads, 0, st>>>(src_coords, src_depth, dst_depth, src_scale, dst_scale, records, out_coords,
                                                   static_cast<int>(num_points));
  g_launch_count += 2;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}
