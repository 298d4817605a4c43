// Small HBM-bound helper kernels around the block kernels:
//   sr_rope_table_build  cos/sin tables of RotaryPositionEmbedding2D (sailrecon/layers/rope.py:107-123)
//   sr_rope2d_f32        standalone 2D RoPE on a [B, H, N, d] tensor (rope.py:165-207) — the fused path applies RoPE
//                        inside the qkv GEMM epilogue; this entry point backs the module's own forward()
//   sr_patchify          ImageNet normalisation (models/aggregator.py:267) + im2col of the stride-14 patch
//                        convolution (layers/patch_embed.py:62-64,78) so the projection runs on the GEMM kernels
//   sr_pose_update       pred += delta and activate_pose (heads/camera_head.py:172-184, heads/head_act.py:12-37)
#include "common.cuh"
#include "sr_internal.h"

namespace sr {
extern int g_launch_count;

__global__ void rope_table_kernel(float* __restrict__ cos_t, float* __restrict__ sin_t, int max_pos, int quarter,
                                  float frequency) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= max_pos * quarter) return;
  const int p = i / quarter, f = i % quarter;
  // rope.py:110-111: exponents = arange(0, D, 2) / D with D = 2 * quarter; inv_freq = 1 / frequency ** exponents
  const float exponent = static_cast<float>(2 * f) / static_cast<float>(2 * quarter);
  const float inv_freq = 1.0f / powf(frequency, exponent);
  const float angle = static_cast<float>(p) * inv_freq;
  cos_t[i] = cosf(angle);
  sin_t[i] = sinf(angle);
}

__global__ void __launch_bounds__(256)
rope2d_kernel(const float* __restrict__ x, float* __restrict__ out, const int32_t* __restrict__ pos,
              const float* __restrict__ cos_t, const float* __restrict__ sin_t, int64_t total, int H, int64_t N, int d,
              int max_pos) {
  const int64_t e = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (e >= total) return;
  const int i = static_cast<int>(e % d);
  const int64_t row = e / d;                 // (b, h, n)
  const int64_t n = row % N;
  const int64_t b = row / (N * H);
  const int half_d = d >> 1, quarter = d >> 2;
  const int half = i / half_d, j = i % half_d, f = j % quarter;
  int p = pos[(b * N + n) * 2 + half];
  p = p < 0 ? 0 : (p >= max_pos ? max_pos - 1 : p);
  const float c = cos_t[p * quarter + f], s = sin_t[p * quarter + f];
  const float self = x[e];
  const float partner = j < quarter ? -x[e + quarter] : x[e - quarter];     // rotate-half, rope.py:136-138
  out[e] = self * c + partner * s;
}

template <bool OUT_BF16>
__global__ void __launch_bounds__(256)
patchify_kernel(const float* __restrict__ img, void* __restrict__ out, int64_t ldo, int S, int H, int W, int ps, int K,
                int Kpad, float m0, float m1, float m2, float sd0, float sd1, float sd2) {
  const int hp = H / ps, wp = W / ps;
  const int64_t total = static_cast<int64_t>(S) * hp * wp * Kpad;
  const int64_t e = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (e >= total) return;
  const int k = static_cast<int>(e % Kpad);
  const int64_t row = e / Kpad;
  float v = 0.f;
  if (k < K) {
    const int c = k / (ps * ps), ky = (k / ps) % ps, kx = k % ps;
    const int px = static_cast<int>(row % wp), py = static_cast<int>((row / wp) % hp);
    const int64_t s = row / (static_cast<int64_t>(wp) * hp);
    const float pix = __ldg(img + ((s * 3 + c) * H + (py * ps + ky)) * static_cast<int64_t>(W) + (px * ps + kx));
    const float mean = c == 0 ? m0 : (c == 1 ? m1 : m2);
    const float sd = c == 0 ? sd0 : (c == 1 ? sd1 : sd2);
    v = (pix - mean) / sd;
  }
  if (OUT_BF16) reinterpret_cast<__nv_bfloat16*>(out)[row * ldo + k] = __float2bfloat16_rn(v);
  else reinterpret_cast<float*>(out)[row * ldo + k] = v;
}

__global__ void pose_update_kernel(float* __restrict__ pred, const float* __restrict__ delta, float* __restrict__ act,
                                   int64_t n, int first) {
  const int64_t e = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (e >= n) return;
  const float v = first ? delta[e] : pred[e] + delta[e];
  pred[e] = v;
  act[e] = (e % 9) >= 7 ? fmaxf(v, 0.f) : v;     // trans / quat linear, FoV relu
}

__global__ void __launch_bounds__(256)
scale_cols_kernel(const float* __restrict__ x, const float* __restrict__ gamma, float* __restrict__ out, int64_t total,
                  int C) {
  const int64_t e = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (e >= total) return;
  out[e] = x[e] * __ldg(gamma + (e % C));
}

}  // namespace sr

extern "C" int sr_scale_cols_f32(const float* x, const float* gamma, float* out, int64_t rows, int64_t C, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(x && gamma && out && C > 0, "sr_scale_cols_f32: bad arguments");
  const int64_t total = rows * C;
  if (total == 0) return SR_OK;
  SR_REQUIRE((total + 255) / 256 < (1ll << 31), "sr_scale_cols_f32: tensor too large");
  scale_cols_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      x, gamma, out, total, static_cast<int>(C));
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

extern "C" int sr_rope_table_build(float* cos_out, float* sin_out, int32_t max_pos, int32_t quarter_dim, float frequency,
                                   void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(cos_out && sin_out && max_pos > 0 && quarter_dim > 0 && frequency > 0.f, "sr_rope_table_build: bad arguments");
  const int n = max_pos * quarter_dim;
  rope_table_kernel<<<(n + 255) / 256, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(cos_out, sin_out, max_pos,
                                                                                         quarter_dim, frequency);
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

extern "C" int sr_rope2d_f32(const float* x, float* out, const int32_t* pos, const float* rope_cos, const float* rope_sin,
                             int32_t max_pos, int64_t B, int64_t H, int64_t N, int64_t d, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(x && out && pos && rope_cos && rope_sin, "sr_rope2d_f32: null pointer");
  SR_REQUIRE(d > 0 && d % 4 == 0 && max_pos > 0, "sr_rope2d_f32: feature dim must be a multiple of 4");
  const int64_t total = B * H * N * d;
  if (total == 0) return SR_OK;
  SR_REQUIRE((total + 255) / 256 < (1ll << 31), "sr_rope2d_f32: tensor too large");
  rope2d_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      x, out, pos, rope_cos, rope_sin, total, static_cast<int>(H), N, static_cast<int>(d), max_pos);
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

extern "C" int sr_patchify(const float* images, void* out, int out_bf16, int64_t ldo, int64_t S, int64_t H, int64_t W,
                           int32_t patch, const float* mean3, const float* std3, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(images && out && mean3 && std3, "sr_patchify: null pointer (mean3/std3 are HOST pointers to 3 floats)");
  SR_REQUIRE(patch > 0 && H % patch == 0 && W % patch == 0 && S > 0, "sr_patchify: H, W must be multiples of the patch size");
  const int K = 3 * patch * patch;
  SR_REQUIRE(ldo >= K && ldo < (1 << 30), "sr_patchify: ldo must be >= 3*patch*patch");
  const int Kpad = static_cast<int>(ldo);
  const int64_t total = S * (H / patch) * (W / patch) * Kpad;
  SR_REQUIRE((total + 255) / 256 < (1ll << 31), "sr_patchify: too many elements");
  const unsigned grid = static_cast<unsigned>((total + 255) / 256);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (out_bf16)
    patchify_kernel<true><<<grid, 256, 0, st>>>(images, out, ldo, static_cast<int>(S), static_cast<int>(H),
                                                static_cast<int>(W), patch, K, Kpad, mean3[0], mean3[1], mean3[2], std3[0],
                                                std3[1], std3[2]);
  else
    patchify_kernel<false><<<grid, 256, 0, st>>>(images, out, ldo, static_cast<int>(S), static_cast<int>(H),
                                                 static_cast<int>(W), patch, K, Kpad, mean3[0], mean3[1], mean3[2], std3[0],
                                                 std3[1], std3[2]);
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

extern "C" int sr_pose_update(float* pred, const float* delta, float* activated, int64_t rows, int32_t first, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(pred && delta && activated, "sr_pose_update: null pointer");
  if (rows == 0) return SR_OK;
  const int64_t n = rows * 9;
  pose_update_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      pred, delta, activated, n, first);
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}
