// Row LayerNorm (HBM-bound): fp32 in -> bf16 (feeds the next tensor-core GEMM) or fp32 out.
// Replaces ATen layer_norm at sailrecon/layers/block.py:50,70 (norm1/norm2) and
// sailrecon/heads/camera_head.py:64,65,77 (token_norm, trunk_norm, adaln_norm), and the adaLN
// modulate + gate + residual chain of camera_head.py:151-160.
//
// One warp per row; the row lives in registers (float4 per lane, coalesced 512 B per warp per step);
// mean and variance are two-pass in fp32 with warp-shuffle reductions.  Algorithmic bytes per row:
// C*4 read + C*2 (bf16) or C*4 (fp32) written.
#include "common.cuh"
#include "sr_internal.h"

namespace sr {
extern int g_launch_count;

constexpr int kLnWarps = 8;

template <int NV, bool OUT_BF16>
__global__ void __launch_bounds__(kLnWarps * 32)
layernorm_kernel(const float* __restrict__ x, int64_t ldx, const int32_t* __restrict__ idx, const float* __restrict__ w,
                 const float* __restrict__ b, void* __restrict__ out, int64_t ldo, int64_t rows, int C, float eps) {
  const int lane = threadIdx.x & 31;
  const int64_t row = static_cast<int64_t>(blockIdx.x) * kLnWarps + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int nvec = C >> 2;
  const int64_t src_row = idx ? static_cast<int64_t>(__ldg(idx + row)) : row;   // fused row gather (scene tokens)
  const float4* xr = reinterpret_cast<const float4*>(x + src_row * ldx);
  float4 v[NV];
  float sum = 0.f;
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    const int i = lane + 32 * j;
    v[j] = i < nvec ? __ldg(xr + i) : make_float4(0.f, 0.f, 0.f, 0.f);
    sum += (v[j].x + v[j].y) + (v[j].z + v[j].w);
  }
  const float mean = warp_sum(sum) / static_cast<float>(C);
  float sq = 0.f;
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    const int i = lane + 32 * j;
    if (i < nvec) {
      float a = v[j].x - mean, bb = v[j].y - mean, c = v[j].z - mean, d = v[j].w - mean;
      sq += (a * a + bb * bb) + (c * c + d * d);
    }
  }
  const float rstd = rsqrtf(warp_sum(sq) / static_cast<float>(C) + eps);
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    const int i = lane + 32 * j;
    if (i >= nvec) continue;
    float4 g = w ? __ldg(reinterpret_cast<const float4*>(w) + i) : make_float4(1.f, 1.f, 1.f, 1.f);
    float4 s = b ? __ldg(reinterpret_cast<const float4*>(b) + i) : make_float4(0.f, 0.f, 0.f, 0.f);
    float4 y;
    y.x = (v[j].x - mean) * rstd * g.x + s.x;
    y.y = (v[j].y - mean) * rstd * g.y + s.y;
    y.z = (v[j].z - mean) * rstd * g.z + s.z;
    y.w = (v[j].w - mean) * rstd * g.w + s.w;
    if (OUT_BF16) {
      uint2 pk = make_uint2(pack_bf16x2(y.x, y.y), pack_bf16x2(y.z, y.w));
      reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(out) + row * ldo)[i] = pk;
    } else {
      reinterpret_cast<float4*>(reinterpret_cast<float*>(out) + row * ldo)[i] = y;
    }
  }
}

// camera_head.py:151-160: out = gate * (LN_noaffine(x) * (1 + scale) + shift) + x ; mod = [shift | scale | gate]
template <int NV>
__global__ void __launch_bounds__(kLnWarps * 32)
adaln_modulate_kernel(const float* __restrict__ x, const float* __restrict__ mod, float* __restrict__ out,
                      int64_t rows, int C, float eps) {
  const int lane = threadIdx.x & 31;
  const int64_t row = static_cast<int64_t>(blockIdx.x) * kLnWarps + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int nvec = C >> 2;
  const float4* xr = reinterpret_cast<const float4*>(x + row * C);
  float4 v[NV];
  float sum = 0.f;
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    const int i = lane + 32 * j;
    v[j] = i < nvec ? __ldg(xr + i) : make_float4(0.f, 0.f, 0.f, 0.f);
    sum += (v[j].x + v[j].y) + (v[j].z + v[j].w);
  }
  const float mean = warp_sum(sum) / static_cast<float>(C);
  float sq = 0.f;
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    const int i = lane + 32 * j;
    if (i < nvec) {
      float a = v[j].x - mean, bb = v[j].y - mean, c = v[j].z - mean, d = v[j].w - mean;
      sq += (a * a + bb * bb) + (c * c + d * d);
    }
  }
  const float rstd = rsqrtf(warp_sum(sq) / static_cast<float>(C) + eps);
  const float4* sh = reinterpret_cast<const float4*>(mod + row * 3 * C);
  const float4* sc = sh + nvec;
  const float4* gt = sc + nvec;
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    const int i = lane + 32 * j;
    if (i >= nvec) continue;
    float4 s = __ldg(sh + i), c = __ldg(sc + i), g = __ldg(gt + i);
    float4 y;
    y.x = g.x * ((v[j].x - mean) * rstd * (1.f + c.x) + s.x) + v[j].x;
    y.y = g.y * ((v[j].y - mean) * rstd * (1.f + c.y) + s.y) + v[j].y;
    y.z = g.z * ((v[j].z - mean) * rstd * (1.f + c.z) + s.z) + v[j].z;
    y.w = g.w * ((v[j].w - mean) * rstd * (1.f + c.w) + s.w) + v[j].w;
    reinterpret_cast<float4*>(out + row * C)[i] = y;
  }
}

template <int NV>
static int launch_ln(const float* x, int64_t ldx, const int32_t* idx, const float* w, const float* b, void* out,
                     int out_bf16, int64_t ldo, int64_t rows, int C, float eps, cudaStream_t st) {
  const unsigned grid = static_cast<unsigned>((rows + kLnWarps - 1) / kLnWarps);
  if (out_bf16)
    layernorm_kernel<NV, true><<<grid, kLnWarps * 32, 0, st>>>(x, ldx, idx, w, b, out, ldo, rows, C, eps);
  else
    layernorm_kernel<NV, false><<<grid, kLnWarps * 32, 0, st>>>(x, ldx, idx, w, b, out, ldo, rows, C, eps);
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

}  // namespace sr

extern "C" int sr_layernorm_gather_fwd(const float* x, int64_t ldx, const int32_t* idx, const float* weight,
                                       const float* bias, void* out, int out_bf16, int64_t ldo, int64_t rows, int64_t C,
                                       float eps, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  if (rows == 0) return SR_OK;
  SR_REQUIRE(x && out, "sr_layernorm_fwd: null pointer");
  SR_REQUIRE(C > 0 && C % 4 == 0 && C <= 4096, "sr_layernorm_fwd: C=%lld must be a multiple of 4 and <= 4096", (long long)C);
  SR_REQUIRE(ldx % 4 == 0 && ldx >= C && ldo % 4 == 0 && ldo >= C, "sr_layernorm_fwd: row strides must be >= C and multiples of 4");
  SR_REQUIRE((reinterpret_cast<uintptr_t>(x) & 15) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0,
             "sr_layernorm_fwd: pointers must be 16-byte aligned");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int c = static_cast<int>(C);
  if (C <= 128) return launch_ln<1>(x, ldx, idx, weight, bias, out, out_bf16, ldo, rows, c, eps, st);
  if (C <= 512) return launch_ln<4>(x, ldx, idx, weight, bias, out, out_bf16, ldo, rows, c, eps, st);
  if (C <= 1024) return launch_ln<8>(x, ldx, idx, weight, bias, out, out_bf16, ldo, rows, c, eps, st);
  if (C <= 2048) return launch_ln<16>(x, ldx, idx, weight, bias, out, out_bf16, ldo, rows, c, eps, st);
  return launch_ln<32>(x, ldx, idx, weight, bias, out, out_bf16, ldo, rows, c, eps, st);
}

extern "C" int sr_adaln_modulate_f32(const float* x, const float* mod, float* out, int64_t rows, int64_t C, float eps,
                                     void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  if (rows == 0) return SR_OK;
  SR_REQUIRE(x && mod && out, "sr_adaln_modulate_f32: null pointer");
  SR_REQUIRE(C > 0 && C % 4 == 0 && C <= 2048, "sr_adaln_modulate_f32: C=%lld must be a multiple of 4 and <= 2048", (long long)C);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const unsigned grid = static_cast<unsigned>((rows + kLnWarps - 1) / kLnWarps);
  const int c = static_cast<int>(C);
  if (C <= 512)
    adaln_modulate_kernel<4><<<grid, kLnWarps * 32, 0, st>>>(x, mod, out, rows, c, eps);
  else if (C <= 1024)
    adaln_modulate_kernel<8><<<grid, kLnWarps * 32, 0, st>>>(x, mod, out, rows, c, eps);
  else
    adaln_modulate_kernel<16><<<grid, kLnWarps * 32, 0, st>>>(x, mod, out, rows, c, eps);
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

extern "C" int sr_layernorm_fwd(const float* x, int64_t ldx, const float* weight, const float* bias, void* out,
                                int out_bf16, int64_t ldo, int64_t rows, int64_t C, float eps, void* stream) {
  return sr_layernorm_gather_fwd(x, ldx, nullptr, weight, bias, out, out_bf16, ldo, rows, C, eps, stream);
}
