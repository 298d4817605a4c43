// fp32 SIMT GEMM with the same epilogues as gemm_tc.cu but no intermediate rounding.
// Used (a) by the camera head, which the reference runs with autocast disabled
// (sailrecon/models/sail_recon.py:119-121; camera_head.py: Linear layers at :69,:73, trunk Blocks, pose_branch)
// and (b) by the fp32 check mode of the aggregator blocks (north_star: 1e-4 relative L2).
// Arbitrary M, N, K (pose_branch.fc2 has N = 9, embed_pose has K = 9).
//
// 64x64 output tile, BK = 16, 256 threads, 4x4 register micro-tile, smem tiles stored K-major-transposed.
#include "common.cuh"
#include "sr_internal.h"

namespace sr {
extern int g_launch_count;

constexpr int kFBM = 64, kFBN = 64, kFBK = 16;

struct GemmF32Params {
  const float* A;
  const float* W;
  int64_t lda, ldw;
  int M, N, K;
  sr_epilogue_t e;
};

template <int EPI>
__global__ void __launch_bounds__(256)
gemm_f32_kernel(const GemmF32Params p) {
  __shared__ float As[kFBK][kFBM + 4];
  __shared__ float Ws[kFBK][kFBN + 4];
  const int tid = threadIdx.x;
  const int m0 = blockIdx.y * kFBM, n0 = blockIdx.x * kFBN;
  const int tx = tid & 15, ty = tid >> 4;      // 16 x 16 threads, each 4 (m) x 4 (n)
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  const int lk = tid & 15;       // k index inside the tile
  const int lr = tid >> 4;       // row 0..15 (+16 per pass)
  for (int k0 = 0; k0 < p.K; k0 += kFBK) {
#pragma unroll
    for (int pass = 0; pass < 4; ++pass) {
      const int r = lr + pass * 16;
      const int gk = k0 + lk;
      const int gm = m0 + r, gn = n0 + r;
      As[lk][r] = (gm < p.M && gk < p.K) ? __ldg(p.A + static_cast<int64_t>(gm) * p.lda + gk) : 0.f;
      Ws[lk][r] = (gn < p.N && gk < p.K) ? __ldg(p.W + static_cast<int64_t>(gn) * p.ldw + gk) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < kFBK; ++k) {
      const float4 a = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
      const float4 w = *reinterpret_cast<const float4*>(&Ws[k][tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w}, wv[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], wv[j], acc[i][j]);
    }
    __syncthreads();
  }

  const sr_epilogue_t& e = p.e;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int gm = m0 + ty * 4 + i;
    if (gm >= p.M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int gn = n0 + tx * 4 + j;
      if (gn >= p.N) continue;
      float y = acc[i][j] + (e.bias ? __ldg(e.bias + gn) : 0.f);
      if (EPI == SR_EPI_BIAS_GELU) y = gelu_erf(y);
      if (EPI == SR_EPI_BIAS_SILU) y = y / (1.f + expf(-y));
      if (EPI == SR_EPI_LS_RESIDUAL) {
        const float g = e.gamma ? __ldg(e.gamma + gn) : 1.f;
        y = e.resid[static_cast<int64_t>(gm) * e.ldr + gn] + g * y;
      }
      reinterpret_cast<float*>(e.out)[static_cast<int64_t>(gm) * e.ldo + gn] = y;
    }
  }
}

// In-place q/k LayerNorm(d) + 2D RoPE on an fp32 [M, 3C] qkv buffer (check mode; attention.py:78-82).
// One warp per (token, head, q|k); head_dim 64: lane holds features lane and lane+32.
__global__ void __launch_bounds__(256)
qknorm_rope_f32_kernel(float* __restrict__ qkv, int64_t ld, int64_t M, int C, int qk_norm, const float* qn_w,
                       const float* qn_b, const float* kn_w, const float* kn_b, float eps, int rope,
                       const float* __restrict__ cos_t, const float* __restrict__ sin_t, const int32_t* __restrict__ pos) {
  const int lane = threadIdx.x & 31;
  const int heads = C / 64;
  const int64_t w = static_cast<int64_t>(blockIdx.x) * 8 + (threadIdx.x >> 5);
  const int64_t total = M * heads * 2;
  if (w >= total) return;
  const int which = static_cast<int>(w % 2);
  const int head = static_cast<int>((w / 2) % heads);
  const int64_t row = w / (2 * heads);
  float* x = qkv + row * ld + which * C + head * 64;
  float a = x[lane], b = x[lane + 32];
  if (qk_norm) {
    const float mean = warp_sum(a + b) * (1.f / 64.f);
    const float da = a - mean, db = b - mean;
    const float rstd = rsqrtf(warp_sum(da * da + db * db) * (1.f / 64.f) + eps);
    const float* nw = which == 0 ? qn_w : kn_w;
    const float* nb = which == 0 ? qn_b : kn_b;
    a = da * rstd * nw[lane] + nb[lane];
    b = db * rstd * nw[lane + 32] + nb[lane + 32];
  }
  if (rope) {
    // a is feature `lane` of the y half, b feature `lane` of the x half; partner inside a half is lane ^ 16
    const int py = pos[row * 2], px = pos[row * 2 + 1];
    const int f = lane & 15;
    const float cy = cos_t[py * 16 + f], sy = sin_t[py * 16 + f];
    const float cx = cos_t[px * 16 + f], sx = sin_t[px * 16 + f];
    const float pa = __shfl_xor_sync(0xffffffffu, a, 16), pb = __shfl_xor_sync(0xffffffffu, b, 16);
    const float sgn = lane < 16 ? -1.f : 1.f;      // rotate_half: first 16 get -x2, last 16 get +x1
    a = a * cy + sgn * pa * sy;
    b = b * cx + sgn * pb * sx;
  }
  x[lane] = a;
  x[lane + 32] = b;
}

}  // namespace sr

extern "C" int sr_gemm_f32(const float* A, int64_t lda, const float* W, int64_t ldw, int64_t M, int64_t N, int64_t K,
                           const sr_epilogue_t* epi, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(A && W && epi && epi->out, "sr_gemm_f32: null pointer");
  SR_REQUIRE(M > 0 && N > 0 && K > 0 && M < (1ll << 31) && N < (1ll << 31) && K < (1ll << 31), "sr_gemm_f32: bad sizes");
  SR_REQUIRE(lda >= K && ldw >= K && epi->ldo >= N, "sr_gemm_f32: leading dimensions too small");
  SR_REQUIRE(!epi->out_bf16, "sr_gemm_f32: output is fp32");
  if (epi->kind == SR_EPI_LS_RESIDUAL) SR_REQUIRE(epi->resid && epi->ldr >= N, "sr_gemm_f32: LS_RESIDUAL needs resid");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  GemmF32Params p{A, W, lda, ldw, static_cast<int>(M), static_cast<int>(N), static_cast<int>(K), *epi};
  dim3 grid(static_cast<unsigned>((N + kFBN - 1) / kFBN), static_cast<unsigned>((M + kFBM - 1) / kFBM));
  SR_REQUIRE(grid.y <= 65535, "sr_gemm_f32: M too large for this path");
  const int kind = epi->kind == SR_EPI_QKV ? SR_EPI_BIAS : epi->kind;
  switch (kind) {
    case SR_EPI_BIAS: gemm_f32_kernel<SR_EPI_BIAS><<<grid, 256, 0, st>>>(p); break;
    case SR_EPI_BIAS_GELU: gemm_f32_kernel<SR_EPI_BIAS_GELU><<<grid, 256, 0, st>>>(p); break;
    case SR_EPI_LS_RESIDUAL: gemm_f32_kernel<SR_EPI_LS_RESIDUAL><<<grid, 256, 0, st>>>(p); break;
    case SR_EPI_BIAS_SILU: gemm_f32_kernel<SR_EPI_BIAS_SILU><<<grid, 256, 0, st>>>(p); break;
    default: return fail(SR_ERR_SHAPE, "sr_gemm_f32: unknown epilogue kind %d", epi->kind);
  }
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  if (epi->kind == SR_EPI_QKV && (epi->qk_norm || epi->rope)) {
    const sr_epilogue_t& e = *epi;
    SR_REQUIRE(e.head_dim == 64 && e.col_offset == 0 && N == 3 * e.qkv_dim, "sr_gemm_f32: QKV epilogue needs the full [3C] output, head_dim 64");
    if (e.rope) SR_REQUIRE(e.rope_cos && e.rope_sin && e.pos, "sr_gemm_f32: rope needs tables and positions");
    const int64_t warps = M * (e.qkv_dim / 64) * 2;
    qknorm_rope_f32_kernel<<<static_cast<unsigned>((warps + 7) / 8), 256, 0, st>>>(
        reinterpret_cast<float*>(e.out), e.ldo, M, static_cast<int>(e.qkv_dim), e.qk_norm, e.qn_w, e.qn_b, e.kn_w, e.kn_b,
        e.qk_eps, e.rope, e.rope_cos, e.rope_sin, e.pos);
    ++g_launch_count;
    SR_CHECK_CUDA(cudaGetLastError());
  }
  return SR_OK;
}
