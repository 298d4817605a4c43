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

// ---------------------------------------------------------------------------------------------- skinny path
// Few rows (the camera head: M = Na + Nq camera tokens, tens of rows) against big weights (216 M fp32 parameters,
// camera_head.py:59-83): the work is streaming W once from HBM.  Warp-per-output-column: a warp owns 4 rows of W, its
// lanes split K in float4s (fully coalesced 512 B per row per step), A[32 rows, 256 k] is staged in shared memory by
// cp.async (double buffered) and every A float4 read from smem feeds 16 FMAs.  The 4 x 32 per-lane partial sums are
// reduced with a 31-shuffle transpose-reduce so that lane l ends up owning row m0 + l.  Grid = N/16 x ceil(M/32) CTAs
// (128 ... 512 CTAs for the camera-head shapes) instead of the 32 ... 128 CTAs of the 64x64 tiling above.
constexpr int kSkM = 32, kSkKC = 256, kSkWarps = 4, kSkNPerWarp = 4;   // 16 FMAs per LDS.128: shared-memory reads no longer bound the loop
constexpr int kSkSmem = 2 * kSkM * kSkKC * 4;

__device__ __forceinline__ void cp_async16_zfill(void* smem_dst, const void* gsrc, bool valid) {
  const int sz = valid ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// lane l of the warp returns sum over lanes of v[l]  (v is consumed)
__device__ __forceinline__ float warp_transpose_reduce32(float (&v)[32], int lane) {
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    const bool hi = (lane & off) != 0;
#pragma unroll
    for (int i = 0; i < off; ++i) {
      const float send = hi ? v[i] : v[i + off];
      const float keep = hi ? v[i + off] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
    }
  }
  return v[0];
}

template <int EPI>
__global__ void __launch_bounds__(kSkWarps * 32, 2)
gemm_f32_skinny_kernel(const GemmF32Params p) {
  extern __shared__ __align__(16) uint8_t sk_smem[];
  float* As = reinterpret_cast<float*>(sk_smem);            // [2][kSkM][kSkKC]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int m0 = blockIdx.y * kSkM;
  const int n_base = (blockIdx.x * kSkWarps + warp) * kSkNPerWarp;
  const int num_kc = (p.K + kSkKC - 1) / kSkKC;

  auto stage_a = [&](int kc, int buf) {
    // 32 rows x 64 float4 = 2048 float4 over the CTA's threads; a warp covers 512 contiguous bytes of one row
#pragma unroll
    for (int i = 0; i < kSkM * (kSkKC / 4) / (kSkWarps * 32); ++i) {
      const int f = tid + i * kSkWarps * 32;
      const int r = f >> 6, c4 = f & 63;
      const int gm = m0 + r, gk = kc * kSkKC + c4 * 4;
      const bool ok = gm < p.M && gk < p.K;
      const float* src = ok ? p.A + static_cast<int64_t>(gm) * p.lda + gk : p.A;
      cp_async16_zfill(As + (buf * kSkM + r) * kSkKC + c4 * 4, src, ok);
    }
    cp_async_commit();
  };
  const float* wrow[kSkNPerWarp];
  bool wok[kSkNPerWarp];
#pragma unroll
  for (int j = 0; j < kSkNPerWarp; ++j) {
    wok[j] = n_base + j < p.N;
    wrow[j] = p.W + static_cast<int64_t>(wok[j] ? n_base + j : 0) * p.ldw;
  }
  auto load_w = [&](int kc, float4 (&w)[kSkNPerWarp][2]) {
#pragma unroll
    for (int j = 0; j < kSkNPerWarp; ++j)
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int gk = kc * kSkKC + h * 128 + lane * 4;
        w[j][h] = (wok[j] && gk < p.K) ? __ldg(reinterpret_cast<const float4*>(wrow[j] + gk)) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
  };

  float acc[kSkNPerWarp][kSkM];
#pragma unroll
  for (int j = 0; j < kSkNPerWarp; ++j)
#pragma unroll
    for (int m = 0; m < kSkM; ++m) acc[j][m] = 0.f;

  float4 wcur[kSkNPerWarp][2], wnext[kSkNPerWarp][2];
  stage_a(0, 0);
  load_w(0, wcur);
  for (int kc = 0; kc < num_kc; ++kc) {
    const int buf = kc & 1;
    cp_async_wait<0>();
    __syncthreads();                       // chunk kc landed for everyone; everyone finished reading buffer buf^1
    if (kc + 1 < num_kc) {
      stage_a(kc + 1, buf ^ 1);
      load_w(kc + 1, wnext);
    }
    const float* a_base = As + buf * kSkM * kSkKC + lane * 4;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
#pragma unroll
      for (int m = 0; m < kSkM; ++m) {
        const float4 a = *reinterpret_cast<const float4*>(a_base + m * kSkKC + h * 128);
#pragma unroll
        for (int j = 0; j < kSkNPerWarp; ++j) {
          acc[j][m] = fmaf(a.x, wcur[j][h].x, acc[j][m]);
          acc[j][m] = fmaf(a.y, wcur[j][h].y, acc[j][m]);
          acc[j][m] = fmaf(a.z, wcur[j][h].z, acc[j][m]);
          acc[j][m] = fmaf(a.w, wcur[j][h].w, acc[j][m]);
        }
      }
    }
    if (kc + 1 < num_kc) {
#pragma unroll
      for (int j = 0; j < kSkNPerWarp; ++j)
#pragma unroll
        for (int h = 0; h < 2; ++h) wcur[j][h] = wnext[j][h];
    }
  }

  const sr_epilogue_t& e = p.e;
  const int gm = m0 + lane;
#pragma unroll
  for (int j = 0; j < kSkNPerWarp; ++j) {
    const float s = warp_transpose_reduce32(acc[j], lane);
    const int gn = n_base + j;
    if (gn >= p.N || gm >= p.M) continue;
    float y = s + (e.bias ? __ldg(e.bias + gn) : 0.f);
    if (EPI == SR_EPI_BIAS_GELU) y = gelu_erf(y);
    if (EPI == SR_EPI_BIAS_SILU) y = y / (1.f + expf(-y));
    if (EPI == SR_EPI_LS_RESIDUAL) {
      const float g = e.gamma ? __ldg(e.gamma + gn) : 1.f;
      y = e.resid[static_cast<int64_t>(gm) * e.ldr + gn] + g * y;
    }
    reinterpret_cast<float*>(e.out)[static_cast<int64_t>(gm) * e.ldo + gn] = y;
  }
}

template <int EPI>
static int launch_skinny(const GemmF32Params& p, cudaStream_t st) {
  auto kern = gemm_f32_skinny_kernel<EPI>;
  static bool configured = false;
  if (!configured) {
    SR_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSkSmem));
    configured = true;
  }
  dim3 grid(static_cast<unsigned>((p.N + kSkWarps * kSkNPerWarp - 1) / (kSkWarps * kSkNPerWarp)),
            static_cast<unsigned>((p.M + kSkM - 1) / kSkM));
  kern<<<grid, kSkWarps * 32, kSkSmem, st>>>(p);
  return SR_OK;
}

// ---------------------------------------------------------------------------------------------- tiny-M path (M <= 8)
// One to eight rows against the same big weights: the per-query-image camera head of the localisation workload
// (localize.Localizer: the anchors' rows are cached, only the query rows run — M = 1 per image) and any M <= 8 call.
// Pure weight streaming: a warp owns kTyCols rows of W, lanes split K in float4s (512 contiguous bytes per row and step,
// kTyUnroll steps in flight), the M rows of A come through L1 (they are re-read by every warp and total <= 256 KB), partial
// sums are reduced with shuffles.  Grid = N / (kTyCols * kTyWarps) CTAs = 256 ... 1024 for the camera-head shapes.
constexpr int kTyWarps = 4, kTyUnroll = 4;   // kTyCols (W rows per warp): 2, or 1 when 2 would leave the grid under ~4 CTAs per SM

template <int EPI, int MR, int kTyCols>
__global__ void __launch_bounds__(kTyWarps * 32)
gemm_f32_tiny_kernel(const GemmF32Params p) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int n_base = (blockIdx.x * kTyWarps + warp) * kTyCols;
  if (n_base >= p.N) return;
  const float* wrow[kTyCols];
#pragma unroll
  for (int j = 0; j < kTyCols; ++j) wrow[j] = p.W + static_cast<int64_t>(min(n_base + j, p.N - 1)) * p.ldw;
  float acc[kTyCols][MR];
#pragma unroll
  for (int j = 0; j < kTyCols; ++j)
#pragma unroll
    for (int m = 0; m < MR; ++m) acc[j][m] = 0.f;
  const int k4 = p.K >> 2;                                   // float4 columns (K % 4 == 0 checked by the caller)
  for (int c0 = lane; c0 < k4; c0 += 32 * kTyUnroll) {
    float4 w[kTyUnroll][kTyCols];
#pragma unroll
    for (int u = 0; u < kTyUnroll; ++u) {
      const int c = c0 + u * 32;
#pragma unroll
      for (int j = 0; j < kTyCols; ++j)
        w[u][j] = c < k4 ? __ldcs(reinterpret_cast<const float4*>(wrow[j]) + c) : make_float4(0.f, 0.f, 0.f, 0.f);   // streamed once
    }
#pragma unroll
    for (int u = 0; u < kTyUnroll; ++u) {
      const int c = c0 + u * 32;
      if (c >= k4) break;
#pragma unroll
      for (int m = 0; m < MR; ++m) {
        const float4 a = m < p.M ? __ldg(reinterpret_cast<const float4*>(p.A + static_cast<int64_t>(m) * p.lda) + c)
                                 : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int j = 0; j < kTyCols; ++j)
          acc[j][m] = fmaf(a.w, w[u][j].w, fmaf(a.z, w[u][j].z, fmaf(a.y, w[u][j].y, fmaf(a.x, w[u][j].x, acc[j][m]))));
      }
    }
  }
  const sr_epilogue_t& e = p.e;
#pragma unroll
  for (int j = 0; j < kTyCols; ++j)
#pragma unroll
    for (int m = 0; m < MR; ++m) {
      const float s = warp_sum(acc[j][m]);
      const int gn = n_base + j;
      if (lane != 0 || gn >= p.N || m >= p.M) continue;
      float y = s + (e.bias ? __ldg(e.bias + gn) : 0.f);
      if (EPI == SR_EPI_BIAS_GELU) y = gelu_erf(y);
      if (EPI == SR_EPI_BIAS_SILU) y = y / (1.f + expf(-y));
      if (EPI == SR_EPI_LS_RESIDUAL) {
        const float g = e.gamma ? __ldg(e.gamma + gn) : 1.f;
        y = e.resid[static_cast<int64_t>(m) * e.ldr + gn] + g * y;
      }
      reinterpret_cast<float*>(e.out)[static_cast<int64_t>(m) * e.ldo + gn] = y;
    }
}

template <int EPI, int kTyCols>
static int launch_tiny_c(const GemmF32Params& p, cudaStream_t st) {
  const unsigned grid = static_cast<unsigned>((p.N + kTyCols * kTyWarps - 1) / (kTyCols * kTyWarps));
  if (p.M == 1) gemm_f32_tiny_kernel<EPI, 1, kTyCols><<<grid, kTyWarps * 32, 0, st>>>(p);
  else if (p.M == 2) gemm_f32_tiny_kernel<EPI, 2, kTyCols><<<grid, kTyWarps * 32, 0, st>>>(p);
  else if (p.M <= 4) gemm_f32_tiny_kernel<EPI, 4, kTyCols><<<grid, kTyWarps * 32, 0, st>>>(p);
  else gemm_f32_tiny_kernel<EPI, 8, kTyCols><<<grid, kTyWarps * 32, 0, st>>>(p);
  return SR_OK;
}
template <int EPI>
static int launch_tiny(const GemmF32Params& p, cudaStream_t st) {
  // N = 2048 with two columns per warp is 256 CTAs: too few bytes in flight to fill HBM (ncu: 2.0 TB/s against 3.3 TB/s at 1024 CTAs)
  return p.N <= 4096 ? launch_tiny_c<EPI, 1>(p, st) : launch_tiny_c<EPI, 2>(p, st);
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
  // few rows against a big weight matrix (camera head): stream W once, warp-per-output-column
  const bool skinny = M <= 128 && K >= 256 && K % 4 == 0 && lda % 4 == 0 && ldw % 4 == 0 &&
                      ((reinterpret_cast<uintptr_t>(A) | reinterpret_cast<uintptr_t>(W)) & 15) == 0;
  const bool tiny = skinny && M <= 8;
  if (tiny) {
    switch (kind) {
      case SR_EPI_BIAS: launch_tiny<SR_EPI_BIAS>(p, st); break;
      case SR_EPI_BIAS_GELU: launch_tiny<SR_EPI_BIAS_GELU>(p, st); break;
      case SR_EPI_LS_RESIDUAL: launch_tiny<SR_EPI_LS_RESIDUAL>(p, st); break;
      case SR_EPI_BIAS_SILU: launch_tiny<SR_EPI_BIAS_SILU>(p, st); break;
      default: return fail(SR_ERR_SHAPE, "sr_gemm_f32: unknown epilogue kind %d", epi->kind);
    }
  } else if (skinny) {
    int rc = SR_OK;
    switch (kind) {
      case SR_EPI_BIAS: rc = launch_skinny<SR_EPI_BIAS>(p, st); break;
      case SR_EPI_BIAS_GELU: rc = launch_skinny<SR_EPI_BIAS_GELU>(p, st); break;
      case SR_EPI_LS_RESIDUAL: rc = launch_skinny<SR_EPI_LS_RESIDUAL>(p, st); break;
      case SR_EPI_BIAS_SILU: rc = launch_skinny<SR_EPI_BIAS_SILU>(p, st); break;
      default: return fail(SR_ERR_SHAPE, "sr_gemm_f32: unknown epilogue kind %d", epi->kind);
    }
    if (rc) return rc;
  } else {
    switch (kind) {
      case SR_EPI_BIAS: gemm_f32_kernel<SR_EPI_BIAS><<<grid, 256, 0, st>>>(p); break;
      case SR_EPI_BIAS_GELU: gemm_f32_kernel<SR_EPI_BIAS_GELU><<<grid, 256, 0, st>>>(p); break;
      case SR_EPI_LS_RESIDUAL: gemm_f32_kernel<SR_EPI_LS_RESIDUAL><<<grid, 256, 0, st>>>(p); break;
      case SR_EPI_BIAS_SILU: gemm_f32_kernel<SR_EPI_BIAS_SILU><<<grid, 256, 0, st>>>(p); break;
      default: return fail(SR_ERR_SHAPE, "sr_gemm_f32: unknown epilogue kind %d", epi->kind);
    }
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
