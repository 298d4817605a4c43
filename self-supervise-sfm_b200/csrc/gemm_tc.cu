// bf16 GEMM on tcgen05 tensor cores with fused epilogues (sm_100a).
//
//   out[M,N] = epilogue(A[M,K] @ W[N,K]^T + bias)
//
// Replaces the cuBLASLt calls behind nn.Linear at sailrecon/layers/attention.py:73 (qkv), :120 (proj) and
// sailrecon/layers/mlp.py:35,38 (fc1, fc2) and folds in what the reference runs as separate ATen kernels:
// bias, exact-erf GELU (mlp.py:36), LayerScale*x + residual (layer_scale.py:23, block.py:110-111) and, for
// the qkv projection, the per-head LayerNorm on q,k (attention.py:78) plus 2D RoPE (rope.py:165-207).
//
// Design (one persistent CTA per SM, 384 threads):
//   warp 0      TMA producer: A tile 128x64 and W tile BNx64 (SWIZZLE_128B) into a 4-stage smem ring
//   warp 1      MMA issuer: one thread issues tcgen05.mma (M=128, N=BN, K=16) x4 per stage; accumulators
//               live in TMEM, double buffered (2 x BN columns) so the epilogue of tile i overlaps tile i+1
//   warp 2      TMEM allocator
//   warps 4-11  epilogue: tcgen05.ld gives each thread one output row x 32 columns; warps w and w+4 share a
//               lane quarter and split the BN columns.  A thread holding a whole 64-wide head makes the
//               q/k LayerNorm + RoPE a purely in-register computation (no shuffles).
// Rounding points reproduce CUDA autocast (SURVEY.md §5 "precision policy"): acc+bias is rounded to bf16
// where the reference's Linear returns bf16, before GELU / LayerScale / LayerNorm consume it.
#include "common.cuh"
#include "sr_internal.h"

namespace sr {

constexpr int kBM = 128;
constexpr int kBK = 64;
constexpr int kGemmThreads = 384;
constexpr int kEpiWarp0 = 4;
constexpr int kNumEpiWarps = 8;

struct GemmParams {
  int M, N, K;
  sr_epilogue_t e;
  int conv_c, conv_wp;     // AMN == 2 (3x3 convolution as a shifted GEMM): channels per pixel, padded row width W + 2
};

template <int BN>
struct GemmCfg {
  static constexpr int kStages = (BN == 256) ? 4 : 6;
  static constexpr int kABytes = kBM * kBK * 2;
  static constexpr int kBBytes = BN * kBK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kEpiStageBytes = 8 * 4096;                  // 4 KB transpose buffer per epilogue warp (LS_RESIDUAL)
  static constexpr int kSmemBytes = kStages * kStageBytes + kEpiStageBytes + 1024;  // +1024: manual 1 KB alignment
  static constexpr int kTmemCols = 2 * BN;                         // double-buffered accumulator
};

// Exact-erf GELU (mlp.py:36, nn.GELU()) for the tensor-core epilogue: erf by Abramowitz-Stegun 7.1.26
// (|error| <= 1.5e-7, four orders below the bf16 rounding that follows) = 2 MUFU + ~13 FMA-pipe instructions per value
// instead of the ~27 of erff(); the fp32 check path (gemm_f32.cu) keeps erff().
__device__ __forceinline__ float gelu_erf_fast(float x) {
  const float z = fabsf(x) * 0.70710678118654752440f;
  float t;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(0.3275911f, z, 1.0f)));
  float poly = fmaf(1.061405429f, t, -1.453152027f);
  poly = fmaf(poly, t, 1.421413741f);
  poly = fmaf(poly, t, -0.284496736f);
  poly = fmaf(poly, t, 0.254829592f);
  float e;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(z * z * -1.4426950408889634f));
  const float erf_abs = fmaf(-poly * t, e, 1.0f);          // erf(|x| / sqrt 2)
  const float hx = 0.5f * x;
  return fmaf(fabsf(hx), erf_abs, hx);                     // 0.5 x (1 + sign(x) erf_abs)
}

// ---------------------------------------------------------------------------------------------- epilogues
// v[32] = raw fp32 accumulators of row `row`, columns [col, col+32) of the output.
template <int EPI, bool OUT_BF16>
__device__ __forceinline__ void epi_store32(const GemmParams& p, int row, int col, uint32_t (&v)[32],
                                            const float4* resid_regs = nullptr) {
  const sr_epilogue_t& e = p.e;
  float x[32];
#pragma unroll
  for (int i = 0; i < 32; i += 4) {
    float4 b = e.bias ? __ldg(reinterpret_cast<const float4*>(e.bias + col + i)) : make_float4(0.f, 0.f, 0.f, 0.f);
    x[i + 0] = __uint_as_float(v[i + 0]) + b.x;
    x[i + 1] = __uint_as_float(v[i + 1]) + b.y;
    x[i + 2] = __uint_as_float(v[i + 2]) + b.z;
    x[i + 3] = __uint_as_float(v[i + 3]) + b.w;
  }
  if (EPI == SR_EPI_BIAS_GELU) {
#pragma unroll
    for (int i = 0; i < 32; ++i) x[i] = gelu_erf_fast(bf16_round(x[i]));
  }
  if (EPI == SR_EPI_LS_RESIDUAL) {
    float* o = reinterpret_cast<float*>(e.out) + static_cast<int64_t>(row) * e.ldo + col;
#pragma unroll
    for (int i = 0; i < 32; i += 4) {
      const float4 rr = resid_regs[i >> 2];                  // loaded before the accumulator was ready (see kernel)
      float4 g = e.gamma ? __ldg(reinterpret_cast<const float4*>(e.gamma + col + i)) : make_float4(1.f, 1.f, 1.f, 1.f);
      float4 y;
      const bool ex = e.exact_residual != 0;               // fp32 accumulate across split-precision passes (dpt.cu)
      y.x = rr.x + g.x * (ex ? x[i + 0] : bf16_round(x[i + 0]));
      y.y = rr.y + g.y * (ex ? x[i + 1] : bf16_round(x[i + 1]));
      y.z = rr.z + g.z * (ex ? x[i + 2] : bf16_round(x[i + 2]));
      y.w = rr.w + g.w * (ex ? x[i + 3] : bf16_round(x[i + 3]));
      *reinterpret_cast<float4*>(o + i) = y;
    }
    return;
  }
  if (OUT_BF16) {
    __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(e.out) + static_cast<int64_t>(row) * e.ldo + col;
#pragma unroll
    for (int i = 0; i < 32; i += 8) {
      uint4 w;
      w.x = pack_bf16x2(x[i + 0], x[i + 1]);
      w.y = pack_bf16x2(x[i + 2], x[i + 3]);
      w.z = pack_bf16x2(x[i + 4], x[i + 5]);
      w.w = pack_bf16x2(x[i + 6], x[i + 7]);
      *reinterpret_cast<uint4*>(o + i) = w;
    }
  } else {
    float* o = reinterpret_cast<float*>(e.out) + static_cast<int64_t>(row) * e.ldo + col;
#pragma unroll
    for (int i = 0; i < 32; i += 4) *reinterpret_cast<float4*>(o + i) = make_float4(x[i], x[i + 1], x[i + 2], x[i + 3]);
  }
}

// v[32] = raw accumulators of one row, columns [col, col+32): bias (+ GELU) and pack to 16 bf16x2 words.
template <int EPI>
__device__ __forceinline__ void epi_pack32(const GemmParams& p, int col, const uint32_t (&v)[32], uint32_t* out16) {
  const sr_epilogue_t& e = p.e;
#pragma unroll
  for (int i = 0; i < 32; i += 4) {
    const float4 b = e.bias ? __ldg(reinterpret_cast<const float4*>(e.bias + col + i)) : make_float4(0.f, 0.f, 0.f, 0.f);
    float x0 = __uint_as_float(v[i + 0]) + b.x, x1 = __uint_as_float(v[i + 1]) + b.y;
    float x2 = __uint_as_float(v[i + 2]) + b.z, x3 = __uint_as_float(v[i + 3]) + b.w;
    if (EPI == SR_EPI_BIAS_GELU) {
      x0 = gelu_erf_fast(bf16_round(x0));
      x1 = gelu_erf_fast(bf16_round(x1));
      x2 = gelu_erf_fast(bf16_round(x2));
      x3 = gelu_erf_fast(bf16_round(x3));
    }
    out16[(i >> 1) + 0] = pack_bf16x2(x0, x1);
    out16[(i >> 1) + 1] = pack_bf16x2(x2, x3);
  }
}

// RoPE table rows of one token, fetched before the accumulator is ready (pos -> table is a dependent load chain).
struct RopeRow {
  float c[2][16];   // [y | x][frequency]
  float s[2][16];
};
__device__ __forceinline__ void load_rope_row(const sr_epilogue_t& e, int row, RopeRow& rr) {
  const int2 yx = __ldg(reinterpret_cast<const int2*>(e.pos) + row);
#pragma unroll
  for (int half = 0; half < 2; ++half) {
    const int ppos = half == 0 ? yx.x : yx.y;
    const float4* ct = reinterpret_cast<const float4*>(e.rope_cos + static_cast<int64_t>(ppos) * 16);
    const float4* st = reinterpret_cast<const float4*>(e.rope_sin + static_cast<int64_t>(ppos) * 16);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float4 c = __ldg(ct + i), sn = __ldg(st + i);
      rr.c[half][4 * i + 0] = c.x; rr.c[half][4 * i + 1] = c.y; rr.c[half][4 * i + 2] = c.z; rr.c[half][4 * i + 3] = c.w;
      rr.s[half][4 * i + 0] = sn.x; rr.s[half][4 * i + 1] = sn.y; rr.s[half][4 * i + 2] = sn.z; rr.s[half][4 * i + 3] = sn.w;
    }
  }
}

// One 64-wide head of q or k for one token: bias -> bf16 round -> LayerNorm(64) -> 2D RoPE -> bf16 store.
// `which` = 0 (q) / 1 (k) / 2 (v: bias + store only).  Reductions run as 4 independent chains: with 2 epilogue warps per
// SM sub-partition a 64-long dependent FADD chain is pure exposed latency.
__device__ __forceinline__ void epi_qkv_head(const GemmParams& p, int col, int which, uint32_t (&v0)[32],
                                             uint32_t (&v1)[32], const RopeRow& rope, uint32_t (&pk)[32]) {
  const sr_epilogue_t& e = p.e;
  float x[64];
#pragma unroll
  for (int i = 0; i < 32; ++i) {
    x[i] = __uint_as_float(v0[i]);
    x[32 + i] = __uint_as_float(v1[i]);
  }
  if (e.bias) {
#pragma unroll
    for (int i = 0; i < 64; i += 4) {
      float4 b = __ldg(reinterpret_cast<const float4*>(e.bias + col + i));
      x[i] += b.x; x[i + 1] += b.y; x[i + 2] += b.z; x[i + 3] += b.w;
    }
  }
  if (which < 2 && (e.qk_norm || e.rope)) {
    // the reference's Linear returned bf16; LayerNorm / RoPE then run in fp32 on that rounded value
#pragma unroll
    for (int i = 0; i < 64; ++i) x[i] = bf16_round(x[i]);
    if (e.qk_norm) {
      float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
      for (int i = 0; i < 64; i += 4) { s0 += x[i]; s1 += x[i + 1]; s2 += x[i + 2]; s3 += x[i + 3]; }
      const float mean = ((s0 + s1) + (s2 + s3)) * (1.0f / 64.0f);
      float q0 = 0.f, q1 = 0.f, q2 = 0.f, q3 = 0.f;
#pragma unroll
      for (int i = 0; i < 64; i += 4) {
        const float d0 = x[i] - mean, d1 = x[i + 1] - mean, d2 = x[i + 2] - mean, d3 = x[i + 3] - mean;
        q0 = fmaf(d0, d0, q0); q1 = fmaf(d1, d1, q1); q2 = fmaf(d2, d2, q2); q3 = fmaf(d3, d3, q3);
      }
      const float rstd = rsqrtf(((q0 + q1) + (q2 + q3)) * (1.0f / 64.0f) + e.qk_eps);
      const float* w = which == 0 ? e.qn_w : e.kn_w;
      const float* b = which == 0 ? e.qn_b : e.kn_b;
#pragma unroll
      for (int i = 0; i < 64; i += 4) {
        float4 ww = __ldg(reinterpret_cast<const float4*>(w + i));
        float4 bb = __ldg(reinterpret_cast<const float4*>(b + i));
        x[i + 0] = (x[i + 0] - mean) * rstd * ww.x + bb.x;
        x[i + 1] = (x[i + 1] - mean) * rstd * ww.y + bb.y;
        x[i + 2] = (x[i + 2] - mean) * rstd * ww.z + bb.z;
        x[i + 3] = (x[i + 3] - mean) * rstd * ww.w + bb.w;
      }
    }
    if (e.rope) {
      // rope.py:196-207: first 32 features rotate with the y position, last 32 with x; inside each half the
      // pairs are (i, i+16) (rotate-half, rope.py:136-138) and both members share frequency i (rope.py:119).
#pragma unroll
      for (int half = 0; half < 2; ++half) {
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const float a = x[half * 32 + i], b = x[half * 32 + 16 + i];
          x[half * 32 + i] = a * rope.c[half][i] - b * rope.s[half][i];
          x[half * 32 + 16 + i] = b * rope.c[half][i] + a * rope.s[half][i];
        }
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 64; i += 2) pk[i >> 1] = pack_bf16x2(x[i], x[i + 1]);   // the kernel stores them row-contiguously
}

// ---------------------------------------------------------------------------------------------- kernel
// AMN / BMN = 1: that operand is MN-major in global memory, i.e. stored [K (contraction) rows, M or N contiguous] — the layout
// the gradients of a Linear have (dW = dY^T X reads dY [T, N] and X [T, K] as they lie; dX = dY W reads W [N, K] as it lies),
// so the training step needs no transposed copies.  A stage then holds [64-wide block][64 contraction rows][128 B] per
// operand (SWIZZLE_128B, one 2-D TMA box {64 features, 64 rows} per block); the UMMA descriptors walk it with start + 2048 B per K = 16
// step, SBO = 1024 (8-row atom), LBO = 8192 (next 64-wide block) and the MN-major bits of the instruction descriptor.
template <int BN, int EPI, bool OUT_BF16, int AMN = 0, int BMN = 0>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
               const GemmParams p) {
  using Cfg = GemmCfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[Cfg::kStages];
  __shared__ __align__(8) uint64_t empty_bar[Cfg::kStages];
  __shared__ __align__(8) uint64_t tmem_full_bar[2];
  __shared__ __align__(8) uint64_t tmem_empty_bar[2];
  __shared__ uint32_t tmem_base_slot;

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);   // provably warp-uniform (see elect_one)
  const int lane = threadIdx.x & 31;
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));

  const int num_m = (p.M + kBM - 1) / kBM;
  const int num_n = p.N / BN;
  const int num_tiles = num_m * num_n;
  const int num_kb = (p.K + kBK - 1) / kBK;             // a ragged K tail only with two MN-major operands (TMA zero-fills both)

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tmem_full_bar[s], 1);
      mbar_init(&tmem_empty_bar[s], kNumEpiWarps);
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<Cfg::kTmemCols>(&tmem_base_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_slot;

  // register re-distribution: the 4 control warps keep 72 registers, the 8 epilogue warps (which hold a 128-wide fp32
  // residual row fragment across the accumulator wait) grow to 216:  128 x 72 + 256 x 216 = 64512 <= 65536
  if (warp < kEpiWarp0) {
  asm volatile("setmaxnreg.dec.sync.aligned.u32 72;");
  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer (whole warp, one elected lane issues)
    int stage = 0;
    uint32_t phase = 0;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
      const int m_blk = t / num_n, n_blk = t % num_n;
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait_relaxed(&empty_bar[stage], phase ^ 1);
        uint8_t* sa = smem + stage * Cfg::kStageBytes;
        uint8_t* sb = sa + Cfg::kABytes;
        if (elect_one()) {
          mbar_arrive_expect_tx(&full_bar[stage], Cfg::kStageBytes);
          if (AMN == 1) {   // one {64 features, 64 contraction rows} box per 64-wide block, 8 KB apart
#pragma unroll
            for (int b = 0; b < kBM / 64; ++b) tma_load_2d(sa + b * 8192, &tmap_a, &full_bar[stage], m_blk * kBM + b * 64, kb * kBK);
          } else if (AMN == 2) {
            // 3x3 / stride 1 / pad 1 convolution on a zero-bordered channels-last map [S, H + 2, W + 2, hi(C) | lo(C)]: output
            // row r needs input row r + (ky - 1) (W + 2) + (kx - 1) for tap (ky, kx) — a ROW SHIFT of the same matrix, so the
            // im2col matrix is never materialised.  K loop = 3 split-precision parts (hi x Whi, lo x Whi, hi x Wlo) x 9 taps x
            // C / 64 blocks; rows shifted past either end of the matrix are zero-filled by TMA.
            const int bpt = p.conv_c >> 6;
            const int idx = kb / bpt, cb = kb - idx * bpt;
            const int part = idx / 9, tap = idx - part * 9;
            const int ky = tap / 3, kx = tap - ky * 3;
            tma_load_2d(sa, &tmap_a, &full_bar[stage], (part == 1 ? p.conv_c : 0) + cb * kBK,
                        m_blk * kBM + (ky - 1) * p.conv_wp + (kx - 1));
          } else {
            tma_load_2d(sa, &tmap_a, &full_bar[stage], kb * kBK, m_blk * kBM);
          }
          if (BMN) {
#pragma unroll
            for (int b = 0; b < BN / 64; ++b) tma_load_2d(sb + b * 8192, &tmap_b, &full_bar[stage], n_blk * BN + b * 64, kb * kBK);
          } else {
            tma_load_2d(sb, &tmap_b, &full_bar[stage], kb * kBK, n_blk * BN);
          }
        }
        __syncwarp();
        if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (whole warp, one elected lane issues)
    constexpr uint32_t idesc = umma_idesc_bf16(kBM, BN, AMN == 1 ? 1 : 0, BMN);
    int stage = 0;
    uint32_t phase = 0;
    int it = 0;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x, ++it) {
      const int as = it & 1;
      mbar_wait(&tmem_empty_bar[as], ((it >> 1) & 1) ^ 1);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + as * BN;
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + stage * Cfg::kStageBytes);
        const uint32_t sb = sa + Cfg::kABytes;
        // K-major: +32 bytes per K = 16 slice inside the 128 B swizzle atom -> +2 in the (addr >> 4) field;
        // MN-major: +16 contraction rows x 128 B = 2048 bytes -> +128
        const uint64_t da = AMN == 1 ? umma_desc_sw128(sa, 8192, 1024) : umma_desc_sw128(sa, 16, 1024);
        const uint64_t db = BMN ? umma_desc_sw128(sb, 8192, 1024) : umma_desc_sw128(sb, 16, 1024);
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < kBK / 16; ++k)
            umma_ss(tmem_d, da + (AMN == 1 ? 128 : 2) * k, db + (BMN ? 128 : 2) * k, idesc, (kb | k) != 0 ? 1u : 0u);
          umma_commit(&empty_bar[stage]);
        }
        __syncwarp();
        if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
      }
      if (elect_one()) umma_commit(&tmem_full_bar[as]);
      __syncwarp();
    }
  }
  } else {
    // ------------------------------------------------------------------ epilogue
    asm volatile("setmaxnreg.inc.sync.aligned.u32 216;");
    const int ew = warp - kEpiWarp0;
    const int quarter = warp & 3;          // TMEM lane quarter this warp may access
    const int half = ew >> 2;              // which half of the BN columns
    constexpr int kColsPerWarp = BN / 2;
    int it = 0;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x, ++it) {
      const int m_blk = t / num_n, n_blk = t % num_n;
      const int as = it & 1;
      const int row = m_blk * kBM + quarter * 32 + lane;
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + as * BN + half * kColsPerWarp;
      const int col_base = n_blk * BN + half * kColsPerWarp;
      // LayerScale-residual epilogue (fp32 read-modify-write of the residual stream).  tcgen05.ld hands every lane one
      // ROW, and a row-per-lane float4 access touches 32 cache lines per instruction -- at K = 1024 (proj) that L1
      // traffic, not the MMA, bounded the kernel (596 TFLOP/s).  The accumulator chunk is therefore transposed through a
      // 4 KB XOR-swizzled shared-memory buffer per warp so that global loads / stores are row-contiguous: lane l handles
      // row 4 i + l / 8, float4 column l % 8 -- four full 128 B lines per instruction.  The residual fragment does not
      // depend on the accumulator: it is loaded (in that layout) BEFORE the wait for the MMA and flies under it.
      float4 resid_regs[EPI == SR_EPI_LS_RESIDUAL ? kColsPerWarp / 4 : 1];
      const int trow = lane >> 3, tc4 = lane & 7;
      const int row0 = m_blk * kBM + quarter * 32;
      if (EPI == SR_EPI_LS_RESIDUAL) {
#pragma unroll
        for (int c = 0; c < kColsPerWarp / 32; ++c) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int r = row0 + i * 4 + trow;
            if (r < p.M)
              resid_regs[c * 8 + i] = *reinterpret_cast<const float4*>(p.e.resid + static_cast<int64_t>(r) * p.e.ldr +
                                                                       col_base + c * 32 + tc4 * 4);
          }
        }
      }
      RopeRow rope_row;
      if (EPI == SR_EPI_QKV && p.e.rope && row < p.M && p.e.col_offset + col_base < 2 * p.e.qkv_dim)
        load_rope_row(p.e, row, rope_row);
      mbar_wait_relaxed(&tmem_full_bar[as], (it >> 1) & 1);
      tc_fence_after();
      if (EPI == SR_EPI_QKV) {
        // one head (64 columns = 128 B of bf16) per row and chunk; stored through the swizzled transpose buffer like the
        // other bf16 epilogues (rows past M compute on garbage and are simply not stored)
        const uint32_t tbuf = smem_u32(smem + Cfg::kStages * Cfg::kStageBytes) + ew * 4096;
        __nv_bfloat16* outp = reinterpret_cast<__nv_bfloat16*>(p.e.out);
#pragma unroll 1
        for (int c = 0; c < kColsPerWarp; c += 64) {
          uint32_t v0[32], v1[32], pk[32];
          tmem_ld32(taddr + c, v0);
          tmem_ld32(taddr + c + 32, v1);
          tmem_ld_wait();
          const int col = col_base + c;
          const int which = static_cast<int>((p.e.col_offset + col) / p.e.qkv_dim);
          epi_qkv_head(p, col, which, v0, v1, rope_row, pk);
#pragma unroll
          for (int k = 0; k < 8; ++k)
            sts128(tbuf + lane * 128 + ((k ^ (lane & 7)) << 4), pk[4 * k], pk[4 * k + 1], pk[4 * k + 2], pk[4 * k + 3]);
          __syncwarp();
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int rl = i * 4 + trow;
            const uint4 w = lds128u(tbuf + rl * 128 + ((tc4 ^ (rl & 7)) << 4));
            const int r = row0 + rl;
            if (r < p.M) *reinterpret_cast<uint4*>(outp + static_cast<int64_t>(r) * p.e.ldo + col + tc4 * 8) = w;
          }
          __syncwarp();
        }
      } else {
        if (EPI == SR_EPI_LS_RESIDUAL) {
          const uint32_t tbuf = smem_u32(smem + Cfg::kStages * Cfg::kStageBytes) + ew * 4096;   // shared-space byte address
          const sr_epilogue_t& e = p.e;
#pragma unroll
          for (int c = 0; c < kColsPerWarp / 32; ++c) {
            uint32_t v[32];
            tmem_ld32(taddr + c * 32, v);
            tmem_ld_wait();
            // lane = row: float4 k of the row goes to slot k ^ (lane & 7) (conflict-free both ways)
#pragma unroll
            for (int k = 0; k < 8; ++k)
              sts128(tbuf + lane * 128 + ((k ^ (lane & 7)) << 4), v[4 * k], v[4 * k + 1], v[4 * k + 2], v[4 * k + 3]);
            __syncwarp();
            const int col = col_base + c * 32 + tc4 * 4;
            const float4 b4 = e.bias ? __ldg(reinterpret_cast<const float4*>(e.bias + col)) : make_float4(0.f, 0.f, 0.f, 0.f);
            const float4 g4 = e.gamma ? __ldg(reinterpret_cast<const float4*>(e.gamma + col)) : make_float4(1.f, 1.f, 1.f, 1.f);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const int rl = i * 4 + trow;
              const float4 a = lds128(tbuf + rl * 128 + ((tc4 ^ (rl & 7)) << 4));
              const int r = row0 + rl;
              if (r < p.M) {
                const float4 rr = resid_regs[c * 8 + i];
                float4 y;
                const bool ex = e.exact_residual != 0;     // fp32 accumulate across split-precision passes (dpt.cu)
                y.x = rr.x + g4.x * (ex ? a.x + b4.x : bf16_round(a.x + b4.x));
                y.y = rr.y + g4.y * (ex ? a.y + b4.y : bf16_round(a.y + b4.y));
                y.z = rr.z + g4.z * (ex ? a.z + b4.z : bf16_round(a.z + b4.z));
                y.w = rr.w + g4.w * (ex ? a.w + b4.w : bf16_round(a.w + b4.w));
                *reinterpret_cast<float4*>(reinterpret_cast<float*>(e.out) + static_cast<int64_t>(r) * e.ldo + col) = y;
              }
            }
            __syncwarp();
          }
        } else if (OUT_BF16) {
          // bf16 outputs (bias / bias + GELU): 64 columns = 128 B per row and chunk, transposed through the same swizzled
          // 4 KB/warp buffer so that every store instruction writes four full 128 B lines (lane l: row 4 i + l / 8,
          // 16 B piece l % 8) instead of 16 B in each of 32 rows.
          const uint32_t tbuf = smem_u32(smem + Cfg::kStages * Cfg::kStageBytes) + ew * 4096;
          __nv_bfloat16* outp = reinterpret_cast<__nv_bfloat16*>(p.e.out);
#pragma unroll 1
          for (int c = 0; c < kColsPerWarp; c += 64) {
            uint32_t v0[32], v1[32], pk[32];
            tmem_ld32(taddr + c, v0);
            tmem_ld32(taddr + c + 32, v1);
            tmem_ld_wait();
            epi_pack32<EPI>(p, col_base + c, v0, &pk[0]);
            epi_pack32<EPI>(p, col_base + c + 32, v1, &pk[16]);
#pragma unroll
            for (int k = 0; k < 8; ++k)
              sts128(tbuf + lane * 128 + ((k ^ (lane & 7)) << 4), pk[4 * k], pk[4 * k + 1], pk[4 * k + 2], pk[4 * k + 3]);
            __syncwarp();
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const int rl = i * 4 + trow;
              const uint4 w = lds128u(tbuf + rl * 128 + ((tc4 ^ (rl & 7)) << 4));
              const int r = row0 + rl;
              if (r < p.M) *reinterpret_cast<uint4*>(outp + static_cast<int64_t>(r) * p.e.ldo + col_base + c + tc4 * 8) = w;
            }
            __syncwarp();
          }
        } else {
#pragma unroll 1
          for (int c = 0; c < kColsPerWarp; c += 32) {
            uint32_t v[32];
            tmem_ld32(taddr + c, v);
            tmem_ld_wait();
            if (row < p.M) epi_store32<EPI, OUT_BF16>(p, row, col_base + c, v);
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty_bar[as]);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<Cfg::kTmemCols>(tmem_base);
  }
}

// ---------------------------------------------------------------------------------------------- host
extern int g_launch_count;

template <int BN, int EPI, bool OUT_BF16, int AMN = 0, int BMN = 0>
static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, cudaStream_t st) {
  using Cfg = GemmCfg<BN>;
  auto kern = gemm_tc_kernel<BN, EPI, OUT_BF16, AMN, BMN>;
  static bool configured = false;  // per instantiation
  if (!configured) {
    SR_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    configured = true;
  }
  const int num_tiles = ((p.M + kBM - 1) / kBM) * (p.N / BN);
  const int grid = num_tiles < kNumSMs ? num_tiles : kNumSMs;
  kern<<<grid, kGemmThreads, Cfg::kSmemBytes, st>>>(ta, tb, p);
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

template <int BN>
static int dispatch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, cudaStream_t st) {
  const sr_epilogue_t& e = p.e;
  switch (e.kind) {
    case SR_EPI_BIAS:
      return e.out_bf16 ? launch_gemm<BN, SR_EPI_BIAS, true>(ta, tb, p, st)
                        : launch_gemm<BN, SR_EPI_BIAS, false>(ta, tb, p, st);
    case SR_EPI_BIAS_GELU:
      return e.out_bf16 ? launch_gemm<BN, SR_EPI_BIAS_GELU, true>(ta, tb, p, st)
                        : launch_gemm<BN, SR_EPI_BIAS_GELU, false>(ta, tb, p, st);
    case SR_EPI_LS_RESIDUAL:
      return launch_gemm<BN, SR_EPI_LS_RESIDUAL, false>(ta, tb, p, st);
    case SR_EPI_QKV:
      return launch_gemm<BN, SR_EPI_QKV, true>(ta, tb, p, st);
  }
  return fail(SR_ERR_SHAPE, "sr_gemm_bf16: unknown epilogue kind %d", e.kind);
}

// Linear gradients (operands as they lie in memory): only the epilogues those use.
template <int BN, int AMN, int BMN>
static int dispatch_gemm_t(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, cudaStream_t st) {
  const sr_epilogue_t& e = p.e;
  switch (e.kind) {
    case SR_EPI_BIAS:
      return e.out_bf16 ? launch_gemm<BN, SR_EPI_BIAS, true, AMN, BMN>(ta, tb, p, st)
                        : launch_gemm<BN, SR_EPI_BIAS, false, AMN, BMN>(ta, tb, p, st);
    case SR_EPI_LS_RESIDUAL:
      return launch_gemm<BN, SR_EPI_LS_RESIDUAL, false, AMN, BMN>(ta, tb, p, st);
  }
  return fail(SR_ERR_SHAPE, "sr_gemm_bf16_t: epilogue kind %d is not available with MN-major operands (BIAS, LS_RESIDUAL)", e.kind);
}

}  // namespace sr

extern "C" int sr_conv3x3_bf16(const void* xs, int64_t ld_xs, int64_t C, int64_t padded_width, const void* w, int64_t ldw,
                               int64_t rows, int64_t N, const sr_epilogue_t* epi, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(xs && w && epi && epi->out, "sr_conv3x3_bf16: null pointer");
  SR_REQUIRE(C > 0 && C % 64 == 0 && ld_xs >= 2 * C && ld_xs % 8 == 0, "sr_conv3x3_bf16: C must be a multiple of 64, rows [hi(C) | lo(C)]");
  SR_REQUIRE(N > 0 && N % 128 == 0 && ldw >= 27 * C && ldw % 8 == 0, "sr_conv3x3_bf16: weights [N %% 128 == 0, 27 C] = (Whi | Whi | Wlo) x (tap, c)");
  SR_REQUIRE(rows > 0 && rows < (1ll << 31) && padded_width >= 3 && padded_width < (1 << 20), "sr_conv3x3_bf16: bad map size");
  SR_REQUIRE((reinterpret_cast<uintptr_t>(xs) & 15) == 0 && (reinterpret_cast<uintptr_t>(w) & 15) == 0, "sr_conv3x3_bf16: 16-byte alignment");
  const sr_epilogue_t& e = *epi;
  SR_REQUIRE((e.kind == SR_EPI_BIAS && !e.out_bf16) || e.kind == SR_EPI_LS_RESIDUAL, "sr_conv3x3_bf16: fp32 output (BIAS or LS_RESIDUAL)");
  SR_REQUIRE((reinterpret_cast<uintptr_t>(e.out) & 15) == 0 && e.ldo % 4 == 0 && e.ldo >= N, "sr_conv3x3_bf16: bad output");
  if (e.kind == SR_EPI_LS_RESIDUAL)
    SR_REQUIRE(e.resid && (reinterpret_cast<uintptr_t>(e.resid) & 15) == 0 && e.ldr % 4 == 0, "sr_conv3x3_bf16: bad residual");
  const int64_t tiles256 = ((rows + kBM - 1) / kBM) * (N / 256);
  const int BN = (N % 256 == 0 && tiles256 >= kNumSMs / 2) ? 256 : 128;
  CUtensorMap ta, tb;
  {
    uint64_t dims[2] = {static_cast<uint64_t>(2 * C), static_cast<uint64_t>(rows)};
    uint64_t strides[1] = {static_cast<uint64_t>(ld_xs) * 2};
    uint32_t box[2] = {kBK, kBM};
    if (int rc = make_tmap_bf16(&ta, xs, 2, dims, strides, box)) return rc;
  }
  {
    uint64_t dims[2] = {static_cast<uint64_t>(27 * C), static_cast<uint64_t>(N)};
    uint64_t strides[1] = {static_cast<uint64_t>(ldw) * 2};
    uint32_t box[2] = {kBK, static_cast<uint32_t>(BN)};
    if (int rc = make_tmap_bf16(&tb, w, 2, dims, strides, box)) return rc;
  }
  GemmParams p;
  p.M = static_cast<int>(rows);
  p.N = static_cast<int>(N);
  p.K = static_cast<int>(27 * C);
  p.e = e;
  p.conv_c = static_cast<int>(C);
  p.conv_wp = static_cast<int>(padded_width);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (e.kind == SR_EPI_BIAS)
    return BN == 256 ? launch_gemm<256, SR_EPI_BIAS, false, 2, 0>(ta, tb, p, st) : launch_gemm<128, SR_EPI_BIAS, false, 2, 0>(ta, tb, p, st);
  return BN == 256 ? launch_gemm<256, SR_EPI_LS_RESIDUAL, false, 2, 0>(ta, tb, p, st)
                   : launch_gemm<128, SR_EPI_LS_RESIDUAL, false, 2, 0>(ta, tb, p, st);
}

extern "C" int sr_gemm_bf16_t(const void* A, int64_t lda, int32_t a_mn_major, const void* W, int64_t ldw, int32_t b_mn_major,
                              int64_t M, int64_t N, int64_t K, const sr_epilogue_t* epi, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(A && W && epi && epi->out, "sr_gemm_bf16_t: null pointer");
  SR_REQUIRE(b_mn_major == 1 && (a_mn_major == 0 || a_mn_major == 1),
             "sr_gemm_bf16_t: W must be MN-major ([K, N] rows); A K-major or MN-major (plain layouts: sr_gemm_bf16)");
  SR_REQUIRE(M > 0 && N > 0 && K > 0 && M < (1ll << 31) && N < (1ll << 31) && K < (1ll << 31), "sr_gemm_bf16_t: bad sizes");
  SR_REQUIRE(N % 128 == 0, "sr_gemm_bf16_t: N=%lld must be a multiple of 128", (long long)N);
  if (a_mn_major) SR_REQUIRE(M % 64 == 0 && lda >= M, "sr_gemm_bf16_t: MN-major A needs M %% 64 == 0 and lda >= M");
  else SR_REQUIRE(K % kBK == 0 && lda >= K, "sr_gemm_bf16_t: K-major A needs K %% 64 == 0 and lda >= K");
  SR_REQUIRE(ldw >= N && lda % 8 == 0 && ldw % 8 == 0, "sr_gemm_bf16_t: row strides must be multiples of 8 and cover the rows");
  SR_REQUIRE((reinterpret_cast<uintptr_t>(A) & 15) == 0 && (reinterpret_cast<uintptr_t>(W) & 15) == 0,
             "sr_gemm_bf16_t: A and W must be 16-byte aligned");
  const sr_epilogue_t& e = *epi;
  const int out_elt = (e.kind == SR_EPI_LS_RESIDUAL || !e.out_bf16) ? 4 : 2;
  SR_REQUIRE((reinterpret_cast<uintptr_t>(e.out) & 15) == 0 && (e.ldo * out_elt) % 16 == 0 && e.ldo >= N,
             "sr_gemm_bf16_t: out must be 16-byte aligned with a 16-byte multiple row stride >= N");
  if (e.kind == SR_EPI_LS_RESIDUAL)
    SR_REQUIRE(e.resid && (reinterpret_cast<uintptr_t>(e.resid) & 15) == 0 && e.ldr % 4 == 0,
               "sr_gemm_bf16_t: LS_RESIDUAL needs a 16-byte aligned fp32 residual");
  const int64_t tiles256 = ((M + kBM - 1) / kBM) * (N / 256);
  const int BN = (N % 256 == 0 && tiles256 >= kNumSMs / 2) ? 256 : 128;
  CUtensorMap ta, tb;
  if (a_mn_major) {       // [K, M] rows: boxes of 64 features x 64 contraction rows (rows / features past the end zero-fill)
    uint64_t dims[2] = {static_cast<uint64_t>(M), static_cast<uint64_t>(K)};
    uint64_t strides[1] = {static_cast<uint64_t>(lda) * 2};
    uint32_t box[2] = {64, kBK};
    if (int rc = make_tmap_bf16(&ta, A, 2, dims, strides, box)) return rc;
  } else {
    uint64_t dims[2] = {static_cast<uint64_t>(K), static_cast<uint64_t>(M)};
    uint64_t strides[1] = {static_cast<uint64_t>(lda) * 2};
    uint32_t box[2] = {kBK, kBM};
    if (int rc = make_tmap_bf16(&ta, A, 2, dims, strides, box)) return rc;
  }
  {
    uint64_t dims[2] = {static_cast<uint64_t>(N), static_cast<uint64_t>(K)};
    uint64_t strides[1] = {static_cast<uint64_t>(ldw) * 2};
    uint32_t box[2] = {64, kBK};
    if (int rc = make_tmap_bf16(&tb, W, 2, dims, strides, box)) return rc;
  }
  GemmParams p;
  p.M = static_cast<int>(M);
  p.N = static_cast<int>(N);
  p.K = static_cast<int>(K);
  p.e = e;
  p.conv_c = p.conv_wp = 0;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (a_mn_major) return BN == 256 ? dispatch_gemm_t<256, 1, 1>(ta, tb, p, st) : dispatch_gemm_t<128, 1, 1>(ta, tb, p, st);
  return BN == 256 ? dispatch_gemm_t<256, 0, 1>(ta, tb, p, st) : dispatch_gemm_t<128, 0, 1>(ta, tb, p, st);
}

extern "C" int sr_gemm_bf16(const void* A, int64_t lda, const void* W, int64_t ldw, int64_t M, int64_t N, int64_t K,
                            const sr_epilogue_t* epi, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(A && W && epi && epi->out, "sr_gemm_bf16: null pointer");
  SR_REQUIRE(M > 0 && N > 0 && K > 0, "sr_gemm_bf16: empty problem M=%lld N=%lld K=%lld", (long long)M, (long long)N,
             (long long)K);
  SR_REQUIRE(K % kBK == 0, "sr_gemm_bf16: K=%lld must be a multiple of 64", (long long)K);
  SR_REQUIRE(N % 128 == 0, "sr_gemm_bf16: N=%lld must be a multiple of 128", (long long)N);
  SR_REQUIRE(lda % 8 == 0 && ldw % 8 == 0 && lda >= K && ldw >= K, "sr_gemm_bf16: lda/ldw must be >= K and multiples of 8");
  SR_REQUIRE((reinterpret_cast<uintptr_t>(A) & 15) == 0 && (reinterpret_cast<uintptr_t>(W) & 15) == 0,
             "sr_gemm_bf16: A and W must be 16-byte aligned");
  SR_REQUIRE(M < (1ll << 31) && N < (1ll << 31), "sr_gemm_bf16: M, N must fit int32");
  const sr_epilogue_t& e = *epi;
  const int out_elt = (e.kind == SR_EPI_LS_RESIDUAL || !e.out_bf16) ? 4 : 2;
  SR_REQUIRE((reinterpret_cast<uintptr_t>(e.out) & 15) == 0 && (e.ldo * out_elt) % 16 == 0 && e.ldo >= N,
             "sr_gemm_bf16: out must be 16-byte aligned with a 16-byte multiple row stride >= N");
  if (e.kind == SR_EPI_LS_RESIDUAL) {
    SR_REQUIRE(e.resid && (reinterpret_cast<uintptr_t>(e.resid) & 15) == 0 && e.ldr % 4 == 0,
               "sr_gemm_bf16: LS_RESIDUAL needs a 16-byte aligned fp32 residual");
  }
  if (e.kind == SR_EPI_QKV) {
    SR_REQUIRE(e.head_dim == 64, "sr_gemm_bf16: QKV epilogue supports head_dim 64 (got %d)", e.head_dim);
    SR_REQUIRE(e.out_bf16, "sr_gemm_bf16: QKV epilogue writes bf16");
    SR_REQUIRE(e.qkv_dim > 0 && e.qkv_dim % 64 == 0 && e.col_offset % 64 == 0 && e.col_offset + N <= 3 * e.qkv_dim,
               "sr_gemm_bf16: QKV epilogue column mapping out of range");
    if (e.qk_norm) SR_REQUIRE(e.qn_w && e.qn_b && e.kn_w && e.kn_b, "sr_gemm_bf16: qk_norm needs q/k norm parameters");
    if (e.rope) SR_REQUIRE(e.rope_cos && e.rope_sin && e.pos, "sr_gemm_bf16: rope needs tables and positions");
  }

  CUtensorMap ta, tb;
  // 128 x 256 tiles by default; when that leaves most SMs idle (few rows: one query image of the localisation workload, the
  // coarse DPT feature maps) the 128 x 128 tiling doubles the tile count
  const int64_t tiles256 = ((M + kBM - 1) / kBM) * (N / 256);
  const int BN = (N % 256 == 0 && tiles256 >= kNumSMs / 2) ? 256 : 128;
  {
    uint64_t dims[2] = {static_cast<uint64_t>(K), static_cast<uint64_t>(M)};
    uint64_t strides[1] = {static_cast<uint64_t>(lda) * 2};
    uint32_t box[2] = {kBK, kBM};
    if (int rc = make_tmap_bf16(&ta, A, 2, dims, strides, box)) return rc;
  }
  {
    uint64_t dims[2] = {static_cast<uint64_t>(K), static_cast<uint64_t>(N)};
    uint64_t strides[1] = {static_cast<uint64_t>(ldw) * 2};
    uint32_t box[2] = {kBK, static_cast<uint32_t>(BN)};
    if (int rc = make_tmap_bf16(&tb, W, 2, dims, strides, box)) return rc;
  }
  GemmParams p;
  p.M = static_cast<int>(M);
  p.N = static_cast<int>(N);
  p.K = static_cast<int>(K);
  p.e = e;
  p.conv_c = p.conv_wp = 0;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  return BN == 256 ? dispatch_gemm<256>(ta, tb, p, st) : dispatch_gemm<128>(ta, tb, p, st);
}
