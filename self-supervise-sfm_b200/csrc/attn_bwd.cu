// Flash attention BACKWARD on tcgen05, head_dim 64 (sm_100a) — the autograd of F.scaled_dot_product_attention at
// sailrecon/layers/attention.py:103-109 for the training step (train/train_imc.py:358-380), on the same structured
// [scene prefix | own segment] layout as the forward kernels (attn_tc4.cu).
//
// Two passes of ONE kernel template, no atomics (deterministic):
//   kMode 0  query-stationary   CTA owns 128 query rows (Q, dO in smem), streams 64-row K/V tiles (prefix, then own):
//              S  = Q K_j^T, dP = dO V_j^T            SS, M=128 N=64 K=64
//              P  = exp2(S * scale_log2 - lse_row);   dS = P * (dP - delta_row) * scale      (bf16, written over dP in TMEM)
//              dQ += dS K_j                            TS, A = dS from TMEM, B = K_j MN-major
//   kMode 1  key-stationary     CTA owns 128 K/V rows (K, V in smem), streams 64-row Q/dO tiles (+ their lse, delta):
//              S^T = K Q_i^T, dP^T = V dO_i^T         SS
//              P^T = exp2(S^T * scale_log2 - lse_col); dS^T = P^T * (dP^T - delta_col) * scale
//              dV += P^T dO_i ;  dK += dS^T Q_i        TS, B = dO_i / Q_i MN-major
//            with over_prefix = 1 the stationary rows are the shared scene prefix and the stream runs over the query
//            tiles of EVERY segment (the prefix gradient is the sum over the segments).
// TMEM per CTA: S 64 | dP 64 | acc0 64 | acc1 64 = 256 columns, 80 KB of smem: two CTAs per SM, whose MMA and softmax
// phases interleave (inside one CTA the tile loop is serial: SS -> softmax -> TS).
//   warp 0 TMA producer, warp 1 MMA issuer, warp 2 TMEM allocator, warps 4-11 softmax gradient: two threads per TMEM lane
//   (warps 4-7 score columns 0-31, warps 8-11 columns 32-63).  kMode 0 has a spare accumulator: dS is double-buffered there
//   and S(t+1) / dP(t+1) are issued under the arithmetic of tile t.
#include <cstdlib>

#include "attn_math.cuh"
#include "common.cuh"
#include "sr_internal.h"

namespace sr {
extern int g_launch_count;

constexpr int kBwStages = 3;
constexpr int kBwXTile = 128 * 64 * 2;                // 16 KB stationary operand
constexpr int kBwSTile = 64 * 64 * 2;                 //  8 KB streamed operand
constexpr int kBwStat = 64 * 4;                       // 64 fp32 row statistics of a streamed query tile
constexpr int kBwStageBytes = 2 * kBwSTile;           // A tile | B tile (1024-byte aligned: SWIZZLE_128B)
constexpr int kBwThreads = 384;                     // 4 control warps + 8 softmax-gradient warps (two per TMEM lane quarter)
constexpr int kBwSmem = 2 * kBwXTile + kBwStages * (kBwStageBytes + 2 * kBwStat) + 1024;

struct BwdParams {
  __nv_bfloat16* d0;             // mode 0: dq            mode 1: dv (own rows) or dpv
  __nv_bfloat16* d1;             //                        mode 1: dk or dpk
  int64_t ldd;
  const float* lse;              // [num_heads, batch, stat_pad]
  const float* delta;
  int stat_pad;
  int batch;
  int x_len;                     // valid stationary rows per segment
  int x_batch_stride;            // rows between segments in the OUTPUT matrices
  int prefix_len, n_prefix_tiles;   // mode 0 only
  int own_len, n_own_tiles;         // streamed rows / 64-row tiles per segment
  int over_prefix;               // mode 1: stream over all segments
  float scale_log2, scale;
  int dbg;                       // timing experiments only (SR_ATTN_BWD_DBG): 1 = skip the exp / dS arithmetic, 2 = also skip TMEM loads
};

__device__ __forceinline__ void bulk_load_1d(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(gsrc)), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

template <int kMode>
__global__ void __launch_bounds__(kBwThreads, 2)
attn_bwd_kernel(const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_y,
                const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b,
                const __grid_constant__ CUtensorMap tm_pa, const __grid_constant__ CUtensorMap tm_pb, const BwdParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t x_full, s_full, p_ready, acc_done, sdp_free;
  __shared__ __align__(8) uint64_t full[kBwStages], empty[kBwStages];
  __shared__ uint32_t tmem_base_slot;

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sX = smem;
  uint8_t* sY = sX + kBwXTile;
  uint8_t* sStage = sY + kBwXTile;                     // per stage: A tile | B tile
  uint8_t* sStat = sStage + kBwStages * kBwStageBytes; // per stage: lse[64] | delta[64] of the streamed query tile

  const int head = blockIdx.y;
  const int b = blockIdx.z;
  const int x0 = blockIdx.x * 128;
  const int n_tiles = kMode == 0 ? p.n_prefix_tiles + p.n_own_tiles : (p.over_prefix ? p.batch * p.n_own_tiles : p.n_own_tiles);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm_x);
    tma_prefetch_desc(&tm_y);
    tma_prefetch_desc(&tm_a);
    tma_prefetch_desc(&tm_b);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(&x_full, 1);
    mbar_init(&s_full, 1);
    mbar_init(&p_ready, 8);
    mbar_init(&sdp_free, 8);
    mbar_init(&acc_done, 1);
    for (int s = 0; s < kBwStages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<256>(&tmem_base_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_slot;
  const uint32_t tmem_S = tmem_base;                   // P (bf16) is written over its first 32 columns
  const uint32_t tmem_dP = tmem_base + 64;             // dS (bf16) likewise
  const uint32_t tmem_acc0 = tmem_base + 128;
  const uint32_t tmem_acc1 = tmem_base + 192;

  if (warp < 4) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 56;");
    if (warp == 0) {
      // ---------------------------------------------------------------- TMA producer
      if (elect_one()) {
        mbar_arrive_expect_tx(&x_full, 2 * kBwXTile);
        tma_load_3d(sX, &tm_x, &x_full, head * 64, x0, b);
        tma_load_3d(sY, &tm_y, &x_full, head * 64, x0, b);
      }
      __syncwarp();
      for (int t = 0; t < n_tiles; ++t) {
        const int s = t % kBwStages;
        mbar_wait_relaxed(&empty[s], ((t / kBwStages) & 1) ^ 1);
        if (elect_one()) {
          uint8_t* st = sStage + s * kBwStageBytes;
          if (kMode == 0) {
            mbar_arrive_expect_tx(&full[s], 2 * kBwSTile);
            if (t < p.n_prefix_tiles) {
              tma_load_3d(st, &tm_pa, &full[s], head * 64, t * 64, 0);
              tma_load_3d(st + kBwSTile, &tm_pb, &full[s], head * 64, t * 64, 0);
            } else {
              tma_load_3d(st, &tm_a, &full[s], head * 64, (t - p.n_prefix_tiles) * 64, b);
              tma_load_3d(st + kBwSTile, &tm_b, &full[s], head * 64, (t - p.n_prefix_tiles) * 64, b);
            }
          } else {
            const int seg = p.over_prefix ? t / p.n_own_tiles : b;
            const int i = p.over_prefix ? t % p.n_own_tiles : t;
            mbar_arrive_expect_tx(&full[s], 2 * kBwSTile + 2 * kBwStat);
            tma_load_3d(st, &tm_a, &full[s], head * 64, i * 64, seg);
            tma_load_3d(st + kBwSTile, &tm_b, &full[s], head * 64, i * 64, seg);
            const int64_t so = (static_cast<int64_t>(head) * p.batch + seg) * p.stat_pad + i * 64;
            bulk_load_1d(sStat + s * 2 * kBwStat, p.lse + so, kBwStat, &full[s]);
            bulk_load_1d(sStat + s * 2 * kBwStat + kBwStat, p.delta + so, kBwStat, &full[s]);
          }
        }
        __syncwarp();
      }
    } else if (warp == 1) {
      // ---------------------------------------------------------------- MMA issuer
      constexpr uint32_t idesc_s = umma_idesc_bf16(128, 64, 0, 0);   // X (K-major) x streamed tile (K-major)
      constexpr uint32_t idesc_t = umma_idesc_bf16(128, 64, 0, 1);   // TMEM x streamed tile (MN-major: d contiguous)
      const uint64_t dx = umma_desc_sw128(smem_u32(sX), 16, 1024);
      const uint64_t dy = umma_desc_sw128(smem_u32(sY), 16, 1024);
      mbar_wait(&x_full, 0);
      auto issue_ss = [&](int t) {                           // S(t) = X A_t^T, dP(t) = Y B_t^T into the S / dP columns
        const int s = t % kBwStages;
        const uint32_t a_addr = smem_u32(sStage + s * kBwStageBytes);
        mbar_wait(&full[s], (t / kBwStages) & 1);
        tc_fence_after();
        if (elect_one()) {
          const uint64_t da = umma_desc_sw128(a_addr, 16, 1024);
          const uint64_t db = umma_desc_sw128(a_addr + kBwSTile, 16, 1024);
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_ss(tmem_S, dx + 2 * k, da + 2 * k, idesc_s, k != 0 ? 1u : 0u);
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_ss(tmem_dP, dy + 2 * k, db + 2 * k, idesc_s, k != 0 ? 1u : 0u);
          umma_commit(&s_full);
        }
        __syncwarp();
      };
      if (kMode == 0) {
        // Query-stationary pass: dS(t) goes to its own double-buffered columns (the unused second accumulator), so
        // S(t+1) / dP(t+1) are issued as soon as the softmax warps have READ S(t) / dP(t) (sdp_free) and run under the
        // arithmetic of tile t; the in-order tensor pipe then gives  SS(t+1), TS(t), SS(t+2), TS(t+1), ...
        issue_ss(0);
        for (int t = 0; t < n_tiles; ++t) {
          const int s = t % kBwStages;
          const uint32_t a_addr = smem_u32(sStage + s * kBwStageBytes);
          if (t + 1 < n_tiles) {
            mbar_wait(&sdp_free, t & 1);
            issue_ss(t + 1);
          }
          mbar_wait(&p_ready, t & 1);
          tc_fence_after();
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < 4; ++k)       // dQ += dS K_j
              umma_ts(tmem_acc0, tmem_acc1 + (t & 1) * 32 + k * 8, umma_desc_sw128(a_addr + k * 2048, 0, 1024), idesc_t,
                      (t > 0 || k != 0) ? 1u : 0u);
            umma_commit(&empty[s]);
            if (t == n_tiles - 1) umma_commit(&acc_done);
          }
          __syncwarp();
        }
      } else {
        for (int t = 0; t < n_tiles; ++t) {
          const int s = t % kBwStages;
          const uint32_t a_addr = smem_u32(sStage + s * kBwStageBytes);
          const uint32_t b_addr = a_addr + kBwSTile;
          issue_ss(t);
          mbar_wait(&p_ready, t & 1);
          tc_fence_after();
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < 4; ++k)       // dV += P^T dO_i
              umma_ts(tmem_acc0, tmem_S + (k >> 1) * 32 + (k & 1) * 8, umma_desc_sw128(b_addr + k * 2048, 0, 1024), idesc_t,
                      (t > 0 || k != 0) ? 1u : 0u);
#pragma unroll
            for (int k = 0; k < 4; ++k)       // dK += dS^T Q_i
              umma_ts(tmem_acc1, tmem_dP + (k >> 1) * 32 + (k & 1) * 8, umma_desc_sw128(a_addr + k * 2048, 0, 1024), idesc_t,
                      (t > 0 || k != 0) ? 1u : 0u);
            umma_commit(&empty[s]);
            if (t == n_tiles - 1) umma_commit(&acc_done);
          }
          __syncwarp();
        }
      }
    }
  } else {
    // ------------------------------------------------------------------ softmax-gradient warps: TWO threads per TMEM lane
    // (warps 4-7 own score columns 0-31 of every tile, warps 8-11 columns 32-63).  Each half writes its bf16 P / dS back
    // into the first 16 columns of ITS OWN 32-column range, so the two warps of a lane quarter never touch each other's
    // columns: the A operand of the TS MMAs is read at column offsets 0, 8, 32, 40 (see the issuer).
    asm volatile("setmaxnreg.inc.sync.aligned.u32 88;");
    const int quarter = warp & 3;
    const int hlf = (warp - 4) >> 2;
    const int r_tile = quarter * 32 + lane;                 // stationary row inside the tile == TMEM lane
    const int x_row = x0 + r_tile;                          // row inside the segment (or the prefix)
    const bool row_ok = x_row < p.x_len;
    const uint32_t lane_off = static_cast<uint32_t>(quarter * 32) << 16;
    const uint32_t tS = tmem_S + lane_off + hlf * 32, tDP = tmem_dP + lane_off + hlf * 32;

    float lse_r = INFINITY, delta_r = 0.f;                  // mode 0: statistics of this thread's query row
    if (kMode == 0 && row_ok) {
      const int64_t so = (static_cast<int64_t>(head) * p.batch + b) * p.stat_pad + x_row;
      lse_r = p.lse[so];
      delta_r = p.delta[so];
    }

    for (int t = 0; t < n_tiles; ++t) {
      const int s = t % kBwStages;
      int valid = 64;                                       // mode 0: valid K/V columns of this tile
      if (kMode == 0) {
        if (t < p.n_prefix_tiles) valid = min(64, p.prefix_len - t * 64);
        else valid = min(64, p.own_len - (t - p.n_prefix_tiles) * 64);
      }
      const uint32_t stat = smem_u32(sStat + s * 2 * kBwStat);
      if (kMode == 1) mbar_wait(&full[s], (t / kBwStages) & 1);   // the bulk-copied statistics are visible to this thread
      mbar_wait(&s_full, t & 1);
      tc_fence_after();
      uint32_t pkp[16], pkd[16];
      if (kDebugBuild && p.dbg != 0) {
#pragma unroll
        for (int i = 0; i < 16; ++i) pkp[i] = pkd[i] = 0u;
      }
#pragma unroll
      for (int c = 0; c < 2; ++c) {                         // two chunks of 16 score columns
        if (kDebugBuild && (p.dbg & 2)) {
          if (kMode == 0 && c == 1 && elect_one()) mbar_arrive(&sdp_free);
          continue;
        }
        uint32_t sv[16], dv[16];
        tmem_ld16(tS + c * 16, sv);
        tmem_ld16(tDP + c * 16, dv);
        tmem_ld_wait();
        if (kMode == 0 && c == 1) {                         // S(t) / dP(t) are in registers: the issuer may overwrite them
          tc_fence_before();
          if (elect_one()) mbar_arrive(&sdp_free);
        }
        if (kDebugBuild && (p.dbg & 1)) { pkp[c] ^= sv[0] ^ dv[15]; continue; }
#pragma unroll
        for (int i = 0; i < 16; i += 4) {
          const int col = hlf * 32 + c * 16 + i;            // column inside the 64-wide tile
          float ls[4], dl[4];
          if (kMode == 1) {
            const float4 a = lds128(stat + col * 4);
            const float4 d = lds128(stat + kBwStat + col * 4);
            ls[0] = a.x; ls[1] = a.y; ls[2] = a.z; ls[3] = a.w;
            dl[0] = d.x; dl[1] = d.y; dl[2] = d.z; dl[3] = d.w;
          } else {
#pragma unroll
            for (int u = 0; u < 4; ++u) { ls[u] = lse_r; dl[u] = delta_r; }
          }
          // dS' = P (dP - delta): the softmax scale is applied once to the dQ / dK accumulators in the epilogue
          float pv[4], ds[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            pv[u] = ex2_approx(fmaf(__uint_as_float(sv[i + u]), p.scale_log2, -ls[u]));
            ds[u] = pv[u] * (__uint_as_float(dv[i + u]) - dl[u]);
          }
          if (kMode == 0 && col + 4 > valid) {              // only the last, partial K/V tile of a segment masks
#pragma unroll
            for (int u = 0; u < 4; ++u)
              if (col + u >= valid) pv[u] = ds[u] = 0.f;
          }
          pkp[c * 8 + (i >> 1)] = cvt_bf16x2(pv[0], pv[1]);
          pkp[c * 8 + (i >> 1) + 1] = cvt_bf16x2(pv[2], pv[3]);
          pkd[c * 8 + (i >> 1)] = cvt_bf16x2(ds[0], ds[1]);
          pkd[c * 8 + (i >> 1) + 1] = cvt_bf16x2(ds[2], ds[3]);
        }
      }
      if (kMode == 1) {
        tmem_st16(tS, pkp);                                 // P^T over the first half of this warp's own S^T columns
        tmem_st16(tDP, pkd);                                // dS^T likewise over dP^T
      } else {
        tmem_st16(tmem_acc1 + lane_off + (t & 1) * 32 + hlf * 16, pkd);   // dS(t): buffer t & 1 of the spare accumulator columns
      }
      tmem_st_wait();
      tc_fence_before();
      if (elect_one()) mbar_arrive(&p_ready);
    }

    // ------------------------------------------------------------------ epilogue: fp32 accumulators -> bf16 rows
    mbar_wait(&acc_done, 0);
    tc_fence_after();
    const int64_t g_row = static_cast<int64_t>(b) * p.x_batch_stride + x_row;
#pragma unroll
    for (int which = 0; which < (kMode == 0 ? 1 : 2); ++which) {
      __nv_bfloat16* base = which == 0 ? p.d0 : p.d1;
      uint32_t o[32];
      tmem_ld32((which == 0 ? tmem_acc0 : tmem_acc1) + lane_off + hlf * 32, o);
      tmem_ld_wait();
      if (kMode == 0 || which == 1) {                       // dQ / dK were accumulated from the unscaled dS'
#pragma unroll
        for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * p.scale);
      }
      if (row_ok) {
        __nv_bfloat16* dst = base + g_row * p.ldd + head * 64 + hlf * 32;
#pragma unroll
        for (int i = 0; i < 32; i += 8) {
          uint4 w;
          w.x = pack_bf16x2(__uint_as_float(o[i + 0]), __uint_as_float(o[i + 1]));
          w.y = pack_bf16x2(__uint_as_float(o[i + 2]), __uint_as_float(o[i + 3]));
          w.z = pack_bf16x2(__uint_as_float(o[i + 4]), __uint_as_float(o[i + 5]));
          w.w = pack_bf16x2(__uint_as_float(o[i + 6]), __uint_as_float(o[i + 7]));
          *reinterpret_cast<uint4*>(dst + i) = w;
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<256>(tmem_base);
  }
}

// delta[h, b, r] = sum_d out[row, 64h + d] * dout[row, 64h + d]: one thread per (row, head), 128-byte lines.
__global__ void attn_bwd_delta_kernel(const __nv_bfloat16* __restrict__ out, int64_t ldo, const __nv_bfloat16* __restrict__ dout,
                                      int64_t lddo, float* __restrict__ delta, int num_heads, int batch, int q_len,
                                      int q_batch_stride, int stat_pad) {
  const int64_t idx = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const int64_t total = static_cast<int64_t>(batch) * q_len * num_heads;
  if (idx >= total) return;
  const int h = static_cast<int>(idx % num_heads);
  const int64_t rq = idx / num_heads;
  const int r = static_cast<int>(rq % q_len);
  const int b = static_cast<int>(rq / q_len);
  const int64_t row = static_cast<int64_t>(b) * q_batch_stride + r;
  const uint4* po = reinterpret_cast<const uint4*>(out + row * ldo + h * 64);
  const uint4* pd = reinterpret_cast<const uint4*>(dout + row * lddo + h * 64);
  float acc = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const uint4 a = po[i], d = pd[i];
    const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, dw[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      acc = fmaf(__uint_as_float(aw[u] << 16), __uint_as_float(dw[u] << 16), acc);
      acc = fmaf(__uint_as_float(aw[u] & 0xffff0000u), __uint_as_float(dw[u] & 0xffff0000u), acc);
    }
  }
  delta[(static_cast<int64_t>(h) * batch + b) * stat_pad + r] = acc;
}

template <int kMode>
static int launch_bwd(dim3 grid, cudaStream_t st, const CUtensorMap& tx, const CUtensorMap& ty, const CUtensorMap& ta,
                      const CUtensorMap& tb, const CUtensorMap& tpa, const CUtensorMap& tpb, const BwdParams& p) {
  auto kern = attn_bwd_kernel<kMode>;
  static bool configured = false;
  if (!configured) {
    SR_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kBwSmem));
    configured = true;
  }
  kern<<<grid, kBwThreads, kBwSmem, st>>>(tx, ty, ta, tb, tpa, tpb, p);
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

// 3-D bf16 map over rows of `base` ([hd columns, len rows, nseg segments]), box = 64 columns x box_rows rows.
static int seg_tmap(CUtensorMap* out, const void* base, int64_t hd, int64_t len, int64_t nseg, int64_t ld, int64_t seg_stride,
                    uint32_t box_rows) {
  uint32_t box[3] = {64, box_rows, 1};
  uint64_t dims[3] = {static_cast<uint64_t>(hd), static_cast<uint64_t>(len), static_cast<uint64_t>(nseg)};
  uint64_t str[2] = {static_cast<uint64_t>(ld) * 2, static_cast<uint64_t>(seg_stride) * static_cast<uint64_t>(ld) * 2};
  return make_tmap_bf16(out, base, 3, dims, str, box);
}

}  // namespace sr

using namespace sr;

extern "C" int sr_attn_bwd_bf16(const sr_attn_bwd_t* a, void* stream) {
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(a && a->q && a->k && a->v && a->out && a->dout && a->lse && a->delta && a->dq && a->dk && a->dv,
             "sr_attn_bwd_bf16: null pointer");
  SR_REQUIRE(a->head_dim == 64, "sr_attn_bwd_bf16: head_dim must be 64 (got %d)", a->head_dim);
  SR_REQUIRE(a->num_heads > 0 && a->batch > 0 && a->q_len > 0 && a->kv_len > 0, "sr_attn_bwd_bf16: empty problem");
  SR_REQUIRE(a->batch <= 65535 && a->num_heads <= 65535, "sr_attn_bwd_bf16: batch / heads exceed grid limits");
  const int64_t hd = static_cast<int64_t>(a->num_heads) * 64;
  const bool has_prefix = a->pk != nullptr && a->prefix_len > 0;
  auto ok_ld = [&](int64_t ld) { return ld >= hd && ld % 8 == 0; };
  SR_REQUIRE(ok_ld(a->ldq) && ok_ld(a->ldkv) && ok_ld(a->ldo) && ok_ld(a->lddo) && ok_ld(a->lddq) && ok_ld(a->lddkv),
             "sr_attn_bwd_bf16: row strides must cover heads*64 columns and be multiples of 8");
  uintptr_t bits = reinterpret_cast<uintptr_t>(a->q) | reinterpret_cast<uintptr_t>(a->k) | reinterpret_cast<uintptr_t>(a->v) |
                   reinterpret_cast<uintptr_t>(a->out) | reinterpret_cast<uintptr_t>(a->dout) | reinterpret_cast<uintptr_t>(a->dq) |
                   reinterpret_cast<uintptr_t>(a->dk) | reinterpret_cast<uintptr_t>(a->dv) | reinterpret_cast<uintptr_t>(a->lse) |
                   reinterpret_cast<uintptr_t>(a->delta);
  if (has_prefix) {
    SR_REQUIRE(a->pv && ok_ld(a->ldp), "sr_attn_bwd_bf16: bad prefix K/V");
    bits |= reinterpret_cast<uintptr_t>(a->pk) | reinterpret_cast<uintptr_t>(a->pv);
    if (a->dpk) {
      SR_REQUIRE(a->dpv && ok_ld(a->lddp), "sr_attn_bwd_bf16: bad prefix gradient buffers");
      bits |= reinterpret_cast<uintptr_t>(a->dpk) | reinterpret_cast<uintptr_t>(a->dpv);
    }
  }
  SR_REQUIRE((bits & 15) == 0, "sr_attn_bwd_bf16: pointers must be 16-byte aligned");
  SR_REQUIRE(a->lse_pad >= a->q_len && a->lse_pad % 64 == 0, "sr_attn_bwd_bf16: lse_pad must be a multiple of 64 >= q_len");
  SR_REQUIRE(a->batch == 1 || (a->q_batch_stride >= a->q_len && a->kv_batch_stride >= a->kv_len),
             "sr_attn_bwd_bf16: segment strides smaller than the segment lengths");
  SR_REQUIRE(a->q_len < (1ll << 30) && a->kv_len < (1ll << 30) && a->prefix_len < (1ll << 30), "sr_attn_bwd_bf16: sequence too long");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int64_t qbs = a->batch > 1 ? a->q_batch_stride : a->q_len;
  const int64_t kbs = a->batch > 1 ? a->kv_batch_stride : a->kv_len;

  // 1. delta = rowsum(out * dout) per head (padding rows stay 0)
  const size_t stat_bytes = static_cast<size_t>(a->num_heads) * a->batch * a->lse_pad * sizeof(float);
  SR_CHECK_CUDA(cudaMemsetAsync(a->delta, 0, stat_bytes, st));
  {
    const int64_t total = static_cast<int64_t>(a->batch) * a->q_len * a->num_heads;
    attn_bwd_delta_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, st>>>(
        reinterpret_cast<const __nv_bfloat16*>(a->out), a->ldo, reinterpret_cast<const __nv_bfloat16*>(a->dout), a->lddo, a->delta,
        a->num_heads, a->batch, static_cast<int>(a->q_len), static_cast<int>(qbs), static_cast<int>(a->lse_pad));
    ++g_launch_count;
    SR_CHECK_CUDA(cudaGetLastError());
  }

  CUtensorMap q_st, do_st, q_sm, do_sm, k_st, v_st, k_sm, v_sm, pk_st, pv_st, pk_sm, pv_sm;
  if (int rc = seg_tmap(&q_st, a->q, hd, a->q_len, a->batch, a->ldq, qbs, 128)) return rc;
  if (int rc = seg_tmap(&q_sm, a->q, hd, a->q_len, a->batch, a->ldq, qbs, 64)) return rc;
  if (int rc = seg_tmap(&do_st, a->dout, hd, a->q_len, a->batch, a->lddo, qbs, 128)) return rc;
  if (int rc = seg_tmap(&do_sm, a->dout, hd, a->q_len, a->batch, a->lddo, qbs, 64)) return rc;
  if (int rc = seg_tmap(&k_st, a->k, hd, a->kv_len, a->batch, a->ldkv, kbs, 128)) return rc;
  if (int rc = seg_tmap(&v_st, a->v, hd, a->kv_len, a->batch, a->ldkv, kbs, 128)) return rc;
  if (int rc = seg_tmap(&k_sm, a->k, hd, a->kv_len, a->batch, a->ldkv, kbs, 64)) return rc;
  if (int rc = seg_tmap(&v_sm, a->v, hd, a->kv_len, a->batch, a->ldkv, kbs, 64)) return rc;
  if (has_prefix) {
    if (int rc = seg_tmap(&pk_st, a->pk, hd, a->prefix_len, 1, a->ldp, a->prefix_len, 128)) return rc;
    if (int rc = seg_tmap(&pv_st, a->pv, hd, a->prefix_len, 1, a->ldp, a->prefix_len, 128)) return rc;
    if (int rc = seg_tmap(&pk_sm, a->pk, hd, a->prefix_len, 1, a->ldp, a->prefix_len, 64)) return rc;
    if (int rc = seg_tmap(&pv_sm, a->pv, hd, a->prefix_len, 1, a->ldp, a->prefix_len, 64)) return rc;
  } else {
    pk_st = k_st; pv_st = v_st; pk_sm = k_sm; pv_sm = v_sm;
  }

  BwdParams p{};
  p.lse = a->lse;
  p.delta = a->delta;
  p.stat_pad = static_cast<int>(a->lse_pad);
  p.batch = a->batch;
  p.scale = a->scale;
  p.scale_log2 = a->scale * 1.4426950408889634f;
  {
    static const int dbg = [] { const char* e = dbg_env("SR_ATTN_BWD_DBG"); return e ? atoi(e) : 0; }();
    p.dbg = dbg;
  }

  // 2. dq: query-stationary pass over [prefix | own] K/V tiles
  p.d0 = reinterpret_cast<__nv_bfloat16*>(a->dq);
  p.d1 = nullptr;
  p.ldd = a->lddq;
  p.x_len = static_cast<int>(a->q_len);
  p.x_batch_stride = static_cast<int>(qbs);
  p.prefix_len = has_prefix ? static_cast<int>(a->prefix_len) : 0;
  p.n_prefix_tiles = (p.prefix_len + 63) / 64;
  p.own_len = static_cast<int>(a->kv_len);
  p.n_own_tiles = (p.own_len + 63) / 64;
  p.over_prefix = 0;
  {
    dim3 grid(static_cast<unsigned>((a->q_len + 127) / 128), static_cast<unsigned>(a->num_heads), static_cast<unsigned>(a->batch));
    if (int rc = launch_bwd<0>(grid, st, q_st, do_st, k_sm, v_sm, pk_sm, pv_sm, p)) return rc;
  }
  // 3. dk, dv of the segments' own rows: key-stationary pass over the segment's query tiles
  p.d0 = reinterpret_cast<__nv_bfloat16*>(a->dv);
  p.d1 = reinterpret_cast<__nv_bfloat16*>(a->dk);
  p.ldd = a->lddkv;
  p.x_len = static_cast<int>(a->kv_len);
  p.x_batch_stride = static_cast<int>(kbs);
  p.prefix_len = 0;
  p.n_prefix_tiles = 0;
  p.own_len = static_cast<int>(a->q_len);
  p.n_own_tiles = (p.own_len + 63) / 64;
  {
    dim3 grid(static_cast<unsigned>((a->kv_len + 127) / 128), static_cast<unsigned>(a->num_heads), static_cast<unsigned>(a->batch));
    if (int rc = launch_bwd<1>(grid, st, k_st, v_st, q_sm, do_sm, q_sm, do_sm, p)) return rc;
  }
  // 4. gradient of the shared prefix: key-stationary over the query tiles of every segment
  if (has_prefix && a->dpk) {
    p.d0 = reinterpret_cast<__nv_bfloat16*>(a->dpv);
    p.d1 = reinterpret_cast<__nv_bfloat16*>(a->dpk);
    p.ldd = a->lddp;
    p.x_len = static_cast<int>(a->prefix_len);
    p.x_batch_stride = 0;
    p.over_prefix = 1;
    dim3 grid(static_cast<unsigned>((a->prefix_len + 127) / 128), static_cast<unsigned>(a->num_heads), 1);
    if (int rc = launch_bwd<1>(grid, st, pk_st, pv_st, q_sm, do_sm, q_sm, do_sm, p)) return rc;
  }
  return SR_OK;
}
