// Flash attention forward on tcgen05, head_dim 64 — FOUR query tiles per CTA over 64-row K/V tiles (sm_100a).
//
// Same contract as attn_tc.cu (replaces F.scaled_dot_product_attention at sailrecon/layers/attention.py:103-109 with the
// structured [scene prefix | own frame] K/V walk of sailrecon/models/aggregator.py:302-311,832-847), different schedule.
// Why: the two-tile kernel is bound by the serial per-warp softmax chain (TMEM load -> row max -> 128 exponentials ->
// P store): with two softmax warps per sub-partition MUFU is busy only 79 % of the time.  Here every
// sub-partition runs FOUR independent softmax streams (one per query tile), each on half-width tiles:
//   TMEM (512 columns)  S0..S3 (64 fp32 columns each; P_q, 64 bf16 = 32 columns, is written back over S_q) | O0..O3 (64 each)
//   smem                Q 4 x 16 KB | ring of 6 x (V(j) 8 KB + K(j+1) 8 KB) | K(0) 8 KB                  (169 KB)
//   warp 0              TMA producer (Q tiles once, then 64 x 64 K and V tiles)
//   warps 1, 3          MMA issuers (query tiles 0-1 and 2-3), each round-robin over its tiles:
//                           wait p_full[q](j)  ->  O_q += P_q(j) V(j)      TS, M=128 N=64 K=16 x4
//                                                  S_q  = Q_q K(j+1)^T     SS, M=128 N=64 K=16 x4  (after the PV: it overwrites P_q)
//                       the in-order tensor pipe makes "S_q(j+1) complete" imply "PV_q(j) complete", so the softmax warps
//                       need no pv_done / s_free barriers at all
//   warp 2              TMEM allocator
//   warps 4-19          softmax: warp 4 + 4q + r owns rows [32r, 32r + 32) of query tile q, one thread per row
// 640 threads; registers re-distributed with setmaxnreg (control 56, softmax 104).
#include <cstdlib>
#include <algorithm>
#include <functional>
#include <mutex>
#include <queue>
#include <vector>

#include "attn_math.cuh"
#include "common.cuh"
#include "sr_internal.h"

namespace sr {
extern int g_launch_count;
static __device__ long long g_attn_trace[3 * 64 * 8];   // SR_ATTN_TRACE: [role][tile][event] clock64() of CTA (0,0,0)
#define SR_TRACE4(role, j, ev)                                                          \
  do {                                                                                  \
    if (trace_on && (j) < 64) g_attn_trace[((role) * 64 + (j)) * 8 + (ev)] = clock64(); \
  } while (0)

constexpr int k4NQ = 4;
constexpr int k4Stages = 6;
constexpr int k4QTile = 128 * 64 * 2;                 // 16 KB: 128 x 64 bf16
constexpr int k4KvTile = 64 * 64 * 2;                 //  8 KB:  64 x 64 bf16
constexpr int k4Threads = 128 + k4NQ * 128;
constexpr int k4Smem = k4NQ * k4QTile + (k4Stages * 2 + 1) * k4KvTile + 1024;
constexpr float k4LazyRescaleLog2 = 8.0f;

struct Attn4Params {
  __nv_bfloat16* out;
  int64_t ldo;
  float* carry_m;
  float* carry_l;
  float* carry_acc;
  int carry_in, carry_out;
  int num_heads;
  int q_len, q_batch_stride;     // rows
  int kv_len;
  int prefix_len;
  int n_prefix_tiles, n_own_tiles;   // 64-row tiles
  int64_t total_q_rows;
  float scale_log2;
  int trace;
  float* lse;                        // training: [num_heads, batch, lse_pad] log2-domain log-sum-exp (nullptr = off)
  int lse_pad;
  // Work decomposition (1-D grid).  A "row" is one (head, segment) pair; its query tiles are covered by n4 CTAs of four
  // tiles followed by n3 CTAs of three tiles.  ALL four-tile CTAs of the launch come first in the grid, the three-tile
  // CTAs last: the hardware dispatches CTAs in linear order onto SMs as they drain, so the lighter CTAs fill the tail wave
  // (16-anchor global shape: 172 tiles per head = 37 x 4 + 8 x 3 -> 592 heavy CTAs = exactly 4 waves of 148 SMs + 128 light
  // ones = 19 tile-steps of makespan instead of the 20 of 688 uniform four-tile CTAs; chosen per shape by choose_mix()).
  int n_rows, n4, n3, batch;
  // row_major = 1: a row's CTAs are adjacent in the grid (its K/V stay L2-resident between them) — used when the launch has
  // at least one row per SM, where K/V of all rows exceed the L2 and the tail wave is short anyway (frame / reloc shapes:
  // the heavy-first order made every light CTA re-read its row's K/V from HBM, +50 % DRAM traffic, profiles/r02_ncu_attn_tc.txt)
  int row_major;
  // Split-KV (under-filled launches, e.g. ONE query frame against a 100-anchor scene prefix = 48 CTAs on 148 SMs): the
  // K/V tile range is cut into n_split chunks, each (row, query-group, chunk) is its own CTA writing its partial
  // (m, l, un-normalised O) into per-chunk carry buffers inside the caller's workspace; attn_split_merge_kernel combines them.
  int n_split;
  int64_t split_stride_ml, split_stride_acc;   // floats between consecutive chunks' (m | l) and acc buffers
};

template <int kPoly>
__global__ void __launch_bounds__(k4Threads, 1)
attn_tc4_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
                const __grid_constant__ CUtensorMap tmap_v, const __grid_constant__ CUtensorMap tmap_pk,
                const __grid_constant__ CUtensorMap tmap_pv, const Attn4Params p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t q_full;
  __shared__ __align__(8) uint64_t kv_full[k4Stages], kv_empty[k4Stages];
  __shared__ __align__(8) uint64_t s_full[k4NQ], p_full[k4NQ], o_done[k4NQ];
  __shared__ uint32_t tmem_base_slot;

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);   // provably warp-uniform (see elect_one)
  const int lane = threadIdx.x & 31;
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + k4NQ * k4QTile;
  uint8_t* sV = sK + k4Stages * k4KvTile;
  uint8_t* sK0 = sV + k4Stages * k4KvTile;                           // K(0)

  int row, q0, nq_cap;                                               // see Attn4Params: heavy CTAs first, light CTAs last
  const int base_ctas = p.n_rows * (p.n4 + p.n3);
  const int split = blockIdx.x / base_ctas;                          // 0 unless the launch is split over the K/V range
  {
    const int lin = blockIdx.x - split * base_ctas, heavy = p.n_rows * p.n4;
    if (p.row_major) {
      const int per_row = p.n4 + p.n3;
      row = lin / per_row;
      const int x = lin - row * per_row;
      q0 = x < p.n4 ? x * 512 : p.n4 * 512 + (x - p.n4) * 384;
      nq_cap = x < p.n4 ? 4 : 3;
    } else if (lin < heavy) {
      row = lin / p.n4;
      q0 = (lin - row * p.n4) * 512;
      nq_cap = 4;
    } else {
      const int r = lin - heavy;
      row = r / p.n3;
      q0 = p.n4 * 512 + (r - row * p.n3) * 384;
      nq_cap = 3;
    }
  }
  const int b = row / p.num_heads;
  const int head = row - b * p.num_heads;
  const int nq_valid = min(nq_cap, (p.q_len - q0 + 127) / 128);      // query tiles that hold at least one row
  const int n_all = p.n_prefix_tiles + p.n_own_tiles;                // K/V tiles of the whole problem
  const int t_chunk = (n_all + p.n_split - 1) / p.n_split;
  const int t0 = split * t_chunk;                                    // this CTA walks tiles [t0, t0 + n_tiles)
  const int n_tiles = min(n_all, t0 + t_chunk) - t0;
  float* const carry_m = p.carry_m + split * p.split_stride_ml;
  float* const carry_l = p.carry_l + split * p.split_stride_ml;
  float* const carry_acc = p.carry_acc + split * p.split_stride_acc;
  const bool trace_on = kDebugBuild && p.trace && blockIdx.x == 0 && lane == 0;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_q);
    tma_prefetch_desc(&tmap_k);
    tma_prefetch_desc(&tmap_v);
    if (p.n_prefix_tiles) {
      tma_prefetch_desc(&tmap_pk);
      tma_prefetch_desc(&tmap_pv);
    }
  }
  if (warp == 1 && lane == 0) {
    mbar_init(&q_full, 1);
    for (int s = 0; s < k4Stages; ++s) {
      mbar_init(&kv_full[s], 1);
      mbar_init(&kv_empty[s], nq_valid > 2 ? 2 : 1);   // one tcgen05.commit per active MMA issuer
    }
    for (int q = 0; q < k4NQ; ++q) {
      mbar_init(&s_full[q], 1);
      mbar_init(&p_full[q], 4);              // one elected arrive per softmax warp
      mbar_init(&o_done[q], 1);
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<512>(&tmem_base_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_slot;
  const uint32_t tmem_S0 = tmem_base;                         // + q * 64   (P_q aliases the first 32 columns of S_q)
  const uint32_t tmem_O0 = tmem_base + k4NQ * 64;             // + q * 64

  if (warp < 4) {
  asm volatile("setmaxnreg.dec.sync.aligned.u32 56;");
  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer (whole warp; one elected lane issues)
    // Step j of the MMA issuers needs V(j) and K(j+1): both travel on ONE barrier into stage j % k4Stages (one wait and
    // one release per step instead of two each); K(0) has its own buffer and barrier.
    auto load_tile = [&](uint8_t* dst, uint64_t* bar, int t, bool is_k) {
      const bool pre = t < p.n_prefix_tiles;
      const int r0 = (pre ? t : t - p.n_prefix_tiles) * 64;
      if (pre) tma_load_2d(dst, is_k ? &tmap_pk : &tmap_pv, bar, head * 64, r0);
      else     tma_load_3d(dst, is_k ? &tmap_k : &tmap_v, bar, head * 64, r0, b);
    };
    if (elect_one()) {
      mbar_arrive_expect_tx(&q_full, nq_valid * k4QTile + k4KvTile);
      for (int q = 0; q < nq_valid; ++q) tma_load_3d(sQ + q * k4QTile, &tmap_q, &q_full, head * 64, q0 + q * 128, b);
      load_tile(sK0, &q_full, t0, true);
    }
    __syncwarp();
    for (int j = 0; j < n_tiles; ++j) {
      const int s = j % k4Stages;
      mbar_wait_relaxed(&kv_empty[s], ((j / k4Stages) & 1) ^ 1);
      if (elect_one()) {
        const bool has_k = j + 1 < n_tiles;
        mbar_arrive_expect_tx(&kv_full[s], has_k ? 2 * k4KvTile : k4KvTile);
        load_tile(sV + s * k4KvTile, &kv_full[s], t0 + j, false);
        if (has_k) load_tile(sK + s * k4KvTile, &kv_full[s], t0 + j + 1, true);
      }
      __syncwarp();
    }
  } else if (warp == 1 || warp == 3) {
    // ------------------------------------------------------------------ MMA issuers: warp 1 serves query tiles 0-1,
    // warp 3 tiles 2-3 (whole warp runs the loop; one elected lane issues).  One issuer for all four tiles was the
    // bottleneck: ~560 clk per tile served (barrier wait + 8 MMAs + commits on a sub-partition it shares with four busy
    // softmax warps) = 2780 clk per step against 2048 clk of MUFU work.
    const int qa = warp == 1 ? 0 : 2;                        // first query tile of this issuer
    const int nq_mine = min(2, nq_valid - qa);               // <= 0: nothing to do
    if (nq_mine > 0) {
      constexpr uint32_t idesc_s = umma_idesc_bf16(128, 64, 0, 0);   // Q (K-major) x K (K-major)
      constexpr uint32_t idesc_o = umma_idesc_bf16(128, 64, 0, 1);   // P (TMEM)    x V (MN-major: d contiguous)
      auto issue_s = [&](int q, uint32_t k_addr) {
        const uint64_t dq = umma_desc_sw128(smem_u32(sQ + q * k4QTile), 16, 1024);
        const uint64_t db = umma_desc_sw128(k_addr, 16, 1024);
        const uint32_t tS = tmem_S0 + q * 64;
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_ss(tS, dq + 2 * k, db + 2 * k, idesc_s, k != 0 ? 1u : 0u);
        umma_commit(&s_full[q]);
      };
      auto issue_pv = [&](int q, uint32_t v_addr, bool accumulate) {
        const uint32_t tO = tmem_O0 + q * 64, tP = tmem_S0 + q * 64;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          // A: P columns [8k, 8k+8) hold kv [16k, 16k+16) as packed bf16 pairs
          // B: V rows [16k, 16k+16) = two 8-row swizzle atoms (SBO 1024 B); one 64-wide MN block (LBO unused)
          const uint64_t db = umma_desc_sw128(v_addr + k * 2048, 0, 1024);
          umma_ts(tO, tP + k * 8, db, idesc_o, (accumulate || k != 0) ? 1u : 0u);
        }
      };

      mbar_wait(&q_full, 0);
      tc_fence_after();
      if (elect_one()) {
        for (int qi = 0; qi < nq_mine; ++qi) issue_s(qa + qi, smem_u32(sK0));
      }
      __syncwarp();
      for (int j = 0; j < n_tiles; ++j) {
        const int s = j % k4Stages;
        const bool has_next = j + 1 < n_tiles;
        mbar_wait(&kv_full[s], (j / k4Stages) & 1);
        if (qa == 0) SR_TRACE4(1, j, 0);
#pragma unroll
        for (int qi = 0; qi < 2; ++qi) {
          if (qi < nq_mine) {
            const int q = qa + qi;
            mbar_wait(&p_full[q], j & 1);            // P_q(j) in TMEM, O_q rescaled
            if (qa == 0) SR_TRACE4(1, j, 1 + qi);    // events 1..2: p_full[q] seen
            tc_fence_after();
            if (elect_one()) {
              issue_pv(q, smem_u32(sV + s * k4KvTile), p.carry_in != 0 || j > 0);
              if (has_next) issue_s(q, smem_u32(sK + s * k4KvTile));
              else umma_commit(&o_done[q]);
              if (qi == nq_mine - 1) umma_commit(&kv_empty[s]);
            }
            __syncwarp();
          }
        }
        if (qa == 0) SR_TRACE4(1, j, 5);
      }
    }
  }
  } else {
    // ------------------------------------------------------------------ softmax warps: one thread = one query row
    asm volatile("setmaxnreg.inc.sync.aligned.u32 104;");
    const int q = (warp - 4) >> 2;
    if (q < nq_valid) {
      const int quarter = warp & 3;
      const int r_tile = quarter * 32 + lane;                 // row inside the 128-row tile == TMEM lane
      const int q_row = q0 + q * 128 + r_tile;                // row inside the segment
      const bool row_ok = q_row < p.q_len;
      const int64_t g_row = static_cast<int64_t>(b) * p.q_batch_stride + q_row;   // row in the q / out matrices
      const uint32_t lane_off = static_cast<uint32_t>(quarter * 32) << 16;
      const uint32_t tS = tmem_S0 + q * 64 + lane_off;        // also P_q (first 32 columns)
      const uint32_t tO = tmem_O0 + q * 64 + lane_off;

      float m = -INFINITY, l = 0.f;
      if (p.carry_in) {
        if (row_ok) {
          m = carry_m[static_cast<int64_t>(head) * p.total_q_rows + g_row];
          l = carry_l[static_cast<int64_t>(head) * p.total_q_rows + g_row];
        }
        // seed O with the carried un-normalised accumulator
        const float* src = carry_acc + g_row * (static_cast<int64_t>(p.num_heads) * 64) + head * 64;
#pragma unroll
        for (int hlf = 0; hlf < 2; ++hlf) {
          uint32_t o[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) o[i] = row_ok ? __float_as_uint(src[hlf * 32 + i]) : 0u;
          tmem_st32(tO + hlf * 32, o);
        }
        tmem_st_wait();
      }
      const uint64_t scale2 = pack2(p.scale_log2, p.scale_log2);

      // valid kv columns of tile j (only the last tile of the prefix / own range is partial)
      auto valid_cols = [&](int t) {                       // t = tile index in the whole problem
        if (t < p.n_prefix_tiles) return t == p.n_prefix_tiles - 1 ? p.prefix_len - t * 64 : 64;
        return t == n_all - 1 ? p.kv_len - (t - p.n_prefix_tiles) * 64 : 64;
      };
      uint32_t sraw[64];
      auto row_max = [&](int valid) {
        if (valid < 64) {
#pragma unroll
          for (int i = 0; i < 64; ++i)
            if (i >= valid) sraw[i] = 0xff800000u;   // -inf
        }
        float mx0 = max3(__uint_as_float(sraw[0]), __uint_as_float(sraw[1]), __uint_as_float(sraw[2]));
        float mx1 = max3(__uint_as_float(sraw[3]), __uint_as_float(sraw[4]), __uint_as_float(sraw[5]));
        float mx2 = max3(__uint_as_float(sraw[6]), __uint_as_float(sraw[7]), __uint_as_float(sraw[8]));
        float mx3 = max3(__uint_as_float(sraw[9]), __uint_as_float(sraw[10]), __uint_as_float(sraw[11]));
#pragma unroll
        for (int i = 12; i < 60; i += 8) {
          mx0 = max3(mx0, __uint_as_float(sraw[i + 0]), __uint_as_float(sraw[i + 1]));
          mx1 = max3(mx1, __uint_as_float(sraw[i + 2]), __uint_as_float(sraw[i + 3]));
          mx2 = max3(mx2, __uint_as_float(sraw[i + 4]), __uint_as_float(sraw[i + 5]));
          mx3 = max3(mx3, __uint_as_float(sraw[i + 6]), __uint_as_float(sraw[i + 7]));
        }
        mx0 = max3(mx0, __uint_as_float(sraw[60]), __uint_as_float(sraw[61]));
        mx1 = max3(mx1, __uint_as_float(sraw[62]), __uint_as_float(sraw[63]));
        return fmaxf(max3(mx0, mx1, mx2), mx3);
      };

      const int tr_role = warp == 4 ? 0 : (warp == 16 ? 2 : -1);   // softmax of query tile 0 / 3
      for (int j = 0; j < n_tiles; ++j) {
        mbar_wait(&s_full[q], j & 1);            // S_q(j) complete; implies PV_q(j-1) complete (in-order tensor pipe)
        if (tr_role >= 0) SR_TRACE4(tr_role, j, 0);
        tc_fence_after();
        tmem_ld32(tS, *reinterpret_cast<uint32_t(*)[32]>(&sraw[0]));
        tmem_ld32(tS + 32, *reinterpret_cast<uint32_t(*)[32]>(&sraw[32]));
        tmem_ld_wait();
        const float mx = row_max(valid_cols(t0 + j)) * p.scale_log2;       // scale > 0
        if (tr_role >= 0 && mx == mx) SR_TRACE4(tr_role, j, 2);
        float m_new = m;
        if (mx > m + k4LazyRescaleLog2) m_new = mx;                        // also the first tile (m = -inf)
        const float alpha = ex2_approx(m - m_new);                         // 1 when unchanged, 0 on the first tile
        const uint64_t negm2 = pack2(-m_new, -m_new);
        uint64_t sum0 = 0, sum1 = 0, sum2 = 0, sum3 = 0;                   // packed (0.f, 0.f)
        uint32_t pk[32];
#pragma unroll
        for (int i = 0; i < 64; i += 8) {
          float e[8];
#pragma unroll
          for (int t = 0; t < 4; ++t) {
            const uint64_t x2 = fma2(pack2(__uint_as_float(sraw[i + 2 * t]), __uint_as_float(sraw[i + 2 * t + 1])), scale2, negm2);
            if (kPoly == 8) {   // timing experiment: no exponential at all
              unpack2(x2, e[2 * t], e[2 * t + 1]);
            } else if ((((i >> 1) + t) & 7) < kPoly) {
              ex2_poly2(x2, e[2 * t], e[2 * t + 1]);
            } else {
              float x0, x1;
              unpack2(x2, x0, x1);
              e[2 * t] = ex2_approx(x0);
              e[2 * t + 1] = ex2_approx(x1);
            }
            pk[(i >> 1) + t] = cvt_bf16x2(e[2 * t], e[2 * t + 1]);
          }
          sum0 = add2(sum0, pack2(e[0], e[1]));
          sum1 = add2(sum1, pack2(e[2], e[3]));
          sum2 = add2(sum2, pack2(e[4], e[5]));
          sum3 = add2(sum3, pack2(e[6], e[7]));
        }
        float r0, r1;
        unpack2(add2(add2(sum0, sum1), add2(sum2, sum3)), r0, r1);
        l = l * alpha + (r0 + r1);
        if (tr_role >= 0 && l >= 0.f) SR_TRACE4(tr_role, j, 3);

        tmem_st32(tS, pk);                       // P_q(j) over the first 32 columns of S_q (this warp's lanes are in registers)
        const bool have_acc = (j > 0) || p.carry_in;
        if (have_acc && __any_sync(0xffffffffu, m_new != m)) {
          uint32_t o[32];
#pragma unroll
          for (int hlf = 0; hlf < 2; ++hlf) {
            tmem_ld32(tO + hlf * 32, o);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
            tmem_st32(tO + hlf * 32, o);
          }
        }
        tmem_st_wait();
        m = m_new;
        tc_fence_before();
        if (elect_one()) mbar_arrive(&p_full[q]);
        if (tr_role >= 0) SR_TRACE4(tr_role, j, 5);
      }

      // ---------------------------------------------------------------- epilogue
      mbar_wait(&o_done[q], 0);
      tc_fence_after();
      const float inv = 1.0f / l;
#pragma unroll
      for (int hlf = 0; hlf < 2; ++hlf) {
        uint32_t o[32];
        tmem_ld32(tO + hlf * 32, o);
        tmem_ld_wait();
        if (row_ok) {
          if (p.carry_out) {
            float* dst = carry_acc + g_row * (static_cast<int64_t>(p.num_heads) * 64) + head * 64 + hlf * 32;
#pragma unroll
            for (int i = 0; i < 32; i += 4) *reinterpret_cast<uint4*>(dst + i) = make_uint4(o[i], o[i + 1], o[i + 2], o[i + 3]);
          } else {
            __nv_bfloat16* dst = p.out + g_row * p.ldo + head * 64 + hlf * 32;
#pragma unroll
            for (int i = 0; i < 32; i += 8) {
              uint4 w;
              w.x = pack_bf16x2(__uint_as_float(o[i + 0]) * inv, __uint_as_float(o[i + 1]) * inv);
              w.y = pack_bf16x2(__uint_as_float(o[i + 2]) * inv, __uint_as_float(o[i + 3]) * inv);
              w.z = pack_bf16x2(__uint_as_float(o[i + 4]) * inv, __uint_as_float(o[i + 5]) * inv);
              w.w = pack_bf16x2(__uint_as_float(o[i + 6]) * inv, __uint_as_float(o[i + 7]) * inv);
              *reinterpret_cast<uint4*>(dst + i) = w;
            }
          }
        }
      }
      if (row_ok && p.lse != nullptr && p.n_split == 1)
        p.lse[(static_cast<int64_t>(head) * p.batch + b) * p.lse_pad + q_row] = m + log2f(l);
      if (row_ok && p.carry_out) {
        carry_m[static_cast<int64_t>(head) * p.total_q_rows + g_row] = m;
        carry_l[static_cast<int64_t>(head) * p.total_q_rows + g_row] = l;
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<512>(tmem_base);
  }
}

// Combines the n_split partial results of a split-KV launch: one warp per (segment, query row, head), lane = 2 of the 64
// output features.  m, l in the log2 domain; acc un-normalised.
__global__ void __launch_bounds__(256)
attn_split_merge_kernel(const float* __restrict__ ws_m, const float* __restrict__ ws_l, const float* __restrict__ ws_acc,
                        int n_split, int64_t stride_ml, int64_t stride_acc, __nv_bfloat16* __restrict__ out, int64_t ldo,
                        float* __restrict__ lse, int lse_pad, int num_heads, int batch, int q_len, int q_batch_stride,
                        int64_t total_q_rows) {
  const int lane = threadIdx.x & 31;
  const int64_t w = static_cast<int64_t>(blockIdx.x) * 8 + (threadIdx.x >> 5);
  if (w >= static_cast<int64_t>(batch) * q_len * num_heads) return;
  const int head = static_cast<int>(w % num_heads);
  const int64_t r = w / num_heads;
  const int q_row = static_cast<int>(r % q_len), b = static_cast<int>(r / q_len);
  const int64_t g_row = static_cast<int64_t>(b) * q_batch_stride + q_row;
  const int64_t ml = static_cast<int64_t>(head) * total_q_rows + g_row;
  float M = -INFINITY;
  for (int s = 0; s < n_split; ++s) M = fmaxf(M, ws_m[s * stride_ml + ml]);
  float L = 0.f, o0 = 0.f, o1 = 0.f;
  for (int s = 0; s < n_split; ++s) {
    const float wgt = exp2f(ws_m[s * stride_ml + ml] - M);
    L = fmaf(ws_l[s * stride_ml + ml], wgt, L);
    const float2 a = *reinterpret_cast<const float2*>(ws_acc + s * stride_acc + g_row * (static_cast<int64_t>(num_heads) * 64) +
                                                      head * 64 + lane * 2);
    o0 = fmaf(a.x, wgt, o0);
    o1 = fmaf(a.y, wgt, o1);
  }
  const float inv = 1.0f / L;
  *reinterpret_cast<uint32_t*>(out + g_row * ldo + head * 64 + lane * 2) = pack_bf16x2(o0 * inv, o1 * inv);
  if (lse != nullptr && lane == 0) lse[(static_cast<int64_t>(head) * batch + b) * lse_pad + q_row] = M + log2f(L);
}

// Number of K/V chunks for a launch: only when the un-split grid leaves more than half of the SMs idle and every chunk
// keeps >= 8 tiles (the per-CTA prologue / epilogue is ~2 tile periods).
static int choose_split(const sr_attn_t* a, int base_ctas, int n_all_tiles, int num_sms) {
  if (a->carry_in || a->carry_out || base_ctas * 2 > num_sms) return 1;
  int ks = num_sms / base_ctas;
  ks = std::min(ks, n_all_tiles / 8);
  if (ks <= 1) return 1;
  const int chunk = (n_all_tiles + ks - 1) / ks;           // tiles per chunk; re-derive the count so that no chunk is empty
  return (n_all_tiles + chunk - 1) / chunk;
}

template <int kPoly>
static int launch4(dim3 grid, cudaStream_t st, const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv,
                   const CUtensorMap& tpk, const CUtensorMap& tpv, const Attn4Params& p) {
  auto kern = attn_tc4_kernel<kPoly>;
  static bool configured = false;  // per instantiation
  if (!configured) {
    SR_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, k4Smem));
    configured = true;
  }
  kern<<<grid, k4Threads, k4Smem, st>>>(tq, tk, tv, tpk, tpv, p);
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

int attn_tc4_trace_read(long long* host_dst, int64_t count) {
  SR_CHECK_CUDA(cudaMemcpyFromSymbol(host_dst, g_attn_trace, static_cast<size_t>(count) * sizeof(long long)));
  return SR_OK;
}

// Picks (n4, n3) for a launch of `rows` (head, segment) pairs with `t` query tiles each on `sms` SMs: simulates the in-order
// dispatch of [all four-tile CTAs, then all three-tile CTAs] onto the earliest-free SM with CTA time ~ its active query
// tiles (the kernel is bound by the per-tile softmax work, 1 CTA/SM) and keeps the smallest makespan.  Cached per shape.
static void choose_mix(int t, int rows, int sms, int force_uniform, int* n4_out, int* n3_out) {
  struct Key { int t, rows, sms, n4, n3; };
  static std::mutex mu;
  static std::vector<Key> cache;
  const int a_max = (t + 3) / 4;
  if (force_uniform) { *n4_out = a_max; *n3_out = 0; return; }
  {
    std::lock_guard<std::mutex> lock(mu);
    for (const Key& k : cache)
      if (k.t == t && k.rows == rows && k.sms == sms) { *n4_out = k.n4; *n3_out = k.n3; return; }
  }
  long best = -1;
  int best_a = a_max, best_b = 0;
  std::vector<long> load(sms);
  for (int a = a_max; a >= 0 && a >= a_max - 24; --a) {
    const int rem = t - 4 * a;
    const int b = rem > 0 ? (rem + 2) / 3 : 0;
    if (a == 0 && b == 0) continue;
    if (static_cast<long>(rows) * b > 16384) break;
    // heavy phase in closed form, light phase with a min-heap
    const long n_heavy = static_cast<long>(rows) * a;
    const int last_heavy = (a == a_max && rem < 0) ? 4 + rem : 4;          // tiles of a row's last heavy CTA when b == 0
    for (int i = 0; i < sms; ++i) load[i] = 4 * (n_heavy / sms + (i < n_heavy % sms ? 1 : 0));
    (void)last_heavy;
    std::priority_queue<long, std::vector<long>, std::greater<long>> heap(load.begin(), load.end());
    for (int r = 0; r < rows; ++r)
      for (int x = 0; x < b; ++x) {
        const int cost = (x == b - 1) ? rem - 3 * (b - 1) : 3;
        const long l = heap.top();
        heap.pop();
        heap.push(l + cost);
      }
    long mk = 0;
    while (!heap.empty()) { mk = std::max(mk, heap.top()); heap.pop(); }
    if (best < 0 || mk < best) { best = mk; best_a = a; best_b = b; }
  }
  std::lock_guard<std::mutex> lock(mu);
  cache.push_back({t, rows, sms, best_a, best_b});
  *n4_out = best_a;
  *n3_out = best_b;
}

// Called by sr_attn_fwd_bf16 (attn_tc.cu) after argument validation.
int attn_tc4_launch(const sr_attn_t* a, int poly, cudaStream_t st) {
  const int64_t hd = static_cast<int64_t>(a->num_heads) * 64;
  const bool has_prefix = a->pk != nullptr && a->prefix_len > 0;
  const uint64_t qbs = a->batch > 1 ? a->q_batch_stride : a->q_len;
  const uint64_t kbs = a->batch > 1 ? a->kv_batch_stride : a->kv_len;
  CUtensorMap tq, tk, tv, tpk, tpv;
  {
    uint32_t box[3] = {64, 128, 1};
    uint64_t dims[3] = {static_cast<uint64_t>(hd), static_cast<uint64_t>(a->q_len), static_cast<uint64_t>(a->batch)};
    uint64_t str[2] = {static_cast<uint64_t>(a->ldq) * 2, qbs * static_cast<uint64_t>(a->ldq) * 2};
    if (int rc = make_tmap_bf16(&tq, a->q, 3, dims, str, box)) return rc;
  }
  {
    uint32_t box[3] = {64, 64, 1};
    uint64_t dims[3] = {static_cast<uint64_t>(hd), static_cast<uint64_t>(a->kv_len), static_cast<uint64_t>(a->batch)};
    uint64_t str[2] = {static_cast<uint64_t>(a->ldkv) * 2, kbs * static_cast<uint64_t>(a->ldkv) * 2};
    if (int rc = make_tmap_bf16(&tk, a->k, 3, dims, str, box)) return rc;
    if (int rc = make_tmap_bf16(&tv, a->v, 3, dims, str, box)) return rc;
  }
  if (has_prefix) {
    uint64_t dims[2] = {static_cast<uint64_t>(hd), static_cast<uint64_t>(a->prefix_len)};
    uint64_t str[1] = {static_cast<uint64_t>(a->ldp) * 2};
    uint32_t box[2] = {64, 64};
    if (int rc = make_tmap_bf16(&tpk, a->pk, 2, dims, str, box)) return rc;
    if (int rc = make_tmap_bf16(&tpv, a->pv, 2, dims, str, box)) return rc;
  } else {
    tpk = tk;
    tpv = tv;
  }
  Attn4Params p;
  p.out = reinterpret_cast<__nv_bfloat16*>(a->out);
  p.ldo = a->ldo;
  p.carry_m = a->carry_m;
  p.carry_l = a->carry_l;
  p.carry_acc = a->carry_acc;
  p.carry_in = a->carry_in;
  p.carry_out = a->carry_out;
  p.num_heads = a->num_heads;
  p.q_len = static_cast<int>(a->q_len);
  p.q_batch_stride = static_cast<int>(qbs);
  p.kv_len = static_cast<int>(a->kv_len);
  p.prefix_len = has_prefix ? static_cast<int>(a->prefix_len) : 0;
  p.n_prefix_tiles = (p.prefix_len + 63) / 64;
  p.n_own_tiles = (p.kv_len + 63) / 64;
  p.total_q_rows = a->batch > 1 ? static_cast<int64_t>(a->batch - 1) * a->q_batch_stride + a->q_len : a->q_len;
  p.scale_log2 = a->scale * 1.4426950408889634f;
  p.trace = dbg_env("SR_ATTN_TRACE") != nullptr;
  p.lse = a->lse;
  p.lse_pad = static_cast<int>(a->lse_pad);
  static int num_sms = 0;
  if (num_sms == 0) {
    int dev = 0;
    SR_CHECK_CUDA(cudaGetDevice(&dev));
    SR_CHECK_CUDA(cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev));
  }
  p.batch = a->batch;
  p.n_rows = a->num_heads * a->batch;
  const int t_tiles = static_cast<int>((a->q_len + 127) / 128);
  choose_mix(t_tiles, p.n_rows, num_sms, a->variant == 4, &p.n4, &p.n3);
  p.row_major = a->variant == 6 ? 1 : (a->variant == 5 ? 0 : (p.n_rows >= num_sms ? 1 : 0));
  const int64_t base_ctas = static_cast<int64_t>(p.n_rows) * (p.n4 + p.n3);
  SR_REQUIRE(base_ctas < (1ll << 31), "sr_attn_fwd_bf16: too many CTAs");
  p.n_split = 1;
  p.split_stride_ml = p.split_stride_acc = 0;
  if (a->workspace != nullptr) {
    const int ks = choose_split(a, static_cast<int>(base_ctas), p.n_prefix_tiles + p.n_own_tiles, num_sms);
    const int64_t per_split = p.total_q_rows * a->num_heads * (2 + 64);            // floats: m, l, acc
    if (ks > 1 && a->workspace_bytes >= ks * per_split * 4) {
      p.n_split = ks;
      p.split_stride_ml = p.total_q_rows * a->num_heads;
      p.split_stride_acc = p.total_q_rows * a->num_heads * 64;
      float* ws = reinterpret_cast<float*>(a->workspace);
      p.carry_m = ws;
      p.carry_l = ws + ks * p.split_stride_ml;
      p.carry_acc = ws + 2 * ks * p.split_stride_ml;
      p.carry_in = 0;
      p.carry_out = 1;
    }
  }
  dim3 grid(static_cast<unsigned>(base_ctas * p.n_split), 1, 1);
  if (p.n_split > 1) {
    SR_REQUIRE(poly == 2, "sr_attn_fwd_bf16: split-KV runs the product kernel only");
    if (int rc = launch4<2>(grid, st, tq, tk, tv, tpk, tpv, p)) return rc;
    const int64_t warps = static_cast<int64_t>(a->batch) * a->q_len * a->num_heads;
    attn_split_merge_kernel<<<static_cast<unsigned>((warps + 7) / 8), 256, 0, st>>>(
        p.carry_m, p.carry_l, p.carry_acc, p.n_split, p.split_stride_ml, p.split_stride_acc,
        reinterpret_cast<__nv_bfloat16*>(a->out), a->ldo, a->lse, static_cast<int>(a->lse_pad), a->num_heads, a->batch,
        static_cast<int>(a->q_len), static_cast<int>(qbs), p.total_q_rows);
    ++g_launch_count;
    SR_CHECK_CUDA(cudaGetLastError());
    return SR_OK;
  }
  switch (poly) {
    case 2: return launch4<2>(grid, st, tq, tk, tv, tpk, tpv, p);
#ifdef SR_DEBUG
    case 0: return launch4<0>(grid, st, tq, tk, tv, tpk, tpv, p);
    case 1: return launch4<1>(grid, st, tq, tk, tv, tpk, tpv, p);
    case 3: return launch4<3>(grid, st, tq, tk, tv, tpk, tpv, p);
    case 4: return launch4<4>(grid, st, tq, tk, tv, tpk, tpv, p);
    case 8: return launch4<8>(grid, st, tq, tk, tv, tpk, tpv, p);
#endif
  }
  return fail(SR_ERR_SHAPE, "sr_attn_fwd_bf16 (4-tile kernel): SR_ATTN_POLY must be 0..4 or 8 (got %d)", poly);
}

}  // namespace sr

extern "C" int64_t sr_attn_workspace_bytes(const sr_attn_t* a, int32_t num_sms) {
  using namespace sr;
  if (a == nullptr || a->head_dim != 64 || a->variant == 3 || num_sms <= 0 || a->q_len <= 0 || a->kv_len <= 0) return 0;
  int n4 = 0, n3 = 0;
  const int rows = a->num_heads * a->batch;
  choose_mix(static_cast<int>((a->q_len + 127) / 128), rows, num_sms, a->variant == 4, &n4, &n3);
  const int64_t pre = a->pk != nullptr && a->prefix_len > 0 ? (a->prefix_len + 63) / 64 : 0;
  const int ks = choose_split(a, rows * (n4 + n3), static_cast<int>(pre + (a->kv_len + 63) / 64), num_sms);
  if (ks <= 1) return 0;
  const int64_t total_q_rows = a->batch > 1 ? static_cast<int64_t>(a->batch - 1) * a->q_batch_stride + a->q_len : a->q_len;
  return ks * total_q_rows * a->num_heads * (2 + 64) * 4;
}

extern "C" int sr_attn_plan(int64_t q_len, int32_t num_heads, int32_t batch, int32_t num_sms, int32_t* n4, int32_t* n3) {
  using namespace sr;
  SR_REQUIRE(q_len > 0 && num_heads > 0 && batch > 0 && num_sms > 0 && n4 && n3, "sr_attn_plan: bad argument");
  int a = 0, b = 0;
  choose_mix(static_cast<int>((q_len + 127) / 128), num_heads * batch, num_sms, 0, &a, &b);
  *n4 = a;
  *n3 = b;
  return SR_OK;
}
