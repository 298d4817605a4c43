// Flash attention forward on tcgen05 tensor cores, head_dim 64, bf16 in / fp32 softmax / bf16 out (sm_100a).
//
// Replaces F.scaled_dot_product_attention at sailrecon/layers/attention.py:103-109 for the three aggregator
// block stacks (frame / global / global-reloc, sailrecon/models/aggregator.py:643-769).  The reference's dense
// boolean mask (aggregator.py:302-311, build_allow_block :832-847) is never materialised: its structure is
// "shared scene prefix + own frame", which this kernel walks as two K/V tile ranges.
//
// One CTA = one head x 2 query tiles of 128 rows.  384 threads:
//   warp 0        TMA producer: Q tiles once, then K/V tiles (128 x 64, SWIZZLE_128B) through a kAtStages ring
//   warp 1, 3     MMA issuers, one per query tile q (one thread each), each an independent in-order pipeline:
//                     S_q(j+1) = Q_q K(j+1)^T   tcgen05.mma SS, M=128 N=128 K=16 x4 -> TMEM   as soon as the softmax
//                                               warpgroup has pulled S_q(j) into registers (s_free)
//                     O_q     += P_q(j) V(j)    tcgen05.mma TS (A = P from TMEM), M=128 N=64 K=16 x8 -> TMEM
//   warp 2        TMEM allocator: S0 S1 (128 cols each) | O0 O1 (64 each) | P0 P1 (64 each: 128 bf16) = 512 columns
//   warps 4-7     softmax warpgroup of Q tile 0, warps 8-11 of Q tile 1: one thread = one query row
//                 (tcgen05.ld 32x32b), online softmax in the log2 domain with lazy rescale of O
//                 (rescale only when the row max grows by more than 2^8), P written back to TMEM as packed bf16
//                 (tcgen05.st) so the second MMA reads it from TMEM and shared memory only carries Q, K, V.
// The softmax inner loop is written for instruction count (it is the critical resource at head_dim 64, where one
// 128x128 tile costs 16K exponentials against 512 tensor-core cycles): FMNMX3 row max, FFMA2 / FADD2 packed fp32,
// MUFU.EX2 without range wrappers, F2FP packed conversion.
// Optional carry of (m, l, un-normalised O) in/out turns one call into a step of ring attention
// (global blocks sharded over ranks).
#include <cstdlib>

#include "attn_math.cuh"
#include "common.cuh"
#include "sr_internal.h"

namespace sr {
extern int g_launch_count;
int attn_tc4_launch(const sr_attn_t* a, int poly, cudaStream_t st);   // attn_tc4.cu
int attn_tc4_trace_read(long long* host_dst, int64_t count);
int attn_tc3_launch(const sr_attn_t* a, int poly, cudaStream_t st);   // attn_tc3.cu
constexpr int kAtKernelDefault = 4;   // query tiles per CTA: 2 (this file) or 4 (attn_tc4.cu); SR_ATTN_KERNEL overrides

constexpr int kAtNQ = 2;
constexpr int kAtStages = 4;
constexpr int kAtTile = 128 * 64 * 2;                 // 16 KB: one 128 x 64 bf16 tile
constexpr int kAtThreads = 128 + kAtNQ * 128;
constexpr int kAtSmem = kAtNQ * kAtTile               // Q
                        + kAtStages * 2 * kAtTile     // K, V ring
                        + 1024;
constexpr int kAtTmemCols = 512;                      // S0 S1 (128 each) | O0 O1 (64 each) | P0 P1 (64 each)
constexpr float kLazyRescaleLog2 = 8.0f;

struct AttnParams {
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
  int n_prefix_tiles, n_own_tiles;
  int64_t total_q_rows;
  float scale_log2;
  int trace;                     // SR_ATTN_TRACE
  int dbg;                       // SR_ATTN_DBG (timing experiments only): 1 = skip PV MMAs, 2 = skip QK MMAs
  int lag_ns;                    // start-up lag of the second softmax warpgroup (de-phases the two MUFU streams)
};

#ifdef SR_DEBUG   // the original two-tile kernel: kept for A/B timing experiments, not part of the shipped library
// Profiling aid: SR_ATTN_TRACE=1 makes CTA (0,0,0) record clock64() at the pipeline hand-offs of its first 64 K/V tiles
// (read back with sr_debug_trace_read, printed by scripts/attn_trace.py).  Layout: g_attn_trace[role][tile][event],
// role 0 = softmax warp 4, 1 = MMA issuer of query tile 0, 2 = TMA producer.
static __device__ long long g_attn_trace[3 * 64 * 8];
#define SR_TRACE(role, j, ev)                                                           \
  do {                                                                                  \
    if (trace_on && (j) < 64) g_attn_trace[((role) * 64 + (j)) * 8 + (ev)] = clock64(); \
  } while (0)

// kPoly = how many of every 8 element pairs take the polynomial exp2 (0 = all on MUFU.EX2)
template <int kPoly>
__global__ void __launch_bounds__(kAtThreads, 1)
attn_tc_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
               const __grid_constant__ CUtensorMap tmap_v, const __grid_constant__ CUtensorMap tmap_pk,
               const __grid_constant__ CUtensorMap tmap_pv, const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t q_full;
  __shared__ __align__(8) uint64_t k_full[kAtStages], k_empty[kAtStages], v_full[kAtStages], v_empty[kAtStages];
  __shared__ __align__(8) uint64_t s_full[kAtNQ], s_free[kAtNQ], p_full[kAtNQ], pv_done[kAtNQ];
  __shared__ uint32_t tmem_base_slot;

  // warp index through a shuffle: the compiler then knows it is warp-uniform, keeps the producer / MMA-issuer address
  // arithmetic in uniform registers and emits UTMALDG / UTCHMMA directly (a `lane == 0` branch made every issue a
  // 12-instruction R2UR waterfall loop: ~100 clk per MMA, which was the bottleneck of the whole kernel)
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + kAtNQ * kAtTile;
  uint8_t* sV = sK + kAtStages * kAtTile;

  const int head = blockIdx.y;
  const int b = blockIdx.z;
  const int q0 = blockIdx.x * (128 * kAtNQ);                 // first query row (within the segment) of this CTA
  const int nq_valid = (q0 + 128 < p.q_len) ? 2 : 1;         // second tile entirely past q_len -> skipped
  const int n_tiles = p.n_prefix_tiles + p.n_own_tiles;
  const bool trace_on = p.trace && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && lane == 0;

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
    for (int s = 0; s < kAtStages; ++s) {
      mbar_init(&k_full[s], 1);
      mbar_init(&k_empty[s], nq_valid);      // one tcgen05.commit per MMA issuer
      mbar_init(&v_full[s], 1);
      mbar_init(&v_empty[s], nq_valid);
    }
    for (int q = 0; q < kAtNQ; ++q) {
      mbar_init(&s_full[q], 1);
      mbar_init(&s_free[q], 4);              // one elected arrive per softmax warp
      mbar_init(&p_full[q], 4);
      mbar_init(&pv_done[q], 1);
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<kAtTmemCols>(&tmem_base_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_slot;
  const uint32_t tmem_S0 = tmem_base;                         // + q * 128
  const uint32_t tmem_O0 = tmem_base + kAtNQ * 128;           // + q * 64
  const uint32_t tmem_P0 = tmem_O0 + kAtNQ * 64;              // + q * 64   (128 bf16 per row = 64 columns)

  // register re-distribution: the 4 control warps keep 72 registers, the 8 softmax warps (a 128-value fp32 row
  // fragment per thread) grow to 216:  128 x 72 + 256 x 216 = 64512 <= 65536
  if (warp < 4) {
  asm volatile("setmaxnreg.dec.sync.aligned.u32 72;");
  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer (whole warp; one elected lane issues)
    if (elect_one()) {
      mbar_arrive_expect_tx(&q_full, nq_valid * kAtTile);
      for (int q = 0; q < nq_valid; ++q) tma_load_3d(sQ + q * kAtTile, &tmap_q, &q_full, head * 64, q0 + q * 128, b);
    }
    __syncwarp();
    for (int j = 0; j < n_tiles; ++j) {
      const int s = j % kAtStages;
      const uint32_t ph = ((j / kAtStages) & 1) ^ 1;
      const bool pre = j < p.n_prefix_tiles;
      const int r0 = (pre ? j : j - p.n_prefix_tiles) * 128;
      mbar_wait(&k_empty[s], ph);
      SR_TRACE(2, j, 0);
      if (elect_one()) {
        mbar_arrive_expect_tx(&k_full[s], kAtTile);
        if (pre) tma_load_2d(sK + s * kAtTile, &tmap_pk, &k_full[s], head * 64, r0);
        else     tma_load_3d(sK + s * kAtTile, &tmap_k, &k_full[s], head * 64, r0, b);
      }
      __syncwarp();
      mbar_wait(&v_empty[s], ph);
      SR_TRACE(2, j, 1);
      if (elect_one()) {
        mbar_arrive_expect_tx(&v_full[s], kAtTile);
        if (pre) tma_load_2d(sV + s * kAtTile, &tmap_pv, &v_full[s], head * 64, r0);
        else     tma_load_3d(sV + s * kAtTile, &tmap_v, &v_full[s], head * 64, r0, b);
      }
      __syncwarp();
    }
  } else if (warp == 1 || warp == 3) {
    // ------------------------------------------------------------------ MMA issuer of query tile q
    const int q = warp >> 1;
    if (q < nq_valid) {   // whole warp runs the loop (uniform control flow); one elected lane issues
      constexpr uint32_t idesc_s = umma_idesc_bf16(128, 128, 0, 0);   // Q (K-major) x K (K-major)
      constexpr uint32_t idesc_o = umma_idesc_bf16(128, 64, 0, 1);    // P (TMEM)    x V (MN-major: d contiguous)
      const uint32_t tS = tmem_S0 + q * 128, tO = tmem_O0 + q * 64, tP = tmem_P0 + q * 64;
      const uint64_t dq = umma_desc_sw128(smem_u32(sQ + q * kAtTile), 16, 1024);
      auto issue_s = [&](int stage) {
        const uint64_t db = umma_desc_sw128(smem_u32(sK + stage * kAtTile), 16, 1024);
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < 4; ++k)
            if (!(p.dbg & 2)) umma_ss(tS, dq + 2 * k, db + 2 * k, idesc_s, k != 0 ? 1u : 0u);
          umma_commit(&s_full[q]);
          umma_commit(&k_empty[stage]);
        }
        __syncwarp();
      };
      auto issue_pv = [&](int stage, bool accumulate) {
        const uint32_t vbase = smem_u32(sV + stage * kAtTile);
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            // A: P columns [8k, 8k+8) hold kv [16k, 16k+16) as packed bf16 pairs
            // B: V rows [16k, 16k+16) = two 8-row swizzle atoms (SBO 1024 B); one 64-wide MN block (LBO unused)
            const uint64_t db = umma_desc_sw128(vbase + k * 2048, 0, 1024);
            if (!(p.dbg & 1)) umma_ts(tO, tP + k * 8, db, idesc_o, (accumulate || k != 0) ? 1u : 0u);
          }
          umma_commit(&pv_done[q]);
          umma_commit(&v_empty[stage]);
        }
        __syncwarp();
      };

      mbar_wait(&q_full, 0);
      mbar_wait(&k_full[0], 0);
      tc_fence_after();
      issue_s(0);
      for (int j = 0; j < n_tiles; ++j) {
        const int s = j % kAtStages;
        if (j + 1 < n_tiles) {
          const int sn = (j + 1) % kAtStages;
          mbar_wait(&k_full[sn], ((j + 1) / kAtStages) & 1);
          if (q == 0) SR_TRACE(1, j, 0);
          mbar_wait(&s_free[q], j & 1);          // softmax q holds S_q(j) in registers
          if (q == 0) SR_TRACE(1, j, 1);
          tc_fence_after();
          issue_s(sn);
          if (q == 0) SR_TRACE(1, j, 2);
        }
        mbar_wait(&v_full[s], (j / kAtStages) & 1);
        if (q == 0) SR_TRACE(1, j, 3);
        mbar_wait(&p_full[q], j & 1);            // P_q(j) in TMEM, O_q rescaled
        if (q == 0) SR_TRACE(1, j, 4);
        tc_fence_after();
        issue_pv(s, p.carry_in != 0 || j > 0);
        if (q == 0) SR_TRACE(1, j, 5);
      }
    }
  }
  } else {
    // ------------------------------------------------------------------ softmax warpgroups
    asm volatile("setmaxnreg.inc.sync.aligned.u32 216;");
    const int q = (warp - 4) >> 2;
    if (q < nq_valid) {
      const int quarter = warp & 3;
      const int r_tile = quarter * 32 + lane;                 // row inside the 128-row tile == TMEM lane
      const int q_row = q0 + q * 128 + r_tile;                // row inside the segment
      const bool row_ok = q_row < p.q_len;
      const int64_t g_row = static_cast<int64_t>(b) * p.q_batch_stride + q_row;   // row in the q / out matrices
      const uint32_t lane_off = static_cast<uint32_t>(quarter * 32) << 16;
      const uint32_t tS = tmem_S0 + q * 128 + lane_off;
      const uint32_t tO = tmem_O0 + q * 64 + lane_off;
      const uint32_t tP = tmem_P0 + q * 64 + lane_off;

      float m = -INFINITY, l = 0.f;
      if (p.carry_in) {
        if (row_ok) {
          m = p.carry_m[static_cast<int64_t>(head) * p.total_q_rows + g_row];
          l = p.carry_l[static_cast<int64_t>(head) * p.total_q_rows + g_row];
        }
        // seed O with the carried un-normalised accumulator
        uint32_t o0[32], o1[32];
        const float* src = p.carry_acc + g_row * (static_cast<int64_t>(p.num_heads) * 64) + head * 64;
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          o0[i] = row_ok ? __float_as_uint(src[i]) : 0u;
          o1[i] = row_ok ? __float_as_uint(src[32 + i]) : 0u;
        }
        tmem_st32(tO, o0);
        tmem_st32(tO + 32, o1);
        tmem_st_wait();
      }
      const uint64_t scale2 = pack2(p.scale_log2, p.scale_log2);

      // valid kv columns of tile j (only the last tile of the prefix / own range is partial)
      auto valid_cols = [&](int j) {
        if (j < p.n_prefix_tiles) return j == p.n_prefix_tiles - 1 ? p.prefix_len - j * 128 : 128;
        return j == n_tiles - 1 ? p.kv_len - (j - p.n_prefix_tiles) * 128 : 128;
      };
      // row max of sraw (masked to `valid` columns), in raw score units
      uint32_t sraw[128];
      auto row_max = [&](int valid) {
        if (valid < 128) {
#pragma unroll
          for (int i = 0; i < 128; ++i)
            if (i >= valid) sraw[i] = 0xff800000u;   // -inf
        }
        float mx0 = max3(__uint_as_float(sraw[0]), __uint_as_float(sraw[1]), __uint_as_float(sraw[2]));
        float mx1 = max3(__uint_as_float(sraw[3]), __uint_as_float(sraw[4]), __uint_as_float(sraw[5]));
        float mx2 = max3(__uint_as_float(sraw[6]), __uint_as_float(sraw[7]), __uint_as_float(sraw[8]));
        float mx3 = max3(__uint_as_float(sraw[9]), __uint_as_float(sraw[10]), __uint_as_float(sraw[11]));
#pragma unroll
        for (int i = 12; i < 124; i += 8) {
          mx0 = max3(mx0, __uint_as_float(sraw[i + 0]), __uint_as_float(sraw[i + 1]));
          mx1 = max3(mx1, __uint_as_float(sraw[i + 2]), __uint_as_float(sraw[i + 3]));
          mx2 = max3(mx2, __uint_as_float(sraw[i + 4]), __uint_as_float(sraw[i + 5]));
          mx3 = max3(mx3, __uint_as_float(sraw[i + 6]), __uint_as_float(sraw[i + 7]));
        }
        mx0 = max3(mx0, __uint_as_float(sraw[124]), __uint_as_float(sraw[125]));
        mx1 = max3(mx1, __uint_as_float(sraw[126]), __uint_as_float(sraw[127]));
        return fmaxf(max3(mx0, mx1, mx2), mx3);
      };

      // Barrier probes: a (non-blocking) mbarrier.test_wait is issued one phase ahead of where its answer is needed,
      // so the ~60-150 clk of a barrier round trip overlaps the exponentials instead of sitting between two phases.
      bool s_ready = false;
      if (q == 1 && p.lag_ns > 0) __nanosleep(p.lag_ns);
      for (int j = 0; j < n_tiles; ++j) {
        if (!s_ready) mbar_wait(&s_full[q], j & 1);
        if (warp == 4) SR_TRACE(0, j, 0);
        tc_fence_after();
#pragma unroll
        for (int c = 0; c < 4; ++c) tmem_ld32(tS + c * 32, *reinterpret_cast<uint32_t(*)[32]>(&sraw[c * 32]));
        tmem_ld_wait();
        tc_fence_before();
        if (elect_one()) mbar_arrive(&s_free[q]);   // S_q may be overwritten by Q_q K(j+1)^T (wait::ld is warp-wide)
        const float mx = row_max(valid_cols(j)) * p.scale_log2;            // scale > 0
        if (warp == 4 && mx == mx) SR_TRACE(0, j, 2);                      // (data-dependent so it cannot be hoisted)
        const bool pv_ready = j > 0 && mbar_test_wait(&pv_done[q], (j - 1) & 1);
        float m_new = m;
        if (mx > m + kLazyRescaleLog2) m_new = mx;                         // also the first tile (m = -inf)
        const float alpha = ex2_approx(m - m_new);                         // 1 when unchanged, 0 on the first tile
        const uint64_t negm2 = pack2(-m_new, -m_new);
        uint64_t sum0 = 0, sum1 = 0, sum2 = 0, sum3 = 0;                   // packed (0.f, 0.f)
        uint32_t pk[64];
#pragma unroll
        for (int i = 0; i < 128; i += 8) {
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
        if (warp == 4 && l >= 0.f) SR_TRACE(0, j, 3);

        if (j > 0 && !pv_ready) mbar_wait(&pv_done[q], (j - 1) & 1);   // P_q and O_q no longer in use by the tensor core
        if (warp == 4) SR_TRACE(0, j, 4);
        tc_fence_after();
        {
          uint32_t(&h0)[32] = *reinterpret_cast<uint32_t(*)[32]>(&pk[0]);
          uint32_t(&h1)[32] = *reinterpret_cast<uint32_t(*)[32]>(&pk[32]);
          tmem_st32(tP, h0);
          tmem_st32(tP + 32, h1);
        }
        s_ready = j + 1 < n_tiles && mbar_test_wait(&s_full[q], (j + 1) & 1);
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
        if (warp == 4) SR_TRACE(0, j, 5);
      }

      // ---------------------------------------------------------------- epilogue
      mbar_wait(&pv_done[q], (n_tiles - 1) & 1);
      tc_fence_after();
      uint32_t o0[32], o1[32];
      tmem_ld32(tO, o0);
      tmem_ld32(tO + 32, o1);
      tmem_ld_wait();
      if (row_ok) {
        if (p.carry_out) {
          p.carry_m[static_cast<int64_t>(head) * p.total_q_rows + g_row] = m;
          p.carry_l[static_cast<int64_t>(head) * p.total_q_rows + g_row] = l;
          float* dst = p.carry_acc + g_row * (static_cast<int64_t>(p.num_heads) * 64) + head * 64;
#pragma unroll
          for (int i = 0; i < 32; i += 4) {
            *reinterpret_cast<uint4*>(dst + i) = make_uint4(o0[i], o0[i + 1], o0[i + 2], o0[i + 3]);
            *reinterpret_cast<uint4*>(dst + 32 + i) = make_uint4(o1[i], o1[i + 1], o1[i + 2], o1[i + 3]);
          }
        } else {
          const float inv = 1.0f / l;
          __nv_bfloat16* dst = p.out + g_row * p.ldo + head * 64;
#pragma unroll
          for (int i = 0; i < 32; i += 8) {
            uint4 w;
            w.x = pack_bf16x2(__uint_as_float(o0[i + 0]) * inv, __uint_as_float(o0[i + 1]) * inv);
            w.y = pack_bf16x2(__uint_as_float(o0[i + 2]) * inv, __uint_as_float(o0[i + 3]) * inv);
            w.z = pack_bf16x2(__uint_as_float(o0[i + 4]) * inv, __uint_as_float(o0[i + 5]) * inv);
            w.w = pack_bf16x2(__uint_as_float(o0[i + 6]) * inv, __uint_as_float(o0[i + 7]) * inv);
            *reinterpret_cast<uint4*>(dst + i) = w;
          }
#pragma unroll
          for (int i = 0; i < 32; i += 8) {
            uint4 w;
            w.x = pack_bf16x2(__uint_as_float(o1[i + 0]) * inv, __uint_as_float(o1[i + 1]) * inv);
            w.y = pack_bf16x2(__uint_as_float(o1[i + 2]) * inv, __uint_as_float(o1[i + 3]) * inv);
            w.z = pack_bf16x2(__uint_as_float(o1[i + 4]) * inv, __uint_as_float(o1[i + 5]) * inv);
            w.w = pack_bf16x2(__uint_as_float(o1[i + 6]) * inv, __uint_as_float(o1[i + 7]) * inv);
            *reinterpret_cast<uint4*>(dst + 32 + i) = w;
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<kAtTmemCols>(tmem_base);
  }
}

constexpr int kAtPolyDefault = 0;   // measured on B200: the kernel is latency- not MUFU-throughput-bound; poly does not pay yet

template <int kPoly>
static int launch_attn(dim3 grid, cudaStream_t st, const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv,
                       const CUtensorMap& tpk, const CUtensorMap& tpv, const AttnParams& p) {
  auto kern = attn_tc_kernel<kPoly>;
  static bool configured = false;  // per instantiation
  if (!configured) {
    SR_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kAtSmem));
    configured = true;
  }
  kern<<<grid, kAtThreads, kAtSmem, st>>>(tq, tk, tv, tpk, tpv, p);
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

#endif  // SR_DEBUG

}  // namespace sr

extern "C" int sr_debug_trace_read(long long* host_dst, int64_t count) {
  using namespace sr;
  SR_REQUIRE(host_dst && count > 0 && count <= 3 * 64 * 8, "sr_debug_trace_read: count out of range");
#ifdef SR_DEBUG
  {
    int variant = kAtKernelDefault;
    if (const char* env = dbg_env("SR_ATTN_KERNEL")) variant = atoi(env);
    if (variant == 4) return attn_tc4_trace_read(host_dst, count);
  }
  SR_CHECK_CUDA(cudaMemcpyFromSymbol(host_dst, g_attn_trace, static_cast<size_t>(count) * sizeof(long long)));
  return SR_OK;
#else
  return fail(SR_ERR_SHAPE, "sr_debug_trace_read: the pipeline trace exists only in -DSR_DEBUG builds of the library");
#endif
}

extern "C" int sr_attn_fwd_bf16(const sr_attn_t* a, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(a && a->q && a->k && a->v, "sr_attn_fwd_bf16: null pointer");
  SR_REQUIRE(a->head_dim == 64, "sr_attn_fwd_bf16: head_dim must be 64 (got %d)", a->head_dim);
  SR_REQUIRE(a->num_heads > 0 && a->batch > 0 && a->q_len > 0 && a->kv_len > 0,
             "sr_attn_fwd_bf16: empty problem (heads %d batch %d q_len %lld kv_len %lld)", a->num_heads, a->batch,
             (long long)a->q_len, (long long)a->kv_len);
  SR_REQUIRE(a->batch <= 65535 && a->num_heads <= 65535, "sr_attn_fwd_bf16: batch / heads exceed grid limits");
  const int64_t hd = static_cast<int64_t>(a->num_heads) * 64;
  SR_REQUIRE(a->ldq >= hd && a->ldkv >= hd && a->ldq % 8 == 0 && a->ldkv % 8 == 0,
             "sr_attn_fwd_bf16: q/kv row strides must be >= heads*64 and multiples of 8");
  SR_REQUIRE(((reinterpret_cast<uintptr_t>(a->q) | reinterpret_cast<uintptr_t>(a->k) | reinterpret_cast<uintptr_t>(a->v)) & 15) == 0,
             "sr_attn_fwd_bf16: q, k, v must be 16-byte aligned");
  SR_REQUIRE(a->batch == 1 || (a->q_batch_stride >= a->q_len && a->kv_batch_stride >= a->kv_len),
             "sr_attn_fwd_bf16: batch strides smaller than segment lengths");
  const bool has_prefix = a->pk != nullptr && a->prefix_len > 0;
  if (has_prefix) {
    SR_REQUIRE(a->pv && a->ldp >= hd && a->ldp % 8 == 0, "sr_attn_fwd_bf16: bad prefix K/V");
    SR_REQUIRE(((reinterpret_cast<uintptr_t>(a->pk) | reinterpret_cast<uintptr_t>(a->pv)) & 15) == 0,
               "sr_attn_fwd_bf16: prefix K/V must be 16-byte aligned");
  }
  const bool carry = a->carry_in || a->carry_out;
  if (carry) SR_REQUIRE(a->carry_m && a->carry_l && a->carry_acc, "sr_attn_fwd_bf16: carry buffers missing");
  if (!a->carry_out) {
    SR_REQUIRE(a->out && a->ldo >= hd && a->ldo % 8 == 0 && (reinterpret_cast<uintptr_t>(a->out) & 15) == 0,
               "sr_attn_fwd_bf16: out must be 16-byte aligned with row stride >= heads*64, multiple of 8");
  }
  SR_REQUIRE(a->q_len < (1ll << 30) && a->kv_len < (1ll << 30) && a->prefix_len < (1ll << 30),
             "sr_attn_fwd_bf16: sequence too long");

  {
    // Kernel choice.  The three- and four-tile kernels run the same ~590 clk per (128-row query tile x 64-row K/V tile)
    // and one CTA per SM, so a launch's time is the makespan of its CTAs weighted by their active query tiles.
    // variant 0 (default): the four-tile kernel with its mixed four- / three-tile work decomposition (attn_tc4.cu,
    // choose_mix); 4: the same kernel with uniform four-tile CTAs; 3: the three-tile kernel (P outside S, one issuer per tile).
    int variant = a->variant != 0 ? a->variant : kAtKernelDefault;
    if (a->variant == 0)
      if (const char* env = dbg_env("SR_ATTN_KERNEL")) variant = atoi(env);
    // polynomial exp2 on the FMA pipe for 2 of every 8 pairs: measured best (0: 851, 1: 905, 2: 921, 3: 892, 4: 843 TF/s)
    int poly = 2;
    if (const char* env = dbg_env("SR_ATTN_POLY")) poly = atoi(env);
    if (a->lse != nullptr) {
      SR_REQUIRE(a->lse_pad >= a->q_len && a->lse_pad % 64 == 0, "sr_attn_fwd_bf16: lse_pad must be a multiple of 64 >= q_len");
      SR_REQUIRE(variant >= 3 && variant <= 6, "sr_attn_fwd_bf16: the two-tile kernel (SR_ATTN_KERNEL=2) does not write lse");
    }
    if (variant == 4 || variant == 5 || variant == 6) return attn_tc4_launch(a, poly, reinterpret_cast<cudaStream_t>(stream));
    if (variant == 3) return attn_tc3_launch(a, poly, reinterpret_cast<cudaStream_t>(stream));
#ifndef SR_DEBUG
    return fail(SR_ERR_SHAPE, "sr_attn_fwd_bf16: kernel variant must be 0 (auto) or 3 .. 6 (got %d)", variant);
#endif
  }
#ifdef SR_DEBUG
  const uint64_t qbs = a->batch > 1 ? a->q_batch_stride : a->q_len;
  const uint64_t kbs = a->batch > 1 ? a->kv_batch_stride : a->kv_len;
  CUtensorMap tq, tk, tv, tpk, tpv;
  uint32_t box3[3] = {64, 128, 1};
  {
    uint64_t dims[3] = {static_cast<uint64_t>(hd), static_cast<uint64_t>(a->q_len), static_cast<uint64_t>(a->batch)};
    uint64_t str[2] = {static_cast<uint64_t>(a->ldq) * 2, qbs * static_cast<uint64_t>(a->ldq) * 2};
    if (int rc = make_tmap_bf16(&tq, a->q, 3, dims, str, box3)) return rc;
  }
  {
    uint64_t dims[3] = {static_cast<uint64_t>(hd), static_cast<uint64_t>(a->kv_len), static_cast<uint64_t>(a->batch)};
    uint64_t str[2] = {static_cast<uint64_t>(a->ldkv) * 2, kbs * static_cast<uint64_t>(a->ldkv) * 2};
    if (int rc = make_tmap_bf16(&tk, a->k, 3, dims, str, box3)) return rc;
    if (int rc = make_tmap_bf16(&tv, a->v, 3, dims, str, box3)) return rc;
  }
  if (has_prefix) {
    uint64_t dims[2] = {static_cast<uint64_t>(hd), static_cast<uint64_t>(a->prefix_len)};
    uint64_t str[1] = {static_cast<uint64_t>(a->ldp) * 2};
    uint32_t box2[2] = {64, 128};
    if (int rc = make_tmap_bf16(&tpk, a->pk, 2, dims, str, box2)) return rc;
    if (int rc = make_tmap_bf16(&tpv, a->pv, 2, dims, str, box2)) return rc;
  } else {
    tpk = tk;
    tpv = tv;
  }

  AttnParams p;
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
  p.n_prefix_tiles = (p.prefix_len + 127) / 128;
  p.n_own_tiles = (p.kv_len + 127) / 128;
  p.total_q_rows = a->batch > 1 ? static_cast<int64_t>(a->batch - 1) * a->q_batch_stride + a->q_len : a->q_len;
  p.scale_log2 = a->scale * 1.4426950408889634f;
  p.lag_ns = 0;
  p.dbg = 0;
  if (const char* env = dbg_env("SR_ATTN_DBG")) p.dbg = atoi(env);
  p.trace = dbg_env("SR_ATTN_TRACE") != nullptr;
  if (const char* env = dbg_env("SR_ATTN_LAG_NS")) p.lag_ns = atoi(env);

  dim3 grid(static_cast<unsigned>((a->q_len + 128 * kAtNQ - 1) / (128 * kAtNQ)), static_cast<unsigned>(a->num_heads),
            static_cast<unsigned>(a->batch));
  // tuning knob (profiling only): SR_ATTN_POLY = 0 / 2 / 3 / 4 polynomial-exp2 pairs out of 8; default kAtPolyDefault
  int poly = kAtPolyDefault;
  if (const char* env = dbg_env("SR_ATTN_POLY")) poly = atoi(env);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  switch (poly) {
    case 0: return launch_attn<0>(grid, st, tq, tk, tv, tpk, tpv, p);
    case 2: return launch_attn<2>(grid, st, tq, tk, tv, tpk, tpv, p);
    case 3: return launch_attn<3>(grid, st, tq, tk, tv, tpk, tpv, p);
    case 4: return launch_attn<4>(grid, st, tq, tk, tv, tpk, tpv, p);
    case 8: return launch_attn<8>(grid, st, tq, tk, tv, tpk, tpv, p);
  }
  return fail(SR_ERR_SHAPE, "sr_attn_fwd_bf16: SR_ATTN_POLY must be 0, 2, 3 or 4 (got %d)", poly);
#endif  // SR_DEBUG
}
