// Flash attention forward on tcgen05, head_dim 64 — THREE query tiles per CTA over 64-row K/V tiles, P outside S (sm_100a).
//
// Same contract as attn_tc.cu / attn_tc4.cu (replaces F.scaled_dot_product_attention at
// sailrecon/layers/attention.py:103-109 with the structured [scene prefix | own frame] K/V walk of
// sailrecon/models/aggregator.py:302-311,832-847).  Experiment against the four-tile kernel: there P_q is written over
// S_q, so S_q(j+1) can only be issued behind PV_q(j) and the ~670 clk MMA turn-around sits inside every stream's serial
// chain.  Three tiles leave room in TMEM for a separate P:
//   TMEM (480 of 512 columns)  S0..S2 (64 fp32 columns each) | P0..P2 (64 bf16 = 32 columns each) | O0..O2 (64 each)
// so S_q(j+1) = Q_q K(j+1)^T is issued as soon as the softmax warps hold S_q(j) in registers and completes under their
// exponentials; the chain of a stream is the softmax alone.
//   smem      Q 3 x 16 KB | K ring 6 x 8 KB | V ring 6 x 8 KB                                  (145 KB)
//   warp 0    TMA producer
//   warps 1-3 MMA issuer of query tile q = warp - 1 (warp 2 allocates TMEM first):
//                 S_q(j+1)   SS, M=128 N=64 K=16 x4   after s_free[q](j)
//                 O_q += P_q(j) V(j)   TS, M=128 N=64 K=16 x4   after p_full[q](j)
//   warps 4-15  softmax: warp 4 + 4q + r owns rows [32r, 32r + 32) of query tile q, one thread per row
// 512 threads; setmaxnreg: control 56, softmax 152.
#include <cstdlib>

#include "attn_math.cuh"
#include "common.cuh"
#include "sr_internal.h"

namespace sr {
extern int g_launch_count;

constexpr int k3NQ = 3;
constexpr int k3Stages = 6;
constexpr int k3QTile = 128 * 64 * 2;                 // 16 KB: 128 x 64 bf16
constexpr int k3KvTile = 64 * 64 * 2;                 //  8 KB:  64 x 64 bf16
constexpr int k3Threads = 128 + k3NQ * 128;
constexpr int k3Smem = k3NQ * k3QTile + k3Stages * 2 * k3KvTile + 1024;
constexpr float k3LazyRescaleLog2 = 8.0f;

struct Attn3Params {
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
  float* lse;                        // training: [num_heads, batch, lse_pad] log2-domain log-sum-exp (nullptr = off)
  int lse_pad;
};

template <int kPoly>
__global__ void __launch_bounds__(k3Threads, 1)
attn_tc3_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
                const __grid_constant__ CUtensorMap tmap_v, const __grid_constant__ CUtensorMap tmap_pk,
                const __grid_constant__ CUtensorMap tmap_pv, const Attn3Params p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t q_full;
  __shared__ __align__(8) uint64_t k_full[k3Stages], k_empty[k3Stages], v_full[k3Stages], v_empty[k3Stages];
  __shared__ __align__(8) uint64_t s_full[k3NQ], s_free[k3NQ], p_full[k3NQ], pv_done[k3NQ];
  __shared__ uint32_t tmem_base_slot;

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);   // provably warp-uniform (see elect_one)
  const int lane = threadIdx.x & 31;
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + k3NQ * k3QTile;
  uint8_t* sV = sK + k3Stages * k3KvTile;

  const int head = blockIdx.y;
  const int b = blockIdx.z;
  const int q0 = blockIdx.x * (128 * k3NQ);                          // first query row (within the segment) of this CTA
  const int nq_valid = min(k3NQ, (p.q_len - q0 + 127) / 128);        // query tiles that hold at least one row
  const int n_tiles = p.n_prefix_tiles + p.n_own_tiles;

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
    for (int s = 0; s < k3Stages; ++s) {
      mbar_init(&k_full[s], 1);
      mbar_init(&k_empty[s], nq_valid);      // one tcgen05.commit per MMA issuer
      mbar_init(&v_full[s], 1);
      mbar_init(&v_empty[s], nq_valid);
    }
    for (int q = 0; q < k3NQ; ++q) {
      mbar_init(&s_full[q], 1);
      mbar_init(&s_free[q], 4);              // one elected arrive per softmax warp
      mbar_init(&p_full[q], 4);
      mbar_init(&pv_done[q], 1);
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<512>(&tmem_base_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_slot;
  const uint32_t tmem_S0 = tmem_base;                         // + q * 64
  const uint32_t tmem_P0 = tmem_base + k3NQ * 64;             // + q * 32   (64 bf16 per row = 32 columns)
  const uint32_t tmem_O0 = tmem_P0 + k3NQ * 32;               // + q * 64

  if (warp < 4) {
  asm volatile("setmaxnreg.dec.sync.aligned.u32 56;");
  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer (whole warp; one elected lane issues)
    if (elect_one()) {
      mbar_arrive_expect_tx(&q_full, nq_valid * k3QTile);
      for (int q = 0; q < nq_valid; ++q) tma_load_3d(sQ + q * k3QTile, &tmap_q, &q_full, head * 64, q0 + q * 128, b);
    }
    __syncwarp();
    for (int j = 0; j < n_tiles; ++j) {
      const int s = j % k3Stages;
      const uint32_t ph = ((j / k3Stages) & 1) ^ 1;
      const bool pre = j < p.n_prefix_tiles;
      const int r0 = (pre ? j : j - p.n_prefix_tiles) * 64;
      mbar_wait_relaxed(&k_empty[s], ph);
      if (elect_one()) {
        mbar_arrive_expect_tx(&k_full[s], k3KvTile);
        if (pre) tma_load_2d(sK + s * k3KvTile, &tmap_pk, &k_full[s], head * 64, r0);
        else     tma_load_3d(sK + s * k3KvTile, &tmap_k, &k_full[s], head * 64, r0, b);
      }
      __syncwarp();
      mbar_wait_relaxed(&v_empty[s], ph);
      if (elect_one()) {
        mbar_arrive_expect_tx(&v_full[s], k3KvTile);
        if (pre) tma_load_2d(sV + s * k3KvTile, &tmap_pv, &v_full[s], head * 64, r0);
        else     tma_load_3d(sV + s * k3KvTile, &tmap_v, &v_full[s], head * 64, r0, b);
      }
      __syncwarp();
    }
  } else {
    // ------------------------------------------------------------------ MMA issuer of query tile q (whole warp runs the
    // loop; one elected lane issues)
    const int q = warp - 1;
    if (q < nq_valid) {
      constexpr uint32_t idesc_s = umma_idesc_bf16(128, 64, 0, 0);   // Q (K-major) x K (K-major)
      constexpr uint32_t idesc_o = umma_idesc_bf16(128, 64, 0, 1);   // P (TMEM)    x V (MN-major: d contiguous)
      const uint32_t tS = tmem_S0 + q * 64, tO = tmem_O0 + q * 64, tP = tmem_P0 + q * 32;
      const uint64_t dq = umma_desc_sw128(smem_u32(sQ + q * k3QTile), 16, 1024);
      auto issue_s = [&](int stage) {
        const uint64_t db = umma_desc_sw128(smem_u32(sK + stage * k3KvTile), 16, 1024);
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_ss(tS, dq + 2 * k, db + 2 * k, idesc_s, k != 0 ? 1u : 0u);
          umma_commit(&s_full[q]);
          umma_commit(&k_empty[stage]);
        }
        __syncwarp();
      };
      auto issue_pv = [&](int stage, bool accumulate) {
        const uint32_t vbase = smem_u32(sV + stage * k3KvTile);
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            // A: P columns [8k, 8k+8) hold kv [16k, 16k+16) as packed bf16 pairs
            // B: V rows [16k, 16k+16) = two 8-row swizzle atoms (SBO 1024 B); one 64-wide MN block (LBO unused)
            const uint64_t db = umma_desc_sw128(vbase + k * 2048, 0, 1024);
            umma_ts(tO, tP + k * 8, db, idesc_o, (accumulate || k != 0) ? 1u : 0u);
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
        const int s = j % k3Stages;
        if (j + 1 < n_tiles) {
          const int sn = (j + 1) % k3Stages;
          mbar_wait(&k_full[sn], ((j + 1) / k3Stages) & 1);
          mbar_wait(&s_free[q], j & 1);          // softmax q holds S_q(j) in registers
          tc_fence_after();
          issue_s(sn);
        }
        mbar_wait(&v_full[s], (j / k3Stages) & 1);
        mbar_wait(&p_full[q], j & 1);            // P_q(j) in TMEM, O_q rescaled
        tc_fence_after();
        issue_pv(s, p.carry_in != 0 || j > 0);
      }
    }
  }
  } else {
    // ------------------------------------------------------------------ softmax warps: one thread = one query row
    asm volatile("setmaxnreg.inc.sync.aligned.u32 152;");
    const int q = (warp - 4) >> 2;
    if (q < nq_valid) {
      const int quarter = warp & 3;
      const int r_tile = quarter * 32 + lane;                 // row inside the 128-row tile == TMEM lane
      const int q_row = q0 + q * 128 + r_tile;                // row inside the segment
      const bool row_ok = q_row < p.q_len;
      const int64_t g_row = static_cast<int64_t>(b) * p.q_batch_stride + q_row;   // row in the q / out matrices
      const uint32_t lane_off = static_cast<uint32_t>(quarter * 32) << 16;
      const uint32_t tS = tmem_S0 + q * 64 + lane_off;
      const uint32_t tP = tmem_P0 + q * 32 + lane_off;
      const uint32_t tO = tmem_O0 + q * 64 + lane_off;

      float m = -INFINITY, l = 0.f;
      if (p.carry_in) {
        if (row_ok) {
          m = p.carry_m[static_cast<int64_t>(head) * p.total_q_rows + g_row];
          l = p.carry_l[static_cast<int64_t>(head) * p.total_q_rows + g_row];
        }
        // seed O with the carried un-normalised accumulator
        const float* src = p.carry_acc + g_row * (static_cast<int64_t>(p.num_heads) * 64) + head * 64;
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
      auto valid_cols = [&](int j) {
        if (j < p.n_prefix_tiles) return j == p.n_prefix_tiles - 1 ? p.prefix_len - j * 64 : 64;
        return j == n_tiles - 1 ? p.kv_len - (j - p.n_prefix_tiles) * 64 : 64;
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

      bool s_ready = false;
      for (int j = 0; j < n_tiles; ++j) {
        if (!s_ready) mbar_wait(&s_full[q], j & 1);
        tc_fence_after();
        tmem_ld32(tS, *reinterpret_cast<uint32_t(*)[32]>(&sraw[0]));
        tmem_ld32(tS + 32, *reinterpret_cast<uint32_t(*)[32]>(&sraw[32]));
        tmem_ld_wait();
        tc_fence_before();
        if (elect_one()) mbar_arrive(&s_free[q]);   // S_q may be overwritten by Q_q K(j+1)^T (wait::ld is warp-wide)
        const float mx = row_max(valid_cols(j)) * p.scale_log2;            // scale > 0
        const bool pv_ready = j > 0 && mbar_test_wait(&pv_done[q], (j - 1) & 1);   // probed early, needed before the P store
        float m_new = m;
        if (mx > m + k3LazyRescaleLog2) m_new = mx;                        // also the first tile (m = -inf)
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

        if (j > 0 && !pv_ready) mbar_wait(&pv_done[q], (j - 1) & 1);   // P_q and O_q no longer in use by the tensor core
        tc_fence_after();
        tmem_st32(tP, pk);
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
      }

      // ---------------------------------------------------------------- epilogue
      mbar_wait(&pv_done[q], (n_tiles - 1) & 1);
      tc_fence_after();
      const float inv = 1.0f / l;
#pragma unroll
      for (int hlf = 0; hlf < 2; ++hlf) {
        uint32_t o[32];
        tmem_ld32(tO + hlf * 32, o);
        tmem_ld_wait();
        if (row_ok) {
          if (p.carry_out) {
            float* dst = p.carry_acc + g_row * (static_cast<int64_t>(p.num_heads) * 64) + head * 64 + hlf * 32;
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
      if (row_ok && p.lse != nullptr)
        p.lse[(static_cast<int64_t>(head) * gridDim.z + b) * p.lse_pad + q_row] = m + log2f(l);
      if (row_ok && p.carry_out) {
        p.carry_m[static_cast<int64_t>(head) * p.total_q_rows + g_row] = m;
        p.carry_l[static_cast<int64_t>(head) * p.total_q_rows + g_row] = l;
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

template <int kPoly>
static int launch3(dim3 grid, cudaStream_t st, const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv,
                   const CUtensorMap& tpk, const CUtensorMap& tpv, const Attn3Params& p) {
  auto kern = attn_tc3_kernel<kPoly>;
  static bool configured = false;  // per instantiation
  if (!configured) {
    SR_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, k3Smem));
    configured = true;
  }
  kern<<<grid, k3Threads, k3Smem, st>>>(tq, tk, tv, tpk, tpv, p);
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

// Called by sr_attn_fwd_bf16 (attn_tc.cu) after argument validation.
int attn_tc3_launch(const sr_attn_t* a, int poly, cudaStream_t st) {
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
  Attn3Params p;
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
  p.lse = a->lse;
  p.lse_pad = static_cast<int>(a->lse_pad);
  dim3 grid(static_cast<unsigned>((a->q_len + 128 * k3NQ - 1) / (128 * k3NQ)), static_cast<unsigned>(a->num_heads),
            static_cast<unsigned>(a->batch));
  switch (poly) {
    case 2: return launch3<2>(grid, st, tq, tk, tv, tpk, tpv, p);
#ifdef SR_DEBUG   // the other polynomial-exp2 ratios are tuning experiments, not part of the shipped library
    case 0: return launch3<0>(grid, st, tq, tk, tv, tpk, tpv, p);
    case 1: return launch3<1>(grid, st, tq, tk, tv, tpk, tpv, p);
    case 3: return launch3<3>(grid, st, tq, tk, tv, tpk, tpv, p);
    case 4: return launch3<4>(grid, st, tq, tk, tv, tpk, tpv, p);
    case 8: return launch3<8>(grid, st, tq, tk, tv, tpk, tpv, p);
#endif
  }
  return fail(SR_ERR_SHAPE, "sr_attn_fwd_bf16 (3-tile kernel): SR_ATTN_POLY must be 0..4 or 8 (got %d)", poly);
}

}  // namespace sr
