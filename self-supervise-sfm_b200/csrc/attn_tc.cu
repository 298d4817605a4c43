// Flash attention forward on tcgen05 tensor cores, head_dim 64, bf16 in / fp32 softmax / bf16 out (sm_100a).
//
// Replaces F.scaled_dot_product_attention at sailrecon/layers/attention.py:103-109 for the three aggregator
// block stacks (frame / global / global-reloc, sailrecon/models/aggregator.py:643-769).  The reference's dense
// boolean mask (aggregator.py:302-311, build_allow_block :832-847) is never materialised: its structure is
// "shared scene prefix + own frame", which this kernel walks as two K/V tile ranges.
//
// One CTA = one head x NQ=2 query tiles of 128 rows.  384 threads:
//   warp 0        TMA producer: Q tiles once, then K/V tiles (128 x 64, SWIZZLE_128B) through a 3-stage ring
//   warp 1        MMA issuer (one thread):  S_q = Q_q K^T   (tcgen05.mma SS, M=128 N=128 K=16 x4)  -> TMEM
//                                           O_q += P_q V    (tcgen05.mma SS, M=128 N=64  K=16 x8)  -> TMEM
//                 S_q(j+1) is issued before P_q(j)V(j) so the tensor pipe works under the softmax
//   warp 2        TMEM allocator (S: 2 x 128 columns, O: 2 x 64 columns)
//   warps 4-7     softmax warpgroup of Q tile 0, warps 8-11 of Q tile 1: one thread = one query row
//                 (tcgen05.ld 32x32b), online softmax in the log2 domain with lazy rescale of O
//                 (rescale only when the row max grows by more than 2^8), P written as bf16 into a
//                 K-major SWIZZLE_128B smem tile that the second MMA consumes.
// Optional carry of (m, l, un-normalised O) in/out turns one call into a step of ring attention
// (global blocks sharded over ranks).
#include "common.cuh"
#include "sr_internal.h"

namespace sr {
extern int g_launch_count;

constexpr int kAtNQ = 2;
constexpr int kAtStages = 3;
constexpr int kAtTile = 128 * 64 * 2;                 // 16 KB: one 128 x 64 bf16 tile
constexpr int kAtThreads = 128 + kAtNQ * 128;
constexpr int kAtSmem = kAtNQ * kAtTile               // Q
                        + kAtNQ * 2 * kAtTile         // P (128 x 128 bf16 = two K-major 64-wide sub-tiles)
                        + kAtStages * 2 * kAtTile     // K, V ring
                        + 1024;
constexpr int kAtTmemCols = 512;                      // S0 S1 (128 each) | O0 O1 (64 each) -> 384, pow2 = 512
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
};

__global__ void __launch_bounds__(kAtThreads, 1)
attn_tc_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
               const __grid_constant__ CUtensorMap tmap_v, const __grid_constant__ CUtensorMap tmap_pk,
               const __grid_constant__ CUtensorMap tmap_pv, const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t q_full;
  __shared__ __align__(8) uint64_t k_full[kAtStages], k_empty[kAtStages], v_full[kAtStages], v_empty[kAtStages];
  __shared__ __align__(8) uint64_t s_full[kAtNQ], p_full[kAtNQ], pv_done[kAtNQ];
  __shared__ uint32_t tmem_base_slot;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sP = sQ + kAtNQ * kAtTile;
  uint8_t* sK = sP + kAtNQ * 2 * kAtTile;
  uint8_t* sV = sK + kAtStages * kAtTile;

  const int head = blockIdx.y;
  const int b = blockIdx.z;
  const int q0 = blockIdx.x * (128 * kAtNQ);                 // first query row (within the segment) of this CTA
  const int nq_valid = (q0 + 128 < p.q_len) ? 2 : 1;         // second tile entirely past q_len -> skipped
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
    for (int s = 0; s < kAtStages; ++s) {
      mbar_init(&k_full[s], 1);
      mbar_init(&k_empty[s], 1);
      mbar_init(&v_full[s], 1);
      mbar_init(&v_empty[s], 1);
    }
    for (int q = 0; q < kAtNQ; ++q) {
      mbar_init(&s_full[q], 1);
      mbar_init(&p_full[q], 128);
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

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      mbar_arrive_expect_tx(&q_full, nq_valid * kAtTile);
      for (int q = 0; q < nq_valid; ++q) tma_load_3d(sQ + q * kAtTile, &tmap_q, &q_full, head * 64, q0 + q * 128, b);
      for (int j = 0; j < n_tiles; ++j) {
        const int s = j % kAtStages;
        const uint32_t ph = ((j / kAtStages) & 1) ^ 1;
        const bool pre = j < p.n_prefix_tiles;
        const int r0 = (pre ? j : j - p.n_prefix_tiles) * 128;
        mbar_wait(&k_empty[s], ph);
        mbar_arrive_expect_tx(&k_full[s], kAtTile);
        if (pre) tma_load_2d(sK + s * kAtTile, &tmap_pk, &k_full[s], head * 64, r0);
        else     tma_load_3d(sK + s * kAtTile, &tmap_k, &k_full[s], head * 64, r0, b);
        mbar_wait(&v_empty[s], ph);
        mbar_arrive_expect_tx(&v_full[s], kAtTile);
        if (pre) tma_load_2d(sV + s * kAtTile, &tmap_pv, &v_full[s], head * 64, r0);
        else     tma_load_3d(sV + s * kAtTile, &tmap_v, &v_full[s], head * 64, r0, b);
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc_s = umma_idesc_bf16(128, 128, 0, 0);   // Q (K-major) x K (K-major)
      constexpr uint32_t idesc_o = umma_idesc_bf16(128, 64, 0, 1);    // P (K-major) x V (MN-major: d contiguous)
      auto issue_s = [&](int q, int stage) {
        const uint64_t da = umma_desc_sw128(smem_u32(sQ + q * kAtTile), 16, 1024);
        const uint64_t db = umma_desc_sw128(smem_u32(sK + stage * kAtTile), 16, 1024);
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_ss(tmem_S0 + q * 128, da + 2 * k, db + 2 * k, idesc_s, k != 0 ? 1u : 0u);
        umma_commit(&s_full[q]);
      };
      auto issue_pv = [&](int q, int stage, bool accumulate) {
        const uint32_t pbase = smem_u32(sP + q * 2 * kAtTile);
        const uint32_t vbase = smem_u32(sV + stage * kAtTile);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          // A: P sub-tile k/4 (64 kv columns each), 32 B per K=16 slice inside the swizzle atom
          const uint64_t da = umma_desc_sw128(pbase + (k >> 2) * kAtTile + (k & 3) * 32, 16, 1024);
          // B: V rows [16k, 16k+16) = two 8-row swizzle atoms (SBO 1024 B); one 64-wide MN block (LBO unused)
          const uint64_t db = umma_desc_sw128(vbase + k * 2048, 0, 1024);
          umma_ss(tmem_O0 + q * 64, da, db, idesc_o, (accumulate || k != 0) ? 1u : 0u);
        }
        umma_commit(&pv_done[q]);
      };

      mbar_wait(&q_full, 0);
      mbar_wait(&k_full[0], 0);
      tc_fence_after();
      for (int q = 0; q < nq_valid; ++q) issue_s(q, 0);
      umma_commit(&k_empty[0]);
      for (int j = 0; j < n_tiles; ++j) {
        const int s = j % kAtStages;
        const int sn = (j + 1) % kAtStages;
        const bool has_next = j + 1 < n_tiles;
        if (has_next) mbar_wait(&k_full[sn], ((j + 1) / kAtStages) & 1);
        for (int q = 0; q < nq_valid; ++q) {
          mbar_wait(&p_full[q], j & 1);          // softmax q done with S_q(j); P_q(j) in smem; O_q rescaled
          tc_fence_after();
          if (has_next) issue_s(q, sn);
          if (q == 0) { mbar_wait(&v_full[s], (j / kAtStages) & 1); tc_fence_after(); }
          issue_pv(q, s, p.carry_in != 0 || j > 0);
        }
        if (has_next) umma_commit(&k_empty[sn]);
        umma_commit(&v_empty[s]);
      }
    }
  } else if (warp >= 4) {
    // ------------------------------------------------------------------ softmax warpgroups
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
      uint8_t* myP = sP + q * 2 * kAtTile + (r_tile >> 3) * 1024 + (r_tile & 7) * 128;
      const int sw = r_tile & 7;

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

      for (int j = 0; j < n_tiles; ++j) {
        mbar_wait(&s_full[q], j & 1);
        tc_fence_after();
        uint32_t sraw[128];
        {
          uint32_t(&c0)[32] = *reinterpret_cast<uint32_t(*)[32]>(&sraw[0]);
          uint32_t(&c1)[32] = *reinterpret_cast<uint32_t(*)[32]>(&sraw[32]);
          uint32_t(&c2)[32] = *reinterpret_cast<uint32_t(*)[32]>(&sraw[64]);
          uint32_t(&c3)[32] = *reinterpret_cast<uint32_t(*)[32]>(&sraw[96]);
          tmem_ld32(tS, c0);
          tmem_ld32(tS + 32, c1);
          tmem_ld32(tS + 64, c2);
          tmem_ld32(tS + 96, c3);
          tmem_ld_wait();
        }
        // valid kv columns of this tile (only the last tile of the prefix / own range is partial)
        int valid = 128;
        if (j < p.n_prefix_tiles) {
          if (j == p.n_prefix_tiles - 1) valid = p.prefix_len - j * 128;
        } else if (j == n_tiles - 1) {
          valid = p.kv_len - (j - p.n_prefix_tiles) * 128;
        }
        float mx = -INFINITY;
        if (valid < 128) {
#pragma unroll
          for (int i = 0; i < 128; ++i) {
            float sv = i < valid ? __uint_as_float(sraw[i]) : -INFINITY;
            sraw[i] = __float_as_uint(sv);
            mx = fmaxf(mx, sv);
          }
        } else {
#pragma unroll
          for (int i = 0; i < 128; ++i) mx = fmaxf(mx, __uint_as_float(sraw[i]));
        }
        mx *= p.scale_log2;
        float m_new = m;
        if (mx > m + kLazyRescaleLog2 || m == -INFINITY) m_new = fmaxf(mx, m);
        const float alpha = (m_new != m) ? exp2f(m - m_new) : 1.0f;   // exp2f(-inf) = 0 on the first tile
        float rowsum = 0.f;
        const float neg_m = -m_new;
#pragma unroll
        for (int i = 0; i < 128; ++i) {
          float pv = exp2f(fmaf(__uint_as_float(sraw[i]), p.scale_log2, neg_m));
          rowsum += pv;
          sraw[i] = __float_as_uint(pv);
        }
        l = l * alpha + rowsum;

        if (j > 0) {
          mbar_wait(&pv_done[q], (j - 1) & 1);     // P_q smem and O_q are no longer in use by the tensor core
          tc_fence_after();
        }
#pragma unroll
        for (int ch = 0; ch < 16; ++ch) {
          uint4 w;
          w.x = pack_bf16x2(__uint_as_float(sraw[ch * 8 + 0]), __uint_as_float(sraw[ch * 8 + 1]));
          w.y = pack_bf16x2(__uint_as_float(sraw[ch * 8 + 2]), __uint_as_float(sraw[ch * 8 + 3]));
          w.z = pack_bf16x2(__uint_as_float(sraw[ch * 8 + 4]), __uint_as_float(sraw[ch * 8 + 5]));
          w.w = pack_bf16x2(__uint_as_float(sraw[ch * 8 + 6]), __uint_as_float(sraw[ch * 8 + 7]));
          *reinterpret_cast<uint4*>(myP + (ch >> 3) * kAtTile + (((ch & 7) ^ sw) << 4)) = w;
        }
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
          tmem_st_wait();
        }
        m = m_new;
        fence_proxy_async_smem();
        tc_fence_before();
        mbar_arrive(&p_full[q]);
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

}  // namespace sr

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

  static bool configured = false;
  if (!configured) {
    SR_CHECK_CUDA(cudaFuncSetAttribute(attn_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kAtSmem));
    configured = true;
  }
  dim3 grid(static_cast<unsigned>((a->q_len + 128 * kAtNQ - 1) / (128 * kAtNQ)), static_cast<unsigned>(a->num_heads),
            static_cast<unsigned>(a->batch));
  attn_tc_kernel<<<grid, kAtThreads, kAtSmem, reinterpret_cast<cudaStream_t>(stream)>>>(tq, tk, tv, tpk, tpv, p);
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}
