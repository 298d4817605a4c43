// fp32 SIMT attention with the same structured (prefix + own segment) semantics as attn_tc.cu.
// Used by the camera-head trunk (16 heads x d=128 over Na+Nq <= a few hundred camera tokens, fp32 because the
// reference disables autocast there: sailrecon/models/sail_recon.py:119, camera_head.py:163-166) and by the
// fp32 check mode.  One warp per (query row, head); keys are consumed 32 at a time with an online softmax.
#include "common.cuh"
#include "sr_internal.h"

namespace sr {
extern int g_launch_count;

struct AttnF32Params {
  const float *q, *k, *v, *pk, *pv;
  float* out;
  int64_t ldq, ldkv, ldp, ldo;
  int num_heads, head_dim, batch;
  int64_t q_len, q_batch_stride, kv_len, kv_batch_stride, prefix_len;
  float scale;
};

template <int DV>  // head_dim / 32
__global__ void __launch_bounds__(256)
attn_f32_kernel(const AttnF32Params p) {
  __shared__ float qs[8][DV * 32];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int64_t w = static_cast<int64_t>(blockIdx.x) * 8 + wib;
  const int64_t total = static_cast<int64_t>(p.batch) * p.q_len * p.num_heads;
  if (w >= total) return;
  const int head = static_cast<int>(w % p.num_heads);
  const int64_t qi = (w / p.num_heads) % p.q_len;
  const int b = static_cast<int>(w / (p.num_heads * p.q_len));
  const int D = DV * 32;
  const int64_t qrow = static_cast<int64_t>(b) * p.q_batch_stride + qi;
  const float* qp = p.q + qrow * p.ldq + head * D;
#pragma unroll
  for (int d = 0; d < DV; ++d) qs[wib][lane + 32 * d] = qp[lane + 32 * d];
  __syncwarp();

  float m = -INFINITY, l = 0.f, o[DV];
#pragma unroll
  for (int d = 0; d < DV; ++d) o[d] = 0.f;
  const int64_t nkeys = p.prefix_len + p.kv_len;
  for (int64_t c0 = 0; c0 < nkeys; c0 += 32) {
    const int64_t kk = c0 + lane;
    float s = -INFINITY;
    if (kk < nkeys) {
      const float* kr = kk < p.prefix_len
                            ? p.pk + kk * p.ldp + head * D
                            : p.k + (static_cast<int64_t>(b) * p.kv_batch_stride + (kk - p.prefix_len)) * p.ldkv + head * D;
      float acc = 0.f;
      for (int d = 0; d < D; d += 4) {
        const float4 kv4 = __ldg(reinterpret_cast<const float4*>(kr + d));
        acc = fmaf(qs[wib][d], kv4.x, acc);
        acc = fmaf(qs[wib][d + 1], kv4.y, acc);
        acc = fmaf(qs[wib][d + 2], kv4.z, acc);
        acc = fmaf(qs[wib][d + 3], kv4.w, acc);
      }
      s = acc * p.scale;
    }
    const float m_new = fmaxf(m, warp_max(s));
    const float pj = (kk < nkeys) ? __expf(s - m_new) : 0.f;
    const float alpha = __expf(m - m_new);
    l = l * alpha + warp_sum(pj);
#pragma unroll
    for (int d = 0; d < DV; ++d) o[d] *= alpha;
    const int cnt = static_cast<int>(nkeys - c0 < 32 ? nkeys - c0 : 32);
    for (int j = 0; j < cnt; ++j) {
      const float pw = __shfl_sync(0xffffffffu, pj, j);
      const int64_t kj = c0 + j;
      const float* vr = kj < p.prefix_len
                            ? p.pv + kj * p.ldp + head * D
                            : p.v + (static_cast<int64_t>(b) * p.kv_batch_stride + (kj - p.prefix_len)) * p.ldkv + head * D;
#pragma unroll
      for (int d = 0; d < DV; ++d) o[d] = fmaf(pw, __ldg(vr + lane + 32 * d), o[d]);
    }
    m = m_new;
  }
  float* op = p.out + qrow * p.ldo + head * D;
  const float inv = 1.f / l;
#pragma unroll
  for (int d = 0; d < DV; ++d) op[lane + 32 * d] = o[d] * inv;
}

}  // namespace sr

extern "C" int sr_attn_fwd_f32(const sr_attn_t* a, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(a && a->q && a->k && a->v && a->out, "sr_attn_fwd_f32: null pointer");
  SR_REQUIRE(a->head_dim == 32 || a->head_dim == 64 || a->head_dim == 128, "sr_attn_fwd_f32: head_dim must be 32, 64 or 128");
  SR_REQUIRE(a->num_heads > 0 && a->batch > 0 && a->q_len > 0 && a->kv_len > 0, "sr_attn_fwd_f32: empty problem");
  SR_REQUIRE(!a->carry_in && !a->carry_out, "sr_attn_fwd_f32: carry is only implemented on the bf16 path");
  const int64_t hd = static_cast<int64_t>(a->num_heads) * a->head_dim;
  SR_REQUIRE(a->ldq >= hd && a->ldkv >= hd && a->ldo >= hd && a->ldkv % 4 == 0, "sr_attn_fwd_f32: bad row strides");
  SR_REQUIRE(((reinterpret_cast<uintptr_t>(a->k) | reinterpret_cast<uintptr_t>(a->v)) & 15) == 0, "sr_attn_fwd_f32: k, v must be 16-byte aligned");
  const bool has_prefix = a->pk != nullptr && a->prefix_len > 0;
  if (has_prefix) SR_REQUIRE(a->pv && a->ldp >= hd && a->ldp % 4 == 0 && ((reinterpret_cast<uintptr_t>(a->pk) | reinterpret_cast<uintptr_t>(a->pv)) & 15) == 0, "sr_attn_fwd_f32: bad prefix");
  AttnF32Params p;
  p.q = reinterpret_cast<const float*>(a->q);
  p.k = reinterpret_cast<const float*>(a->k);
  p.v = reinterpret_cast<const float*>(a->v);
  p.pk = has_prefix ? reinterpret_cast<const float*>(a->pk) : nullptr;
  p.pv = has_prefix ? reinterpret_cast<const float*>(a->pv) : nullptr;
  p.out = reinterpret_cast<float*>(a->out);
  p.ldq = a->ldq; p.ldkv = a->ldkv; p.ldp = a->ldp; p.ldo = a->ldo;
  p.num_heads = a->num_heads; p.head_dim = a->head_dim; p.batch = a->batch;
  p.q_len = a->q_len; p.q_batch_stride = a->batch > 1 ? a->q_batch_stride : a->q_len;
  p.kv_len = a->kv_len; p.kv_batch_stride = a->batch > 1 ? a->kv_batch_stride : a->kv_len;
  p.prefix_len = has_prefix ? a->prefix_len : 0;
  p.scale = a->scale;
  const int64_t warps = static_cast<int64_t>(a->batch) * a->q_len * a->num_heads;
  const unsigned grid = static_cast<unsigned>((warps + 7) / 8);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (a->head_dim == 32) attn_f32_kernel<1><<<grid, 256, 0, st>>>(p);
  else if (a->head_dim == 64) attn_f32_kernel<2><<<grid, 256, 0, st>>>(p);
  else attn_f32_kernel<4><<<grid, 256, 0, st>>>(p);
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}
