// fp32 backward pieces of the camera head (sailrecon/heads/camera_head.py:85-186 under torch autograd, as
// train/train_imc.py:380 differentiates it).  The reference runs the head with autocast DISABLED (sail_recon.py:119-121),
// so its backward is fp32 too; the bf16 training kernels of the aggregator (train_rows.cu, attn_bwd.cu) do not apply.
//
// The head works on T = anchors + queries camera-token rows (10 .. 48 in training) against 216 M fp32 parameters: every
// Linear gradient is either a skinny GEMM (dX = dY W: sr_gemm_f32 on a transposed copy of W made by sr_transpose_f32) or
// an outer-product sum over T rows (dW = dY^T X: sr_gemm_f32 with the contraction over T), both bound by streaming the
// weights / writing their gradients once.  The kernels here are the row-wise rest:
//   sr_transpose_f32            dst[c, r] = src[r, c]  (32 x 32 tiles through shared memory, both sides coalesced)
//   sr_colsum_f32               out[c] (+)= sum_r a[r, c] * b[r, c]   (bias and LayerScale-gamma gradients; b NULL = 1)
//   sr_act_f32 / sr_act_bwd_f32 GELU (erf form, mlp.py:23 nn.GELU) and SiLU (camera_head.py:72) on saved pre-activations
//   sr_ls_residual_f32f32       out = resid + gamma[c] * z  (block.py:110-111 with the Linear output kept for d gamma)
//   sr_layernorm_bwd_f32        d LayerNorm with fp32 incoming gradient (token_norm, trunk_norm, the blocks' norm1 / norm2)
//   sr_adaln_modulate_bwd_f32   backward of gate * (LN_noaffine(x) * (1 + scale) + shift) + x  (camera_head.py:151-161)
//   sr_attn_small_bwd_f32       attention backward for T <= 128 rows with the build_lr_mask structure (camera_head.py:197-228):
//                               column c is visible to row r iff c < n_anchor or c == r; one block per head, P and dS in
//                               shared memory, probabilities recomputed from q, k
//   sr_pose_act_bwd_f32         backward of activate_pose (head_act.py:12-37: translation / quaternion linear, FoV relu)
#include "common.cuh"
#include "sr_internal.h"

namespace sr {
extern int g_launch_count;

__global__ void transpose_f32_kernel(const float* __restrict__ src, int64_t lds, float* __restrict__ dst, int64_t ldd, int R,
                                     int C) {
  __shared__ float tile[32][33];
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += 8) {
    const int r = r0 + i, c = c0 + threadIdx.x;
    if (r < R && c < C) tile[i][threadIdx.x] = src[static_cast<int64_t>(r) * lds + c];
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += 8) {
    const int c = c0 + i, r = r0 + threadIdx.x;
    if (r < R && c < C) dst[static_cast<int64_t>(c) * ldd + r] = tile[threadIdx.x][i];
  }
}

// rows are few (T): one thread per column walks them; consecutive threads on consecutive columns
__global__ void colsum_f32_kernel(const float* __restrict__ a, int64_t lda, const float* __restrict__ b, int64_t ldb,
                                  float* __restrict__ out, int R, int C, int accumulate) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  float s = 0.f;
  for (int r = 0; r < R; ++r) s += a[static_cast<int64_t>(r) * lda + c] * (b ? b[static_cast<int64_t>(r) * ldb + c] : 1.0f);
  out[c] = accumulate ? out[c] + s : s;
}

__device__ __forceinline__ float gelu_erf_grad(float u) {
  return 0.5f * (1.0f + erff(u * 0.70710678118654752f)) + u * 0.39894228040143268f * expf(-0.5f * u * u);
}
__device__ __forceinline__ float sigmoid_f32(float u) { return 1.0f / (1.0f + expf(-u)); }

// kind 0 = GELU, 1 = SiLU.  dg == nullptr: forward out = act(u); else out = dg * act'(u)
__global__ void act_f32_kernel(const float* __restrict__ u, const float* __restrict__ dg, float* __restrict__ out, int64_t n,
                               int kind) {
  const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= n) return;
  const float x = u[i];
  float y;
  if (dg == nullptr) {
    y = kind == 0 ? gelu_erf(x) : x * sigmoid_f32(x);
  } else {
    const float s = sigmoid_f32(x);
    y = dg[i] * (kind == 0 ? gelu_erf_grad(x) : s * (1.0f + x * (1.0f - s)));
  }
  out[i] = y;
}

__global__ void ls_residual_f32f32_kernel(const float* __restrict__ resid, const float* __restrict__ z,
                                          const float* __restrict__ gamma, float* __restrict__ out, int64_t n, int C) {
  const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= n) return;
  out[i] = resid[i] + (gamma ? gamma[i % C] : 1.0f) * z[i];
}

__device__ __forceinline__ float block_sum(float v, float* sh) {     // blockDim.x == 256; all threads get the sum
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
  __syncthreads();
  float s = 0.f;
#pragma unroll
  for (int w = 0; w < 8; ++w) s += sh[w];
  return s;
}

// One block per row.  dx = (dres ? dres : 0) + rstd * (g - mean(g) - xhat * mean(g * xhat)),  g = dy * weight.
// dweight[c] += dy * xhat, dbias[c] += dy (atomics over the few rows).
__global__ void __launch_bounds__(256)
layernorm_bwd_f32_kernel(const float* __restrict__ x, const float* __restrict__ weight, const float* __restrict__ dy,
                         const float* __restrict__ dres, float* __restrict__ dx, float* __restrict__ dweight,
                         float* __restrict__ dbias, int C, float eps) {
  __shared__ float sh[8];
  const int64_t row = blockIdx.x;
  const float* xr = x + row * C;
  const float* gr = dy + row * C;
  float s = 0.f;
  for (int c = threadIdx.x; c < C; c += 256) s += xr[c];
  const float mean = block_sum(s, sh) / C;
  float q = 0.f;
  for (int c = threadIdx.x; c < C; c += 256) { const float d = xr[c] - mean; q += d * d; }
  const float rstd = rsqrtf(block_sum(q, sh) / C + eps);
  float sg = 0.f, sgx = 0.f;
  for (int c = threadIdx.x; c < C; c += 256) {
    const float xh = (xr[c] - mean) * rstd, g = gr[c] * (weight ? weight[c] : 1.0f);
    sg += g;
    sgx += g * xh;
  }
  const float mg = block_sum(sg, sh) / C, mgx = block_sum(sgx, sh) / C;
  for (int c = threadIdx.x; c < C; c += 256) {
    const float xh = (xr[c] - mean) * rstd, d = gr[c], g = d * (weight ? weight[c] : 1.0f);
    const float v = rstd * (g - mg - xh * mgx);
    dx[row * C + c] = dres ? dres[row * C + c] + v : v;
    if (dweight) atomicAdd(&dweight[c], d * xh);
    if (dbias) atomicAdd(&dbias[c], d);
  }
}

// y = gate * (n * (1 + scale) + shift) + x,  n = LN_noaffine(x);  mod = [shift | scale | gate]
// d_mod = [dy * gate | dy * gate * n | dy * m];  dx (+)= LN_noaffine_bwd(dy * gate * (1 + scale)) + dy
__global__ void __launch_bounds__(256)
adaln_modulate_bwd_f32_kernel(const float* __restrict__ x, const float* __restrict__ mod, const float* __restrict__ dy,
                              float* __restrict__ dx, float* __restrict__ dmod, int C, float eps, int accumulate_dx) {
  __shared__ float sh[8];
  const int64_t row = blockIdx.x;
  const float* xr = x + row * C;
  const float* gr = dy + row * C;
  const float* shf = mod + row * 3 * C;
  const float* scl = shf + C;
  const float* gat = scl + C;
  float s = 0.f;
  for (int c = threadIdx.x; c < C; c += 256) s += xr[c];
  const float mean = block_sum(s, sh) / C;
  float q = 0.f;
  for (int c = threadIdx.x; c < C; c += 256) { const float d = xr[c] - mean; q += d * d; }
  const float rstd = rsqrtf(block_sum(q, sh) / C + eps);
  float sg = 0.f, sgx = 0.f;
  for (int c = threadIdx.x; c < C; c += 256) {
    const float n = (xr[c] - mean) * rstd, dm = gr[c] * gat[c], g = dm * (1.0f + scl[c]);
    dmod[row * 3 * C + c] = dm;
    dmod[row * 3 * C + C + c] = dm * n;
    dmod[row * 3 * C + 2 * C + c] = gr[c] * (n * (1.0f + scl[c]) + shf[c]);
    sg += g;
    sgx += g * n;
  }
  const float mg = block_sum(sg, sh) / C, mgx = block_sum(sgx, sh) / C;
  for (int c = threadIdx.x; c < C; c += 256) {
    const float n = (xr[c] - mean) * rstd, g = gr[c] * gat[c] * (1.0f + scl[c]);
    const float v = rstd * (g - mg - n * mgx) + gr[c];
    dx[row * C + c] = accumulate_dx ? dx[row * C + c] + v : v;
  }
}

// One block (256 threads) per head.  q, k, v, dout, dq, dk, dv are row views with the head at columns [h * D, (h + 1) * D).
__global__ void __launch_bounds__(256)
attn_small_bwd_f32_kernel(const float* __restrict__ q, const float* __restrict__ k, const float* __restrict__ v, int64_t ldqkv,
                          const float* __restrict__ dout, int64_t lddo, float* __restrict__ dq, float* __restrict__ dk,
                          float* __restrict__ dv, int64_t ldd, int T, int D, int n_anchor, float scale) {
  extern __shared__ float sm[];
  float* P = sm;               // [T, T]
  float* dS = sm + T * T;      // [T, T]
  const int h = blockIdx.x, tid = threadIdx.x;
  const int64_t off = static_cast<int64_t>(h) * D;
  auto visible = [&](int r, int c) { return c < n_anchor || c == r; };
  // S = scale * q k^T (masked), then row softmax
  for (int e = tid; e < T * T; e += 256) {
    const int r = e / T, c = e - r * T;
    float s = -INFINITY;
    if (visible(r, c)) {
      s = 0.f;
      for (int d = 0; d < D; ++d) s += q[r * ldqkv + off + d] * k[c * ldqkv + off + d];
      s *= scale;
    }
    P[e] = s;
  }
  __syncthreads();
  for (int r = tid; r < T; r += 256) {
    float m = -INFINITY;
    for (int c = 0; c < T; ++c) m = fmaxf(m, P[r * T + c]);
    float l = 0.f;
    for (int c = 0; c < T; ++c) { const float p = expf(P[r * T + c] - m); P[r * T + c] = p; l += p; }
    const float inv = 1.0f / l;
    for (int c = 0; c < T; ++c) P[r * T + c] *= inv;
  }
  __syncthreads();
  // dP = dout v^T ;  dS = P * (dP - rowsum(dP * P))
  for (int e = tid; e < T * T; e += 256) {
    const int r = e / T, c = e - r * T;
    float s = 0.f;
    if (visible(r, c))
      for (int d = 0; d < D; ++d) s += dout[r * lddo + off + d] * v[c * ldqkv + off + d];
    dS[e] = s;
  }
  __syncthreads();
  for (int r = tid; r < T; r += 256) {
    float dot = 0.f;
    for (int c = 0; c < T; ++c) dot += dS[r * T + c] * P[r * T + c];
    for (int c = 0; c < T; ++c) dS[r * T + c] = P[r * T + c] * (dS[r * T + c] - dot) * scale;
  }
  __syncthreads();
  // dv = P^T dout ; dk = dS^T q ; dq = dS k
  for (int e = tid; e < T * D; e += 256) {
    const int r = e / D, d = e - r * D;
    float av = 0.f, ak = 0.f, aq = 0.f;
    for (int c = 0; c < T; ++c) {
      av += P[c * T + r] * dout[c * lddo + off + d];
      ak += dS[c * T + r] * q[c * ldqkv + off + d];
      aq += dS[r * T + c] * k[c * ldqkv + off + d];
    }
    dv[r * ldd + off + d] = av;
    dk[r * ldd + off + d] = ak;
    dq[r * ldd + off + d] = aq;
  }
}

// activate_pose backward: columns 0..6 (translation, quaternion) pass through, FoV columns 7, 8 are relu'd
__global__ void pose_act_bwd_f32_kernel(const float* __restrict__ pred, const float* __restrict__ dact, float* __restrict__ dpred,
                                        int64_t n) {
  const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= n) return;
  const int c = static_cast<int>(i % 9);
  dpred[i] = (c >= 7 && !(pred[i] > 0.f)) ? 0.f : dact[i];
}

}  // namespace sr

namespace {
inline cudaStream_t hs(void* s) { return reinterpret_cast<cudaStream_t>(s); }
inline unsigned nblk(int64_t n, int per) { return static_cast<unsigned>((n + per - 1) / per); }
}  // namespace

extern "C" int sr_transpose_f32(const float* src, int64_t lds, float* dst, int64_t ldd, int64_t R, int64_t C, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  if (R == 0 || C == 0) return SR_OK;
  SR_REQUIRE(src && dst && R > 0 && C > 0 && lds >= C && ldd >= R, "sr_transpose_f32: bad arguments");
  SR_REQUIRE((R + 31) / 32 <= 65535 && C < (1ll << 31), "sr_transpose_f32: matrix too large");
  dim3 grid(nblk(C, 32), nblk(R, 32));
  transpose_f32_kernel<<<grid, dim3(32, 8), 0, hs(stream)>>>(src, lds, dst, ldd, static_cast<int>(R), static_cast<int>(C));
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

extern "C" int sr_colsum_f32(const float* a, int64_t lda, const float* b, int64_t ldb, float* out, int64_t R, int64_t C,
                             int32_t accumulate, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  if (C == 0) return SR_OK;
  SR_REQUIRE(a && out && R >= 0 && R < (1 << 30) && C > 0 && C < (1 << 30) && lda >= C && (b == nullptr || ldb >= C),
             "sr_colsum_f32: bad arguments");
  colsum_f32_kernel<<<nblk(C, 128), 128, 0, hs(stream)>>>(a, lda, b, ldb, out, static_cast<int>(R), static_cast<int>(C), accumulate);
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

extern "C" int sr_act_f32(const float* u, float* out, int64_t n, int32_t kind, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  if (n == 0) return SR_OK;
  SR_REQUIRE(u && out && n > 0 && n < (1ll << 38) && (kind == 0 || kind == 1), "sr_act_f32: bad arguments (kind 0 = GELU, 1 = SiLU)");
  act_f32_kernel<<<nblk(n, 256), 256, 0, hs(stream)>>>(u, nullptr, out, n, kind);
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

extern "C" int sr_act_bwd_f32(const float* u, const float* dg, float* du, int64_t n, int32_t kind, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  if (n == 0) return SR_OK;
  SR_REQUIRE(u && dg && du && n > 0 && n < (1ll << 38) && (kind == 0 || kind == 1), "sr_act_bwd_f32: bad arguments");
  act_f32_kernel<<<nblk(n, 256), 256, 0, hs(stream)>>>(u, dg, du, n, kind);
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

extern "C" int sr_ls_residual_f32f32(const float* resid, const float* z, const float* gamma, float* out, int64_t rows, int64_t C,
                                     void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  const int64_t n = rows * C;
  if (n == 0) return SR_OK;
  SR_REQUIRE(resid && z && out && rows > 0 && C > 0 && C < (1 << 30) && n < (1ll << 38), "sr_ls_residual_f32f32: bad arguments");
  ls_residual_f32f32_kernel<<<nblk(n, 256), 256, 0, hs(stream)>>>(resid, z, gamma, out, n, static_cast<int>(C));
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

extern "C" int sr_layernorm_bwd_f32(const float* x, const float* weight, const float* dy, const float* dres, float* dx,
                                    float* dweight, float* dbias, int64_t rows, int64_t C, float eps, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  if (rows == 0) return SR_OK;
  SR_REQUIRE(x && dy && dx && rows > 0 && rows < (1ll << 31) && C > 0 && C < (1 << 24), "sr_layernorm_bwd_f32: bad arguments");
  layernorm_bwd_f32_kernel<<<static_cast<unsigned>(rows), 256, 0, hs(stream)>>>(x, weight, dy, dres, dx, dweight, dbias,
                                                                                static_cast<int>(C), eps);
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

extern "C" int sr_adaln_modulate_bwd_f32(const float* x, const float* mod, const float* dy, float* dx, float* dmod, int64_t rows,
                                         int64_t C, float eps, int32_t accumulate_dx, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  if (rows == 0) return SR_OK;
  SR_REQUIRE(x && mod && dy && dx && dmod && rows > 0 && rows < (1ll << 31) && C > 0 && C < (1 << 24),
             "sr_adaln_modulate_bwd_f32: bad arguments");
  adaln_modulate_bwd_f32_kernel<<<static_cast<unsigned>(rows), 256, 0, hs(stream)>>>(x, mod, dy, dx, dmod, static_cast<int>(C), eps,
                                                                                     accumulate_dx);
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

extern "C" int sr_attn_small_bwd_f32(const float* q, const float* k, const float* v, int64_t ldqkv, const float* dout, int64_t lddo,
                                     float* dq, float* dk, float* dv, int64_t ldd, int64_t T, int32_t num_heads, int32_t head_dim,
                                     int64_t n_anchor, float scale, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  if (T == 0) return SR_OK;
  SR_REQUIRE(q && k && v && dout && dq && dk && dv, "sr_attn_small_bwd_f32: null pointer");
  SR_REQUIRE(T > 0 && T <= 128, "sr_attn_small_bwd_f32: at most 128 camera-token rows (got %lld)", static_cast<long long>(T));
  SR_REQUIRE(num_heads > 0 && head_dim > 0 && head_dim <= 256 && n_anchor >= 0 && n_anchor <= T, "sr_attn_small_bwd_f32: bad shape");
  const int64_t width = static_cast<int64_t>(num_heads) * head_dim;
  SR_REQUIRE(ldqkv >= width && lddo >= width && ldd >= width, "sr_attn_small_bwd_f32: leading dimensions too small");
  const size_t smem = sizeof(float) * 2 * T * T;
  if (smem > 48 * 1024)
    SR_CHECK_CUDA(cudaFuncSetAttribute(attn_small_bwd_f32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
  attn_small_bwd_f32_kernel<<<num_heads, 256, smem, hs(stream)>>>(q, k, v, ldqkv, dout, lddo, dq, dk, dv, ldd, static_cast<int>(T),
                                                                  head_dim, static_cast<int>(n_anchor), scale);
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

extern "C" int sr_pose_act_bwd_f32(const float* pred, const float* d_activated, float* d_pred, int64_t rows, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  if (rows == 0) return SR_OK;
  SR_REQUIRE(pred && d_activated && d_pred && rows > 0 && rows < (1ll << 31), "sr_pose_act_bwd_f32: bad arguments");
  pose_act_bwd_f32_kernel<<<nblk(rows * 9, 256), 256, 0, hs(stream)>>>(pred, d_activated, d_pred, rows * 9);
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}
