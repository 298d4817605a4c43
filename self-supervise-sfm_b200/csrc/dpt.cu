// Row kernels of the DPT dense heads (sailrecon/heads/dpt_head.py:22-344, called at
// sailrecon/models/sail_recon.py:129-136 and :246-252 with autocast disabled).
//
// Layout: every feature map is CHANNELS-LAST fp32, i.e. a row matrix [N*H*W, C] — the same row-matrix convention as the
// token buffers of the aggregator, so every convolution is a GEMM on the tcgen05 kernel of gemm_tc.cu:
//   1x1 conv / ConvTranspose(k = stride)      rows as they are               x  W[Cout(*k*k), Cin]
//   3x3 conv, stride 1, padding 1             zero-bordered map, SHIFTED rows (gemm_tc.cu AMN == 2, sr_conv3x3_bf16): no im2col
//   3x3 conv, stride 2 (one small layer)      im2col rows [.., 9*Cin]        x  W[Cout, (ky, kx, Cin)]
// fp32 semantics on a bf16 tensor core: each fp32 operand is split into hi = bf16(x) and lo = bf16(x - hi); the product is
// hi*hi + lo*hi + hi*lo accumulated in fp32 (the dropped lo*lo term is 2^-16 relative), evaluated as TWO GEMMs:
//   out  = [A_hi | A_lo] x [W_hi | W_hi]^T (+ bias, + residual)      K' = 2K
//   out += A_hi x W_lo^T                                            exact fp32 accumulate (sr_epilogue.exact_residual)
// so the im2col / split kernel writes rows [hi(K) | lo(K)].  That is MORE accurate than the reference's own default GPU path
// (cuDNN with TF32 allowed: 10-bit mantissa) and meets the 1e-4 fp32 check against the TF32-free reference.
//
// All kernels here are HBM-bound gathers / scatters: 16-byte accesses along the channel axis, one thread per 8 (im2col) or
// 4 channels, grid-stride free (one element group per thread, 256 threads per CTA).
#include "common.cuh"
#include "sr_internal.h"

namespace sr {
extern int g_launch_count;

__device__ __forceinline__ void split8(const float (&v)[8], uint4& hi, uint4& lo) {
  uint32_t h[4], l[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const __nv_bfloat16 h0 = __float2bfloat16_rn(v[2 * i]), h1 = __float2bfloat16_rn(v[2 * i + 1]);
    const float r0 = v[2 * i] - __bfloat162float(h0), r1 = v[2 * i + 1] - __bfloat162float(h1);
    h[i] = pack_bf16x2(__bfloat162float(h0), __bfloat162float(h1));
    l[i] = pack_bf16x2(r0, r1);
  }
  hi = make_uint4(h[0], h[1], h[2], h[3]);
  lo = make_uint4(l[0], l[1], l[2], l[3]);
}

// x fp32 [N, H, W, C] (row stride ldx per pixel) -> out bf16 [N*Ho*Wo, 2*k*k*C]: columns [(ky*k + kx)*C + c] = hi,
// the same + k*k*C = lo.  relu: max(x, 0) first (the ResidualConvUnit activates BEFORE its convolutions, dpt_head.py:461-472).
__global__ void __launch_bounds__(256)
im2col_split_kernel(const float* __restrict__ x, int64_t ldx, __nv_bfloat16* __restrict__ out, int N, int H, int W, int C,
                    int Ho, int Wo, int k, int stride, int pad, int relu) {
  const int c8 = C >> 3;
  const int64_t total = static_cast<int64_t>(N) * Ho * Wo * k * k * c8;
  const int64_t idx = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (idx >= total) return;
  const int cg = static_cast<int>(idx % c8);
  int64_t r = idx / c8;
  const int tap = static_cast<int>(r % (k * k));
  r /= k * k;                                               // output pixel (n, yo, xo)
  const int xo = static_cast<int>(r % Wo);
  const int64_t r2 = r / Wo;
  const int yo = static_cast<int>(r2 % Ho);
  const int n = static_cast<int>(r2 / Ho);
  const int ky = tap / k, kx = tap - ky * k;
  const int yi = yo * stride - pad + ky, xi = xo * stride - pad + kx;
  float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (yi >= 0 && yi < H && xi >= 0 && xi < W) {
    const float4* src = reinterpret_cast<const float4*>(x + ((static_cast<int64_t>(n) * H + yi) * W + xi) * ldx + cg * 8);
    const float4 a = __ldg(src), b = __ldg(src + 1);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    if (relu) {
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] = fmaxf(v[i], 0.f);
    }
  }
  uint4 hi, lo;
  split8(v, hi, lo);
  const int64_t K = static_cast<int64_t>(k) * k * C;
  __nv_bfloat16* dst = out + r * (2 * K) + static_cast<int64_t>(tap) * C + cg * 8;
  *reinterpret_cast<uint4*>(dst) = hi;
  *reinterpret_cast<uint4*>(dst + K) = lo;
}

// F.interpolate(mode="bilinear", align_corners=True) on channels-last maps (custom_interpolate, dpt_head.py:565-594).
__global__ void __launch_bounds__(256)
bilinear_nhwc_kernel(const float* __restrict__ x, float* __restrict__ out, int N, int Hi, int Wi, int Ho, int Wo, int C,
                     float sy, float sx, int in_pad) {
  const int c4 = C >> 2;
  const int64_t total = static_cast<int64_t>(N) * Ho * Wo * c4;
  const int64_t idx = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (idx >= total) return;
  const int cg = static_cast<int>(idx % c4);
  int64_t r = idx / c4;
  const int xo = static_cast<int>(r % Wo);
  r /= Wo;
  const int yo = static_cast<int>(r % Ho);
  const int n = static_cast<int>(r / Ho);
  const float fy = sy * yo, fx = sx * xo;                   // area_pixel_compute_source_index, align_corners = true
  const int y0 = static_cast<int>(fy), x0 = static_cast<int>(fx);
  const int y1 = y0 + (y0 < Hi - 1 ? 1 : 0), x1 = x0 + (x0 < Wi - 1 ? 1 : 0);
  const float ly1 = fy - y0, lx1 = fx - x0, ly0 = 1.f - ly1, lx0 = 1.f - lx1;
  const int Wp = Wi + 2 * in_pad;                           // the source may carry the zero border of the shifted-GEMM convolutions
  const float* base = x + (static_cast<int64_t>(n) * (Hi + 2 * in_pad) * Wp + static_cast<int64_t>(in_pad) * Wp + in_pad) * C + cg * 4;
  const float4 v00 = __ldg(reinterpret_cast<const float4*>(base + (static_cast<int64_t>(y0) * Wp + x0) * C));
  const float4 v01 = __ldg(reinterpret_cast<const float4*>(base + (static_cast<int64_t>(y0) * Wp + x1) * C));
  const float4 v10 = __ldg(reinterpret_cast<const float4*>(base + (static_cast<int64_t>(y1) * Wp + x0) * C));
  const float4 v11 = __ldg(reinterpret_cast<const float4*>(base + (static_cast<int64_t>(y1) * Wp + x1) * C));
  float4 o;
  o.x = ly0 * (lx0 * v00.x + lx1 * v01.x) + ly1 * (lx0 * v10.x + lx1 * v11.x);
  o.y = ly0 * (lx0 * v00.y + lx1 * v01.y) + ly1 * (lx0 * v10.y + lx1 * v11.y);
  o.z = ly0 * (lx0 * v00.z + lx1 * v01.z) + ly1 * (lx0 * v10.z + lx1 * v11.z);
  o.w = ly0 * (lx0 * v00.w + lx1 * v01.w) + ly1 * (lx0 * v10.w + lx1 * v11.w);
  *reinterpret_cast<float4*>(out + ((static_cast<int64_t>(n) * Ho + yo) * Wo + xo) * C + cg * 4) = o;
}

// ConvTranspose2d with kernel = stride = k, padding 0 (dpt_head.py:85-102): g [N*H*W, k*k*C] (columns (ky, kx, c), the GEMM
// output) -> out [N, H*k, W*k, C] + bias[c].
__global__ void __launch_bounds__(256)
pixel_shuffle_kernel(const float* __restrict__ g, const float* __restrict__ bias, float* __restrict__ out, int N, int H, int W,
                     int C, int k) {
  const int c4 = C >> 2;
  const int64_t total = static_cast<int64_t>(N) * H * W * k * k * c4;
  const int64_t idx = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (idx >= total) return;
  const int cg = static_cast<int>(idx % c4);
  int64_t r = idx / c4;
  const int tap = static_cast<int>(r % (k * k));
  r /= k * k;                                               // input pixel
  const int xi = static_cast<int>(r % W);
  const int64_t r2 = r / W;
  const int yi = static_cast<int>(r2 % H);
  const int n = static_cast<int>(r2 / H);
  const int ky = tap / k, kx = tap - ky * k;
  float4 v = __ldg(reinterpret_cast<const float4*>(g + r * (static_cast<int64_t>(k) * k * C) + static_cast<int64_t>(tap) * C + cg * 4));
  if (bias != nullptr) {
    const float4 b = __ldg(reinterpret_cast<const float4*>(bias + cg * 4));
    v.x += b.x; v.y += b.y; v.z += b.z; v.w += b.w;
  }
  const int64_t o = ((static_cast<int64_t>(n) * H * k + yi * k + ky) * (static_cast<int64_t>(W) * k) + xi * k + kx) * C + cg * 4;
  *reinterpret_cast<float4*>(out + o) = v;
}

// out = act(a + alpha * b[i mod period])  (period = n: plain add; period = H*W*C: the broadcast positional embedding; b may be
// NULL; act = ReLU when relu_out: the ResidualConvUnit's in-place nn.ReLU(inplace=True), dpt_head.py:357,461)
__global__ void __launch_bounds__(256)
add_bcast_kernel(const float* __restrict__ a, const float* __restrict__ b, float* __restrict__ out, int64_t n4, int64_t period4,
                 float alpha, int relu_out) {
  const int64_t idx = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (idx >= n4) return;
  float4 x = __ldg(reinterpret_cast<const float4*>(a) + idx);
  if (b != nullptr) {
    const float4 y = __ldg(reinterpret_cast<const float4*>(b) + (idx % period4));
    x = make_float4(x.x + alpha * y.x, x.y + alpha * y.y, x.z + alpha * y.z, x.w + alpha * y.w);
  }
  if (relu_out) x = make_float4(fmaxf(x.x, 0.f), fmaxf(x.y, 0.f), fmaxf(x.z, 0.f), fmaxf(x.w, 0.f));
  reinterpret_cast<float4*>(out)[idx] = x;
}

// create_uv_grid + position_grid_to_embed (heads/utils.py:11-123): table [H, W, C] fp32, evaluated in double like the
// reference (its omega / outer product / sin / cos run in float64, then .float()).
__global__ void uv_posembed_kernel(float* __restrict__ out, int H, int W, int C, double aspect, double omega0) {
  const int64_t idx = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (idx >= static_cast<int64_t>(H) * W * C) return;
  const int c = static_cast<int>(idx % C);
  const int64_t r = idx / C;
  const int xg = static_cast<int>(r % W), yg = static_cast<int>(r / W);
  const double diag = sqrt(aspect * aspect + 1.0);
  const double span_x = aspect / diag, span_y = 1.0 / diag;
  // torch.linspace(left, right, steps) in fp32: start + i * step for the first half, end - (steps - 1 - i) * step after it
  auto lin = [](double lo, double hi, int steps, int i) -> double {
    const float start = static_cast<float>(lo), end = static_cast<float>(hi);
    if (steps == 1) return start;
    const float step = (end - start) / static_cast<float>(steps - 1);
    return i < steps / 2 ? static_cast<double>(start + step * i) : static_cast<double>(end - step * (steps - 1 - i));
  };
  const double u = lin(-span_x * (W - 1) / W, span_x * (W - 1) / W, W, xg);
  const double v = lin(-span_y * (H - 1) / H, span_y * (H - 1) / H, H, yg);
  const int half = C / 2, quarter = C / 4;                  // [sin_x | cos_x | sin_y | cos_y], each C/4 wide
  const int which = c / half, j = c % half;
  const double pos = which == 0 ? u : v;
  const int f = j % quarter;
  const double omega = 1.0 / pow(omega0, static_cast<double>(f) / static_cast<double>(quarter));
  const double ang = pos * omega;
  out[idx] = static_cast<float>(j < quarter ? sin(ang) : cos(ang));
}

// output_conv2[1:] + activate_head (dpt_head.py:150-160,335-340; head_act.py:63-114): h [px, ldh] = output of the 3x3 conv
// (pre-ReLU, `hid` channels) -> ReLU -> 1x1 conv to out_dim -> xyz = act(out[:-1]), conf = conf_act(out[-1]).
// act: 0 = inv_log (sign(y) expm1(|y|)), 1 = exp, 2 = linear, 3 = relu, 4 = sigmoid.  conf_act: 0 = expp1, 1 = expp0, 2 = sigmoid.
template <int OUT>
__global__ void __launch_bounds__(256)
dpt_out_kernel(const float* __restrict__ h, int64_t ldh, const float* __restrict__ w, const float* __restrict__ bias,
               float* __restrict__ pred, float* __restrict__ conf, int64_t px, int hid, int act, int conf_act, int map_h, int map_w,
               int pad) {
  extern __shared__ float sw[];                             // [OUT, hid] + [OUT]
  for (int i = threadIdx.x; i < OUT * hid + OUT; i += blockDim.x) sw[i] = i < OUT * hid ? w[i] : bias[i - OUT * hid];
  __syncthreads();
  const int64_t p = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (p >= px) return;
  float acc[OUT];
#pragma unroll
  for (int o = 0; o < OUT; ++o) acc[o] = 0.f;
  int64_t row = p;                                          // h may be a zero-bordered map [S, H + 2, W + 2, ldh]
  if (pad) {
    const int64_t hw = static_cast<int64_t>(map_h) * map_w;
    const int64_t s_ = p / hw, r_ = p - s_ * hw;
    const int y = static_cast<int>(r_ / map_w), x = static_cast<int>(r_ - static_cast<int64_t>(y) * map_w);
    row = (s_ * (map_h + 2) + y + 1) * (map_w + 2) + x + 1;
  }
  const float4* src = reinterpret_cast<const float4*>(h + row * ldh);
  for (int c = 0; c < hid; c += 4) {
    float4 v = __ldg(src + (c >> 2));
    v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f);
#pragma unroll
    for (int o = 0; o < OUT; ++o)
      acc[o] = fmaf(v.w, sw[o * hid + c + 3], fmaf(v.z, sw[o * hid + c + 2], fmaf(v.y, sw[o * hid + c + 1], fmaf(v.x, sw[o * hid + c], acc[o]))));
  }
#pragma unroll
  for (int o = 0; o < OUT - 1; ++o) {
    const float y = acc[o] + sw[OUT * hid + o];
    float r;
    if (act == 0) r = copysignf(expm1f(fabsf(y)), y);
    else if (act == 1) r = expf(y);
    else if (act == 3) r = fmaxf(y, 0.f);
    else if (act == 4) r = 1.f / (1.f + expf(-y));
    else r = y;
    pred[p * (OUT - 1) + o] = r;
  }
  const float yc = acc[OUT - 1] + sw[OUT * hid + OUT - 1];
  conf[p] = conf_act == 0 ? 1.f + expf(yc) : (conf_act == 1 ? expf(yc) : 1.f / (1.f + expf(-yc)));
}

// x fp32 channels-last [S, H + 2 in_pad, W + 2 in_pad, ldx >= C] -> out bf16 [S, H + 2, W + 2, hi(C) | lo(C)] with a ZERO
// border: the operand layout of the shifted-GEMM 3x3 convolution (gemm_tc.cu, AMN == 2).  8 B per element moved instead of the
// 4 + 36 B of a materialised im2col.  relu: max(x, 0) first.
__global__ void __launch_bounds__(256)
split_pad_kernel(const float* __restrict__ x, int64_t ldx, int in_pad, __nv_bfloat16* __restrict__ out, int S, int H, int W, int C,
                 int relu) {
  const int c8 = C >> 3;
  const int Hp = H + 2, Wp = W + 2;
  const int64_t total = static_cast<int64_t>(S) * Hp * Wp * c8;
  const int64_t idx = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (idx >= total) return;
  const int cg = static_cast<int>(idx % c8);
  const int64_t r = idx / c8;                                // padded output pixel
  const int xp = static_cast<int>(r % Wp);
  const int64_t r2 = r / Wp;
  const int yp = static_cast<int>(r2 % Hp);
  const int s = static_cast<int>(r2 / Hp);
  float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (yp >= 1 && yp <= H && xp >= 1 && xp <= W) {
    const int Wi = W + 2 * in_pad;
    const int64_t src_px = (static_cast<int64_t>(s) * (H + 2 * in_pad) + (yp - 1 + in_pad)) * Wi + (xp - 1 + in_pad);
    const float4* src = reinterpret_cast<const float4*>(x + src_px * ldx + cg * 8);
    const float4 a = __ldg(src), b = __ldg(src + 1);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    if (relu) {
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] = fmaxf(v[i], 0.f);
    }
  }
  uint4 hi, lo;
  split8(v, hi, lo);
  __nv_bfloat16* dst = out + r * (2 * static_cast<int64_t>(C)) + cg * 8;
  *reinterpret_cast<uint4*>(dst) = hi;
  *reinterpret_cast<uint4*>(dst + C) = lo;
}

// out = act(a + b) between channels-last maps of one spatial size that may or may not carry the one-pixel border
// (pad_a / pad_b / pad_out in {0, 1}; b may be NULL; border pixels of a padded output are written as 0).
__global__ void __launch_bounds__(256)
map_add_kernel(const float* __restrict__ a, int pad_a, const float* __restrict__ b, int pad_b, float* __restrict__ out, int pad_out,
               int S, int H, int W, int C, int relu) {
  const int c4 = C >> 2;
  const int Ho = H + 2 * pad_out, Wo = W + 2 * pad_out;
  const int64_t total = static_cast<int64_t>(S) * Ho * Wo * c4;
  const int64_t idx = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (idx >= total) return;
  const int cg = static_cast<int>(idx % c4);
  const int64_t r = idx / c4;
  const int xo = static_cast<int>(r % Wo) - pad_out;
  const int64_t r2 = r / Wo;
  const int yo = static_cast<int>(r2 % Ho) - pad_out;
  const int s = static_cast<int>(r2 / Ho);
  float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
  if (yo >= 0 && yo < H && xo >= 0 && xo < W) {
    const int64_t pa = (static_cast<int64_t>(s) * (H + 2 * pad_a) + yo + pad_a) * (W + 2 * pad_a) + xo + pad_a;
    v = __ldg(reinterpret_cast<const float4*>(a + pa * C + cg * 4));
    if (b != nullptr) {
      const int64_t pb = (static_cast<int64_t>(s) * (H + 2 * pad_b) + yo + pad_b) * (W + 2 * pad_b) + xo + pad_b;
      const float4 w = __ldg(reinterpret_cast<const float4*>(b + pb * C + cg * 4));
      v = make_float4(v.x + w.x, v.y + w.y, v.z + w.z, v.w + w.w);
    }
    if (relu) v = make_float4(fmaxf(v.x, 0.f), fmaxf(v.y, 0.f), fmaxf(v.z, 0.f), fmaxf(v.w, 0.f));
  }
  *reinterpret_cast<float4*>(out + r * C + cg * 4) = v;
}

static inline unsigned blocks_for(int64_t n) { return static_cast<unsigned>((n + 255) / 256); }

}  // namespace sr

extern "C" int sr_im2col_split_bf16(const float* x, int64_t ldx, void* out, int64_t N, int64_t H, int64_t W, int64_t C, int32_t k,
                                    int32_t stride, int32_t pad, int32_t relu, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(x && out, "sr_im2col_split_bf16: null pointer");
  SR_REQUIRE(C > 0 && C % 8 == 0 && ldx >= C && ldx % 4 == 0 && (k == 1 || k == 3) && stride >= 1 && pad >= 0,
             "sr_im2col_split_bf16: C must be a multiple of 8, kernel 1 or 3");
  SR_REQUIRE(((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(out)) & 15) == 0, "sr_im2col_split_bf16: 16-byte alignment");
  const int64_t Ho = (H + 2 * pad - k) / stride + 1, Wo = (W + 2 * pad - k) / stride + 1;
  const int64_t total = N * Ho * Wo * k * k * (C / 8);
  if (total <= 0) return SR_OK;
  SR_REQUIRE(blocks_for(total) < (1u << 31) && (total + 255) / 256 < (1ll << 31), "sr_im2col_split_bf16: too many elements");
  im2col_split_kernel<<<blocks_for(total), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      x, ldx, reinterpret_cast<__nv_bfloat16*>(out), static_cast<int>(N), static_cast<int>(H), static_cast<int>(W),
      static_cast<int>(C), static_cast<int>(Ho), static_cast<int>(Wo), k, stride, pad, relu);
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

extern "C" int sr_split_pad_bf16(const float* x, int64_t ldx, int32_t in_pad, void* out, int64_t S, int64_t H, int64_t W, int64_t C,
                                 int32_t relu, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(x && out && C > 0 && C % 8 == 0 && ldx >= C && ldx % 4 == 0 && S > 0 && H > 0 && W > 0, "sr_split_pad_bf16: bad argument");
  SR_REQUIRE(((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(out)) & 15) == 0, "sr_split_pad_bf16: 16-byte alignment");
  const int64_t total = S * (H + 2) * (W + 2) * (C / 8);
  SR_REQUIRE((total + 255) / 256 < (1ll << 31), "sr_split_pad_bf16: too many elements");
  split_pad_kernel<<<blocks_for(total), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      x, ldx, in_pad ? 1 : 0, reinterpret_cast<__nv_bfloat16*>(out), static_cast<int>(S), static_cast<int>(H), static_cast<int>(W),
      static_cast<int>(C), relu);
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

extern "C" int sr_map_add_f32(const float* a, int32_t pad_a, const float* b, int32_t pad_b, float* out, int32_t pad_out, int64_t S,
                              int64_t H, int64_t W, int64_t C, int32_t relu, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(a && out && C > 0 && C % 4 == 0 && S > 0 && H > 0 && W > 0, "sr_map_add_f32: bad argument");
  SR_REQUIRE(((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b) | reinterpret_cast<uintptr_t>(out)) & 15) == 0,
             "sr_map_add_f32: 16-byte alignment");
  const int po = pad_out ? 1 : 0;
  const int64_t total = S * (H + 2 * po) * (W + 2 * po) * (C / 4);
  SR_REQUIRE((total + 255) / 256 < (1ll << 31), "sr_map_add_f32: too many elements");
  map_add_kernel<<<blocks_for(total), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      a, pad_a ? 1 : 0, b, pad_b ? 1 : 0, out, po, static_cast<int>(S), static_cast<int>(H), static_cast<int>(W), static_cast<int>(C), relu);
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

extern "C" int sr_bilinear_nhwc_f32(const float* x, float* out, int64_t N, int64_t Hi, int64_t Wi, int64_t Ho, int64_t Wo, int64_t C,
                                    int32_t in_pad, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(x && out && C > 0 && C % 4 == 0 && Hi > 0 && Wi > 0 && Ho > 0 && Wo > 0, "sr_bilinear_nhwc_f32: bad argument");
  const int64_t total = N * Ho * Wo * (C / 4);
  if (total <= 0) return SR_OK;
  SR_REQUIRE((total + 255) / 256 < (1ll << 31), "sr_bilinear_nhwc_f32: too many elements");
  const float sy = Ho > 1 ? static_cast<float>(Hi - 1) / static_cast<float>(Ho - 1) : 0.f;
  const float sx = Wo > 1 ? static_cast<float>(Wi - 1) / static_cast<float>(Wo - 1) : 0.f;
  bilinear_nhwc_kernel<<<blocks_for(total), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      x, out, static_cast<int>(N), static_cast<int>(Hi), static_cast<int>(Wi), static_cast<int>(Ho), static_cast<int>(Wo),
      static_cast<int>(C), sy, sx, in_pad ? 1 : 0);
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

extern "C" int sr_pixel_shuffle_f32(const float* g, const float* bias, float* out, int64_t N, int64_t H, int64_t W, int64_t C,
                                    int32_t k, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(g && out && C > 0 && C % 4 == 0 && k >= 1, "sr_pixel_shuffle_f32: bad argument");
  const int64_t total = N * H * W * k * k * (C / 4);
  if (total <= 0) return SR_OK;
  SR_REQUIRE((total + 255) / 256 < (1ll << 31), "sr_pixel_shuffle_f32: too many elements");
  pixel_shuffle_kernel<<<blocks_for(total), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      g, bias, out, static_cast<int>(N), static_cast<int>(H), static_cast<int>(W), static_cast<int>(C), k);
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

extern "C" int sr_add_bcast_f32(const float* a, const float* b, float* out, int64_t n, int64_t period, float alpha, int32_t relu_out,
                                void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(a && out && n >= 0 && period > 0 && n % 4 == 0 && period % 4 == 0, "sr_add_bcast_f32: n and period must be multiples of 4");
  SR_REQUIRE(((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b) | reinterpret_cast<uintptr_t>(out)) & 15) == 0,
             "sr_add_bcast_f32: 16-byte alignment");
  if (n == 0) return SR_OK;
  SR_REQUIRE((n / 4 + 255) / 256 < (1ll << 31), "sr_add_bcast_f32: too many elements");
  add_bcast_kernel<<<blocks_for(n / 4), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(a, b, out, n / 4, period / 4, alpha, relu_out);
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

extern "C" int sr_uv_posembed_table(float* out, int64_t H, int64_t W, int64_t C, double aspect_ratio, double omega0, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(out && H > 0 && W > 0 && C > 0 && C % 4 == 0, "sr_uv_posembed_table: C must be a multiple of 4");
  const int64_t total = H * W * C;
  SR_REQUIRE((total + 255) / 256 < (1ll << 31), "sr_uv_posembed_table: too many elements");
  uv_posembed_kernel<<<blocks_for(total), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      out, static_cast<int>(H), static_cast<int>(W), static_cast<int>(C), aspect_ratio, omega0);
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

extern "C" int sr_dpt_out_head_f32(const float* h, int64_t ldh, const float* w, const float* bias, float* pred, float* conf,
                                   int64_t px, int32_t hid, int32_t out_dim, int32_t act, int32_t conf_act, int32_t map_h,
                                   int32_t map_w, int32_t pad, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(h && w && bias && pred && conf, "sr_dpt_out_head_f32: null pointer");
  SR_REQUIRE(hid > 0 && hid % 4 == 0 && ldh >= hid && ldh % 4 == 0 && (out_dim == 2 || out_dim == 4),
             "sr_dpt_out_head_f32: hid must be a multiple of 4 and out_dim 2 (depth + conf) or 4 (xyz + conf)");
  SR_REQUIRE(act >= 0 && act <= 4 && conf_act >= 0 && conf_act <= 2, "sr_dpt_out_head_f32: unknown activation code");
  SR_REQUIRE(!pad || (map_h > 0 && map_w > 0 && px % (static_cast<int64_t>(map_h) * map_w) == 0), "sr_dpt_out_head_f32: bad padded map");
  if (px == 0) return SR_OK;
  SR_REQUIRE((px + 255) / 256 < (1ll << 31), "sr_dpt_out_head_f32: too many pixels");
  const size_t smem = static_cast<size_t>(out_dim) * (hid + 1) * sizeof(float);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (out_dim == 4) dpt_out_kernel<4><<<blocks_for(px), 256, smem, st>>>(h, ldh, w, bias, pred, conf, px, hid, act, conf_act, map_h, map_w, pad);
  else dpt_out_kernel<2><<<blocks_for(px), 256, smem, st>>>(h, ldh, w, bias, pred, conf, px, hid, act, conf_act, map_h, map_w, pad);
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}
