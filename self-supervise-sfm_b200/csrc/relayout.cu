// Token re-layout kernels (HBM-bound, coalesced float4 per lane).
// Replace the ATen advanced-index / cat kernels of sailrecon/models/aggregator.py:
//   sr_gather_rows_f32  -> random_select_features gather (:617-625) + special-token prepend (:588)
//   sr_copy_cols_f32    -> torch.cat([frame, reloc], dim=-1) halves (:406-423, :568-571)
//   sr_cast_f32_to_bf16 -> the weight cast autocast performs on every Linear call (done once here)
// The frame<->global re-layout itself (:355-357, :393-399) needs no kernel: anchors and queries live in
// one persistent [anchors ; queries] residual buffer (see DESIGN.md "data layout").
#include "common.cuh"
#include "sr_internal.h"

namespace sr {
extern int g_launch_count;

__global__ void __launch_bounds__(256)
gather_rows_kernel(const float* __restrict__ src, int64_t lds, const int32_t* __restrict__ idx, float* __restrict__ dst,
                   int64_t ldd, int64_t n, int nvec) {
  // one warp per destination row
  const int lane = threadIdx.x & 31;
  const int64_t r = static_cast<int64_t>(blockIdx.x) * 8 + (threadIdx.x >> 5);
  if (r >= n) return;
  const int64_t s = __ldg(idx + r);
  const float4* sp = reinterpret_cast<const float4*>(src + s * lds);
  float4* dp = reinterpret_cast<float4*>(dst + r * ldd);
  for (int i = lane; i < nvec; i += 32) dp[i] = __ldg(sp + i);
}

__global__ void __launch_bounds__(256)
copy_cols_kernel(const float* __restrict__ src, int64_t lds, float* __restrict__ dst, int64_t ldd, int64_t rows,
                 int nvec) {
  const int lane = threadIdx.x & 31;
  const int64_t r = static_cast<int64_t>(blockIdx.x) * 8 + (threadIdx.x >> 5);
  if (r >= rows) return;
  const float4* sp = reinterpret_cast<const float4*>(src + r * lds);
  float4* dp = reinterpret_cast<float4*>(dst + r * ldd);
  for (int i = lane; i < nvec; i += 32) dp[i] = __ldg(sp + i);
}

__global__ void __launch_bounds__(256)
cast_bf16_kernel(const float* __restrict__ src, int64_t lds, __nv_bfloat16* __restrict__ dst, int64_t ldd, int64_t rows,
                 int nvec) {
  const int lane = threadIdx.x & 31;
  const int64_t r = static_cast<int64_t>(blockIdx.x) * 8 + (threadIdx.x >> 5);
  if (r >= rows) return;
  const float4* sp = reinterpret_cast<const float4*>(src + r * lds);
  uint2* dp = reinterpret_cast<uint2*>(dst + r * ldd);
  for (int i = lane; i < nvec; i += 32) {
    float4 v = __ldg(sp + i);
    dp[i] = make_uint2(pack_bf16x2(v.x, v.y), pack_bf16x2(v.z, v.w));
  }
}

}  // namespace sr

extern "C" int sr_gather_rows_f32(const float* src, int64_t lds, const int32_t* idx, float* dst, int64_t ldd, int64_t n,
                                  int64_t C, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  if (n == 0) return SR_OK;
  SR_REQUIRE(src && idx && dst, "sr_gather_rows_f32: null pointer");
  SR_REQUIRE(C % 4 == 0 && lds % 4 == 0 && ldd % 4 == 0, "sr_gather_rows_f32: C and row strides must be multiples of 4");
  SR_REQUIRE((reinterpret_cast<uintptr_t>(src) & 15) == 0 && (reinterpret_cast<uintptr_t>(dst) & 15) == 0,
             "sr_gather_rows_f32: pointers must be 16-byte aligned");
  gather_rows_kernel<<<static_cast<unsigned>((n + 7) / 8), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      src, lds, idx, dst, ldd, n, static_cast<int>(C / 4));
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

extern "C" int sr_copy_cols_f32(const float* src, int64_t lds, float* dst, int64_t ldd, int64_t col_off, int64_t rows,
                                int64_t C, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  if (rows == 0) return SR_OK;
  SR_REQUIRE(src && dst, "sr_copy_cols_f32: null pointer");
  SR_REQUIRE(C % 4 == 0 && lds % 4 == 0 && ldd % 4 == 0 && col_off % 4 == 0,
             "sr_copy_cols_f32: C, col_off and row strides must be multiples of 4");
  SR_REQUIRE((reinterpret_cast<uintptr_t>(src) & 15) == 0 && (reinterpret_cast<uintptr_t>(dst) & 15) == 0,
             "sr_copy_cols_f32: pointers must be 16-byte aligned");
  copy_cols_kernel<<<static_cast<unsigned>((rows + 7) / 8), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      src, lds, dst + col_off, ldd, rows, static_cast<int>(C / 4));
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

extern "C" int sr_cast_f32_to_bf16(const float* src, int64_t lds, void* dst, int64_t ldd, int64_t rows, int64_t C,
                                   void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  if (rows == 0) return SR_OK;
  SR_REQUIRE(src && dst, "sr_cast_f32_to_bf16: null pointer");
  SR_REQUIRE(C % 4 == 0 && lds % 4 == 0 && ldd % 4 == 0, "sr_cast_f32_to_bf16: C and row strides must be multiples of 4");
  SR_REQUIRE((reinterpret_cast<uintptr_t>(src) & 15) == 0 && (reinterpret_cast<uintptr_t>(dst) & 7) == 0,
             "sr_cast_f32_to_bf16: pointers must be aligned");
  cast_bf16_kernel<<<static_cast<unsigned>((rows + 7) / 8), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      src, lds, reinterpret_cast<__nv_bfloat16*>(dst), ldd, rows, static_cast<int>(C / 4));
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}
