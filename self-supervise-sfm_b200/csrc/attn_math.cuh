// Softmax arithmetic helpers shared by the tcgen05 attention kernels (attn_tc.cu, attn_tc4.cu).
#pragma once
#include "common.cuh"

namespace sr {

// ---- packed fp32 helpers (FFMA2 / FADD2 on sm_100) and bare MUFU.EX2
__device__ __forceinline__ uint64_t pack2(float lo, float hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void unpack2(uint64_t v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
__device__ __forceinline__ uint64_t add2(uint64_t a, uint64_t b) {
  uint64_t r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ float ex2_approx(float x) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
// exp2 of a packed pair on the FMA / ALU pipes instead of the MUFU unit (16 exponentials / clk / SM is the binding
// resource of this kernel): round-to-nearest split x = i + f via the 1.5 * 2^23 magic add, f in [-0.5, 0.5],
// degree-3 minimax polynomial for 2^f (max relative error 7.5e-5, far below the bf16 rounding of P that follows), and
// the integer part added straight into the exponent field.  x is clamped at -126 so the exponent cannot wrap.
__device__ __forceinline__ void ex2_poly2(uint64_t x2, float& e0, float& e1) {
  float x0, x1;
  unpack2(x2, x0, x1);
  x0 = fmaxf(x0, -126.0f);
  x1 = fmaxf(x1, -126.0f);
  const uint64_t xc = pack2(x0, x1);
  const uint64_t t2 = add2(xc, pack2(12582912.0f, 12582912.0f));
  const uint64_t i2 = add2(t2, pack2(-12582912.0f, -12582912.0f));
  const uint64_t f2 = fma2(i2, pack2(-1.0f, -1.0f), xc);
  uint64_t p2 = fma2(pack2(0.0551716685f, 0.0551716685f), f2, pack2(0.2426111251f, 0.2426111251f));
  p2 = fma2(p2, f2, pack2(0.6932609677f, 0.6932609677f));
  p2 = fma2(p2, f2, pack2(0.9999280572f, 0.9999280572f));
  float t0, t1, p0, p1;
  unpack2(t2, t0, t1);
  unpack2(p2, p0, p1);
  e0 = __int_as_float(__float_as_int(p0) + (__float_as_int(t0) << 23));
  e1 = __int_as_float(__float_as_int(p1) + (__float_as_int(t1) << 23));
}
__device__ __forceinline__ void named_bar_arrive(uint32_t id, uint32_t nthreads) {
  asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
__device__ __forceinline__ float max3(float a, float b, float c) {
  float r;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
  return r;
}
__device__ __forceinline__ uint32_t cvt_bf16x2(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}

}  // namespace sr
