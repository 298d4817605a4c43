// Micro-benchmarks that size the attention softmax pipeline on B200 (sm_100a): TMEM load bandwidth, MUFU.EX2 issue rate
// per warp / per SM sub-partition, mbarrier try_wait latency.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3
// -o tools/ubench.bin tools/ubench.cu ; run on the GPU box.  Prints cycles per operation (clock64 on one SM).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../self-supervise-sfm_b200/csrc/common.cuh"
using namespace sr;

__global__ void k_ldtm(long long* out, int iters, int nwarps) {
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) tmem_alloc<512>(&slot);
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t base = slot + ((uint32_t)((warp & 3) * 32) << 16);
  uint32_t acc = 0;
  __syncthreads();
  long long t0 = clock64();
  if (warp < nwarps) {
    for (int i = 0; i < iters; ++i) {
      uint32_t a[32], b[32], c[32], d[32];
      tmem_ld32(base + ((warp >> 2) * 128), a);
      tmem_ld32(base + ((warp >> 2) * 128) + 32, b);
      tmem_ld32(base + ((warp >> 2) * 128) + 64, c);
      tmem_ld32(base + ((warp >> 2) * 128) + 96, d);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 32; ++j) acc += a[j] ^ b[j] ^ c[j] ^ d[j];
    }
  }
  long long t1 = clock64();
  if (acc == 0x12345) out[1] = acc;
  __syncthreads();
  if (threadIdx.x == 0) out[0] = t1 - t0;
  tc_fence_before(); __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc<512>(slot); }
}

__global__ void k_sttm(long long* out, int iters, int nwarps) {
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) tmem_alloc<512>(&slot);
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t base = slot + ((uint32_t)((warp & 3) * 32) << 16);
  uint32_t a[32];
  for (int j = 0; j < 32; ++j) a[j] = threadIdx.x + j;
  __syncthreads();
  long long t0 = clock64();
  if (warp < nwarps) {
    for (int i = 0; i < iters; ++i) {
      tmem_st32(base + ((warp >> 2) * 128), a);
      tmem_st32(base + ((warp >> 2) * 128) + 32, a);
      tmem_st_wait();
    }
  }
  long long t1 = clock64();
  __syncthreads();
  if (threadIdx.x == 0) out[0] = t1 - t0;
  tc_fence_before(); __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc<512>(slot); }
}

template <int ILP>
__global__ void k_mufu(long long* out, float* sink, int iters, int nwarps) {
  const int warp = threadIdx.x >> 5;
  float x[ILP];
  for (int j = 0; j < ILP; ++j) x[j] = -0.001f * (threadIdx.x + j);
  __syncthreads();
  long long t0 = clock64();
  if (warp < nwarps) {
    for (int i = 0; i < iters; ++i) {
#pragma unroll
      for (int j = 0; j < ILP; ++j) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[j]));
    }
  }
  long long t1 = clock64();
  float s = 0;
  for (int j = 0; j < ILP; ++j) s += x[j];
  if (s == 123.f) sink[0] = s;
  __syncthreads();
  if (threadIdx.x == 0) out[0] = t1 - t0;
}

// mixed: per 2 elements: FFMA2, 2 MUFU, FADD2, F2FP  (the softmax inner loop) on nwarps warps
__global__ void k_softmax_mix(long long* out, float* sink, int iters, int nwarps) {
  const int warp = threadIdx.x >> 5;
  float x[16];
  for (int j = 0; j < 16; ++j) x[j] = -0.001f * (threadIdx.x + j);
  float sum0 = 0, sum1 = 0; uint32_t pk = 0;
  __syncthreads();
  long long t0 = clock64();
  if (warp < nwarps) {
    for (int i = 0; i < iters; ++i) {
#pragma unroll
      for (int j = 0; j < 16; j += 2) {
        float a = fmaf(x[j], 0.99f, -0.01f), b = fmaf(x[j + 1], 0.99f, -0.01f);
        float ea, eb;
        asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(ea) : "f"(a));
        asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(eb) : "f"(b));
        sum0 += ea; sum1 += eb;
        uint32_t p;
        asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(p) : "f"(eb), "f"(ea));
        pk ^= p;
        x[j] = a; x[j + 1] = b;
      }
    }
  }
  long long t1 = clock64();
  if (sum0 + sum1 == 123.f || pk == 77) sink[0] = sum0;
  __syncthreads();
  if (threadIdx.x == 0) out[0] = t1 - t0;
}

// Do TMEM loads and MUFU ops of two warps on the SAME sub-partition overlap or serialise?  warp 0: LDTM loop, warp 4: MUFU
// loop (both on sub-partition 0); mode 1 = LDTM only, 2 = MUFU only, 3 = both; also warp 8 = mbarrier arrive+try_wait (mode & 4).
__global__ void k_mio(long long* out, float* sink, int iters, int mode) {
  __shared__ uint32_t slot;
  __shared__ __align__(8) uint64_t bar;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) tmem_alloc<512>(&slot);
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t base = slot;
  uint32_t acc = 0;
  float x[8];
  for (int j = 0; j < 8; ++j) x[j] = -0.001f * (threadIdx.x + j);
  __syncthreads();
  long long t0 = clock64();
  if (warp == 0 && (mode & 1)) {
    for (int i = 0; i < iters; ++i) {
      uint32_t a[32], b[32], c[32], d[32];
      tmem_ld32(base, a); tmem_ld32(base + 32, b); tmem_ld32(base + 64, c); tmem_ld32(base + 96, d);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 32; ++j) acc += a[j] ^ b[j] ^ c[j] ^ d[j];
    }
  }
  if (warp == 4 && (mode & 2)) {
    for (int i = 0; i < iters * 4; ++i) {
#pragma unroll
      for (int j = 0; j < 8; ++j) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[j]));
    }
  }
  if (warp == 8 && (mode & 4) && (threadIdx.x & 31) == 0) {
    for (int i = 0; i < iters; ++i) { mbar_arrive(&bar); mbar_wait(&bar, i & 1); }
  }
  long long t1 = clock64();
  float s = 0;
  for (int j = 0; j < 8; ++j) s += x[j];
  if (s == 123.f || acc == 0x12345) sink[0] = s;
  if ((threadIdx.x & 31) == 0) out[warp] = t1 - t0;
  tc_fence_before(); __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc<512>(slot); }
}

__global__ void k_trywait(long long* out, int iters) {
  __shared__ __align__(8) uint64_t bar;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
  __syncthreads();
  if (threadIdx.x == 0) {
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
      mbar_arrive(&bar);
      mbar_wait(&bar, i & 1);
    }
    long long t1 = clock64();
    out[0] = t1 - t0;
  }
}

int main() {
  long long* d; float* sink;
  cudaMalloc(&d, 256); cudaMalloc(&sink, 64);
  long long h;
  const int it = 2000;
  for (int nw : {1, 4, 8}) {
    k_ldtm<<<1, 256>>>(d, it, nw); cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
    printf("ldtm 4x(32x32b.x32)+wait  warps=%d : %.1f clk per iter (16 KB/warp-iter) -> %.1f B/clk/SM\n", nw, (double)h / it, 16384.0 * nw / ((double)h / it));
  }
  for (int nw : {1, 4, 8}) {
    k_sttm<<<1, 256>>>(d, it, nw); cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
    printf("sttm 2x(32x32b.x32)+wait  warps=%d : %.1f clk per iter (8 KB/warp-iter) -> %.1f B/clk/SM\n", nw, (double)h / it, 8192.0 * nw / ((double)h / it));
  }
  for (int nw : {1, 2, 4, 8, 16}) {
    k_mufu<8><<<1, 512>>>(d, sink, it, nw); cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
    printf("mufu.ex2 ILP8 warps=%d : %.2f clk per warp-instr (per warp) ; %.2f lanes/clk/SM\n", nw, (double)h / (it * 8), 32.0 * 8 * it * nw / (double)h);
  }
  for (int nw : {1, 4, 8, 16}) {
    k_softmax_mix<<<1, 512>>>(d, sink, it, nw); cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
    printf("softmax mix (16 elem/iter) warps=%d : %.2f clk per element (per warp) ; %.2f elems/clk/SM\n", nw, (double)h / (it * 16), 32.0 * 16 * it * nw / (double)h);
  }
  for (int mode : {1, 2, 3, 4, 5, 6, 7}) {
    long long hh[12];
    cudaMemset(d, 0, 96);
    k_mio<<<1, 384>>>(d, sink, it, mode); cudaMemcpy(hh, d, 96, cudaMemcpyDeviceToHost);
    printf("same-SMSP mix mode %d (1 LDTM 16KB/iter, 2 MUFU 32 instr/iter, 4 mbarrier arrive+wait): LDTM %.1f  MUFU %.1f  mbar %.1f clk per iter\n",
           mode, (double)hh[0] / it, (double)hh[4] / it, (double)hh[8] / it);
  }
  k_trywait<<<1, 32>>>(d, it); cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
  printf("mbarrier arrive + try_wait (complete) : %.1f clk\n", (double)h / it);
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
