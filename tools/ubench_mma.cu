// tcgen05.mma issue / execution rates for the operand modes the attention kernel uses (B200, one CTA on one SM).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/ubench_mma.bin tools/ubench_mma.cu
// One thread issues `iters` batches of `per` MMAs (K = 16 each) on fixed shared-memory / TMEM operands, commits, waits.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../self-supervise-sfm_b200/csrc/common.cuh"
using namespace sr;

// mode 0: SS  A K-major smem, B K-major smem      (S = Q K^T)
// mode 1: SS  A K-major smem, B MN-major smem     (O = P V, P in smem)
// mode 2: TS  A TMEM,         B MN-major smem     (O = P V, P in TMEM)  <- what attn_tc.cu does
// mode 3: TS  A TMEM,         B K-major smem      (O = P V^T-layout)
template <int MODE, int N, int NACC = 1>
__global__ void k_mma(long long* out, int iters, int per) {
  extern __shared__ uint8_t raw[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t slot;
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 64 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
  if (warp == 0) tmem_alloc<512>(&slot);
  fence_proxy_async_smem();
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tm = slot;
  if (__shfl_sync(0xffffffffu, threadIdx.x >> 5, 0) == 0) {   // whole warp, uniform control flow; one elected lane issues
    const uint32_t a_addr = smem_u32(smem), b_addr = smem_u32(smem + 32768);
    constexpr uint32_t idesc = umma_idesc_bf16(128, N, 0, (MODE == 1 || MODE == 2) ? 1 : 0);
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      if (elect_one()) {
        for (int k = 0; k < per; ++k) {
          const uint64_t da = umma_desc_sw128(a_addr + (k & 3) * 32, 16, 1024);
          const uint64_t db = (MODE == 1 || MODE == 2) ? umma_desc_sw128(b_addr + (k & 7) * 2048, 0, 1024)
                                                       : umma_desc_sw128(b_addr + (k & 3) * 32, 16, 1024);
          const uint32_t td = tm + (k % NACC) * 64;     // NACC independent accumulators (64 columns apart)
          if (MODE <= 1) umma_ss(td, da, db, idesc, 1u);
          else           umma_ts(td, tm + 256 + (k & 7) * 8, db, idesc, 1u);
        }
        umma_commit(&bar);
      }
      __syncwarp();
      mbar_wait(&bar, it & 1);
    }
    long long t1 = clock64();
    if (threadIdx.x == 0) out[0] = t1 - t0;
  }
  tc_fence_before(); __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc<512>(tm); }
}

// The attention kernel's issue pattern: per K/V tile and query tile 4 x SS(N=128) + 2 commits, 8 x TS(N=64) + 2 commits,
// issued without waiting (only the tensor pipe throttles).  NISS = 1 or 2 issuer threads (warps 1 and 3) on separate accumulators.
template <int NISS, int NCOMMIT, bool KMAJOR_V>
__global__ void k_pattern(long long* out, int iters) {
  extern __shared__ uint8_t raw[];
  __shared__ __align__(8) uint64_t dummy[2][4], done[2];
  __shared__ uint32_t slot;
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 96 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) {
    for (int q = 0; q < 2; ++q) { for (int i = 0; i < 4; ++i) mbar_init(&dummy[q][i], 1); mbar_init(&done[q], 1); }
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc<512>(&slot);
  fence_proxy_async_smem();
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tm = slot;
  const int q = warp >> 1;
  if ((warp == 1 || (warp == 3 && NISS == 2)) && (threadIdx.x & 31) == 0) {
    constexpr uint32_t idesc_s = umma_idesc_bf16(128, 128, 0, 0);
    constexpr uint32_t idesc_o = umma_idesc_bf16(128, 64, 0, KMAJOR_V ? 0 : 1);
    const uint32_t qa = smem_u32(smem + q * 16384), ka = smem_u32(smem + 32768), va = smem_u32(smem + 65536);
    const uint32_t tS = tm + q * 128, tO = tm + 256 + q * 64, tP = tm + 384 + q * 64;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      for (int k = 0; k < 4; ++k)
        umma_ss(tS, umma_desc_sw128(qa + k * 32, 16, 1024), umma_desc_sw128(ka + (it & 1) * 16384 + k * 32, 16, 1024), idesc_s, k ? 1u : 0u);
      for (int c = 0; c < NCOMMIT; ++c) umma_commit(&dummy[q][c]);
      for (int k = 0; k < 8; ++k) {
        const uint64_t db = KMAJOR_V ? umma_desc_sw128(va + (k & 3) * 32 + (k >> 2) * 8192, 16, 1024)
                                     : umma_desc_sw128(va + k * 2048, 0, 1024);
        umma_ts(tO, tP + k * 8, db, idesc_o, 1u);
      }
      for (int c = 0; c < NCOMMIT; ++c) umma_commit(&dummy[q][2 + c]);
    }
    umma_commit(&done[q]);
    mbar_wait(&done[q], 0);
    long long t1 = clock64();
    out[q] = t1 - t0;
  }
  tc_fence_before(); __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc<512>(tm); }
}

template <int NISS, int NCOMMIT, bool KMAJOR_V>
void run_pattern(const char* name, long long* d) {
  const int iters = 400;
  cudaFuncSetAttribute(k_pattern<NISS, NCOMMIT, KMAJOR_V>, cudaFuncAttributeMaxDynamicSharedMemorySize, 110 * 1024);
  cudaMemset(d, 0, 64);
  k_pattern<NISS, NCOMMIT, KMAJOR_V><<<1, 128, 110 * 1024>>>(d, iters);
  long long h[2] = {0, 0};
  cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
  printf("%-60s : %7.1f clk per tile per issuer (q0), %7.1f (q1); floor 4x64 + 8x48..62 = 640..750   %s\n", name, (double)h[0] / iters,
         (double)h[1] / iters, cudaGetErrorString(cudaGetLastError()));
}

template <int MODE, int N, int NACC = 1>
void run(const char* name, long long* d, int per) {
  const int iters = 200;
  cudaFuncSetAttribute(k_mma<MODE, N, NACC>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  k_mma<MODE, N, NACC><<<1, 128, 100 * 1024>>>(d, iters, per);
  long long h = 0;
  cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
  const double per_mma = (double)h / iters / per;
  printf("%-44s N=%3d batch=%2d : %7.1f clk per MMA (ideal %d) %s\n", name, N, per, per_mma, 128 * N * 16 / 4096,
         cudaGetErrorString(cudaGetLastError()));
}

int main() {
  long long* d;
  cudaMalloc(&d, 64);
  for (int per : {4, 8, 32}) {
    run<0, 128>("SS  A K-major, B K-major  (S = Q K^T)", d, per);
    run<0, 64>("SS  A K-major, B K-major", d, per);
    run<1, 64>("SS  A K-major, B MN-major (P smem, V)", d, per);
    run<2, 64>("TS  A TMEM,    B MN-major (P TMEM, V)", d, per);
    run<3, 64>("TS  A TMEM,    B K-major  (P TMEM, V^T)", d, per);
    run<2, 128>("TS  A TMEM,    B MN-major", d, per);
    run<3, 128>("TS  A TMEM,    B K-major", d, per);
    run<0, 256>("SS  A K-major, B K-major  (GEMM)", d, per);
  }
  printf("--- independent accumulators (is the N=64 floor an accumulate dependency?)\n");
  run<2, 64, 2>("TS N=64 MN-major B, 2 accumulators", d, 32);
  run<2, 64, 4>("TS N=64 MN-major B, 4 accumulators", d, 32);
  run<3, 64, 2>("TS N=64 K-major B, 2 accumulators", d, 32);
  run<0, 64, 2>("SS N=64 K-major, 2 accumulators", d, 32);
  run<0, 64, 4>("SS N=64 K-major, 4 accumulators", d, 32);
  run<0, 32, 4>("SS N=32 K-major, 4 accumulators", d, 32);
  run<0, 128, 2>("SS N=128 K-major, 2 accumulators", d, 32);
  printf("--- attention issue pattern\n");
  run_pattern<1, 0, false>("1 issuer, no commits, V MN-major", d);
  run_pattern<1, 2, false>("1 issuer, 2+2 commits per tile, V MN-major", d);
  run_pattern<2, 0, false>("2 issuers, no commits, V MN-major", d);
  run_pattern<2, 2, false>("2 issuers, 2+2 commits per tile, V MN-major (= attn_tc.cu)", d);
  run_pattern<2, 1, false>("2 issuers, 1+1 commits per tile, V MN-major", d);
  run_pattern<2, 2, true>("2 issuers, 2+2 commits per tile, V K-major (V^T)", d);
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
