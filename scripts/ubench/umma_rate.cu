// Microbenchmark: cycles per tcgen05.mma (kind::f16, K = 16) as a function of N, for M = 128 (one CTA),
// M = 64 (one CTA) and M = 256 (CTA pair, cta_group::2).  One thread issues R back-to-back MMAs on a fixed
// (zero-filled) SW128 K-major smem tile, commits, waits; clock64 around the whole sequence.
//   nvcc -std=c++17 -gencode arch=compute_100a,code=sm_100a -O3 -I vits_b200/csrc -o scripts/ubench/umma_rate scripts/ubench/umma_rate.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include <cuda.h>
#include "ptx.h"
using namespace vitsb;

template <bool TWO>
__global__ void __launch_bounds__(128, 1) rate_kernel(int M, int N, int R, long long* out) {
  extern __shared__ uint8_t raw[];
  __shared__ uint32_t s_tmem;
  __shared__ __align__(8) uint64_t bar;
  const uint32_t base = (ptx::smem_u32(raw) + 1023u) & ~1023u;
  const uint32_t sA = base, sB = base + 16384;            // A: 128 x 64 fp16, B: up to 256 x 64 fp16
  for (uint32_t i = threadIdx.x; i < (16384 + 32768) / 4; i += blockDim.x)
    asm volatile("st.shared.b32 [%0], %1;" ::"r"(base + 4 * i), "r"(0u));
  const int warp = threadIdx.x >> 5;
  const uint32_t bar_a = ptx::smem_u32(&bar);
  if (threadIdx.x == 0) { ptx::mbar_init(bar_a, 1); ptx::mbar_fence_init(); }
  if (warp == 0) { if (TWO) ptx::tmem_alloc_2sm(ptx::smem_u32(&s_tmem), 512); else ptx::tmem_alloc(ptx::smem_u32(&s_tmem), 512); }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  ptx::tc_fence_before();
  if (TWO) ptx::cluster_sync_all(); else __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = s_tmem;
  const int rank = TWO ? (int)ptx::cluster_ctarank() : 0;
  long long t0 = 0, t1 = 0, t2 = 0;
  if (threadIdx.x == 0 && rank == 0) {
    const uint64_t dhi = ptx::umma_smem_desc(0, 128, 0);
    const uint64_t ad = dhi | (uint64_t)((sA & 0x3FFFFu) >> 4), bd = dhi | (uint64_t)((sB & 0x3FFFFu) >> 4);
    const uint32_t idesc = ptx::umma_idesc_f16(0, M, N);
    t0 = clock64();
    for (int r = 0; r < R; r += 4) {
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
        if (TWO) ptx::umma_f16_2sm(tmem, ad + 2 * kk, bd + 2 * kk, idesc, (r | kk) ? 1u : 0u);
        else ptx::umma_f16(tmem, ad + 2 * kk, bd + 2 * kk, idesc, (r | kk) ? 1u : 0u);
      }
    }
    t1 = clock64();
    if (TWO) ptx::umma_commit_2sm(bar_a); else ptx::umma_commit(bar_a);
  }
  if (threadIdx.x == 0) {
    ptx::mbar_wait(bar_a, 0);
    t2 = clock64();
    if (rank == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
  }
  ptx::tc_fence_before();
  if (TWO) ptx::cluster_sync_all(); else __syncthreads();
  if (warp == 0) { ptx::tc_fence_after(); if (TWO) ptx::tmem_dealloc_2sm(tmem, 512); else ptx::tmem_dealloc(tmem, 512); }
}

int main() {
  long long* d; cudaMalloc(&d, 16);
  const int R = 2048;
  const size_t smem = 1024 + 16384 + 32768;
  cudaFuncSetAttribute(rate_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaFuncSetAttribute(rate_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  printf("mode,M,N,issue_cycles_per_mma,total_cycles_per_mma,ideal_cycles_per_sm(rows*N*16/4096)\n");
  for (int mode = 0; mode < 3; ++mode) {
    const int M = mode == 0 ? 128 : (mode == 1 ? 64 : 256);
    for (int N : {16, 32, 64, 96, 128, 192, 256}) {
      if (M == 256 && N % 32) continue;
      if (M == 128 && N % 16) continue;
      long long h[2] = {0, 0};
      for (int rep = 0; rep < 2; ++rep) {
        if (mode == 2) {
          cudaLaunchConfig_t cfg{}; cfg.gridDim = dim3(2); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = smem;
          cudaLaunchAttribute at; at.id = cudaLaunchAttributeClusterDimension; at.val.clusterDim.x = 2; at.val.clusterDim.y = 1; at.val.clusterDim.z = 1;
          cfg.attrs = &at; cfg.numAttrs = 1;
          cudaLaunchKernelEx(&cfg, rate_kernel<true>, M, N, R, d);
        } else {
          rate_kernel<false><<<1, 128, smem>>>(M, N, R, d);
        }
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("error: %s (M=%d N=%d)\n", cudaGetErrorString(e), M, N); return 1; }
        cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
      }
      const int rows_per_sm = M == 256 ? 128 : M;
      printf("%s,%d,%d,%.1f,%.1f,%.1f\n", mode == 2 ? "pair" : "single", M, N, (double)h[0] / R, (double)h[1] / R, rows_per_sm * N * 16 / 4096.0);
    }
  }
  return 0;
}
