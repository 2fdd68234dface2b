// Probe (GPU box): is a tcgen05.mma stream issue-bound or execution-bound? One CTA per SM; `nw` warps (1 or 2) each issue
// `iters` x 16 MMAs (M = 64 or 128, K = 16, SS operands, SW128 K-major) back to back from one elected lane to their own
// accumulators. M = 64 halves the execution time, so its cycles/MMA bound the issue cost from above.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_issue mma_issue.cu && ./mma_issue
#include <cstdio>
#include <cstdlib>
#include <cuda_bf16.h>
#include "../../pdac_pathological_image_segmentation_b200/csrc/ptx.cuh"
using namespace pdac;

__global__ void __launch_bounds__(128, 1) issue(int M, int N, int nw, int iters, long long* out) {
  extern __shared__ uint8_t raw[];
  __shared__ __align__(8) uint64_t bar[2];
  __shared__ uint32_t slot;
  uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  for (int i = threadIdx.x; i < 160 * 1024 / 4; i += blockDim.x) ((uint32_t*)smem)[i] = 0;
  if (threadIdx.x == 0) { mbar_init(&bar[0], 1); mbar_init(&bar[1], 1); fence_barrier_init(); }
  if (threadIdx.x < 32) { tmem_alloc(&slot, 512); tmem_relinquish(); }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tb = slot;
  const int warp = threadIdx.x >> 5;
  if (warp < nw) {
    const uint32_t leader = elect_one();
    const uint32_t idesc = make_idesc_bf16(M, N, 0, 0);
    const uint32_t hi = (uint32_t)(make_smem_desc(0, 16, 8 * 128, 2) >> 32);
    const uint32_t a_lo0 = (((smem_u32(smem) + warp * 32768) >> 4) & 0x3FFF) | (1u << 16);
    const uint32_t b_lo0 = (((smem_u32(smem) + 96 * 1024) >> 4) & 0x3FFF) | (1u << 16);
    const uint32_t tacc = tb + warp * 256;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      if (leader) {
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)
            umma_bf16_lohi(tacc + (j & 1) * 128, a_lo0 + j * 1024 + 2 * kk, hi, b_lo0 + 2 * kk, hi, idesc, 1u);
      }
    }
    long long t_issue = clock64() - t0;
    if (leader) umma_commit(&bar[warp]);
    __syncwarp();
    mbar_wait(&bar[warp], 0, nullptr);
    long long t1 = clock64();
    if ((threadIdx.x & 31) == 0 && blockIdx.x == 0) { out[warp * 2] = t1 - t0; out[warp * 2 + 1] = t_issue; }
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) { tc_fence_after(); tmem_dealloc(tb, 512); }
}

int main() {
  long long* d;
  cudaMalloc(&d, 64);
  cudaFuncSetAttribute(issue, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  for (int M : {64, 128})
    for (int N : {16, 64, 128})
      for (int nw : {1, 2}) {
        const int iters = 200;
        cudaMemset(d, 0, 64);
        issue<<<148, 128, 200 * 1024>>>(M, N, nw, iters, d);
        cudaError_t e = cudaDeviceSynchronize();
        long long h[4] = {0, 0, 0, 0};
        cudaMemcpy(h, d, 32, cudaMemcpyDeviceToHost);
        printf("M %3d N %3d warps %d: warp0 %6.1f cycles/MMA to completion, %6.1f to issue", M, N, nw, (double)h[0] / (iters * 16),
               (double)h[1] / (iters * 16));
        if (nw == 2) printf(" | warp1 %6.1f / %6.1f", (double)h[2] / (iters * 16), (double)h[3] / (iters * 16));
        printf("  (%s)\n", cudaGetErrorString(e));
      }
  return 0;
}
