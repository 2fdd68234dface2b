// Probe (GPU box): cycles per tcgen05.mma (cta_group::1, kind::f16, M=128, K=16, SS operands) as a function of N, the
// swizzle width of the K-major operands and a row shift of the A start address, with no loads in the way. One CTA per SM
// (grid 148) or a single CTA, `iters` x `unroll` MMAs issued back to back by one elected lane, timed with clock64.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_rate mma_rate.cu && ./mma_rate
#include <cstdio>
#include <cstdlib>
#include <cuda_bf16.h>
#include "../../pdac_pathological_image_segmentation_b200/csrc/ptx.cuh"
using namespace pdac;

__global__ void __launch_bounds__(128, 1) rate(int N, int pitch, int pitch_b, int shift, int nacc, int iters, long long* out) {
  extern __shared__ uint8_t raw[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t slot;
  uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  for (int i = threadIdx.x; i < 160 * 1024 / 4; i += blockDim.x) ((uint32_t*)smem)[i] = 0;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
  if (threadIdx.x < 32) { tmem_alloc(&slot, 512); tmem_relinquish(); }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tb = slot;
  long long t0 = 0, t1 = 0;
  if (threadIdx.x < 32) {
    const uint32_t leader = elect_one();
    const uint32_t lt = pitch == 128 ? 2 : (pitch == 64 ? 4 : 6);
    const uint32_t idesc = make_idesc_bf16(128, N, 0, 0);
    const uint32_t hi = (uint32_t)(make_smem_desc(0, 16, 8 * pitch, lt) >> 32);
    const uint32_t ltb = pitch_b == 128 ? 2 : (pitch_b == 64 ? 4 : 6);
    const uint32_t hib = (uint32_t)(make_smem_desc(0, 16, 8 * pitch_b, ltb) >> 32);
    const uint32_t a_lo0 = (((smem_u32(smem) + shift * pitch) >> 4) & 0x3FFF) | (1u << 16);
    const uint32_t b_lo0 = (((smem_u32(smem) + 96 * 1024) >> 4) & 0x3FFF) | (1u << 16);
    const int ksteps = pitch / 32;
    const uint32_t kb_mask = pitch_b / 32 - 1;
    t0 = clock64();
    if (ksteps == 4) {
      for (int it = 0; it < iters; ++it) {
        if (leader) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            if (j < nacc) {
#pragma unroll
              for (int kk = 0; kk < 4; ++kk)
                umma_bf16_lohi(tb + j * N, a_lo0 + j * 1024 + 2 * kk, hi, b_lo0 + 2 * (kk & kb_mask), hib, idesc, 1u);
            }
          }
        }
      }
    } else if (ksteps == 2) {
      for (int it = 0; it < iters; ++it) {
        if (leader) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            if (j < nacc) {
#pragma unroll
              for (int kk = 0; kk < 2; ++kk)
                umma_bf16_lohi(tb + j * N, a_lo0 + j * 512 + 2 * kk, hi, b_lo0 + 2 * (kk & kb_mask), hib, idesc, 1u);
            }
          }
        }
      }
    } else {
      for (int it = 0; it < iters; ++it) {
        if (leader) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            if (j < nacc) umma_bf16_lohi(tb + j * N, a_lo0 + j * 256, hi, b_lo0, hib, idesc, 1u);
          }
        }
      }
    }
    if (leader) umma_commit(&bar);
    __syncwarp();
    mbar_wait(&bar, 0, nullptr);
    t1 = clock64();
    if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = t1 - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) { tc_fence_after(); tmem_dealloc(tb, 512); }
}

int main() {
  long long* d;
  cudaMalloc(&d, 8);
  cudaFuncSetAttribute(rate, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  const int Ns[] = {16, 32, 64, 128, 256};
  const int pitches[] = {32, 64, 128};
  for (int grid : {148})
    for (int pitch : pitches)
      for (int N : Ns)
        for (int pitch_b : {32, 128})
        for (int shift : {3}) {
          const int nacc = 512 / N < 4 ? 512 / N : 4;
          if (nacc < 1) continue;
          const int iters = 200, ksteps = pitch / 32;
          rate<<<grid, 128, 200 * 1024>>>(N, pitch, pitch_b, shift, nacc, iters, d);
          cudaError_t e = cudaDeviceSynchronize();
          long long h = 0;
          cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
          printf("grid %3d pitchA %3d pitchB %3d N %3d shift %2d nacc %d: %7.1f cycles/MMA (%s)\n", grid, pitch, pitch_b, N, shift, nacc,
                 (double)h / (iters * nacc * ksteps), cudaGetErrorString(e));
        }
  return 0;
}
