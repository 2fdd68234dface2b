// Probe (GPU box): does a K-major SWIZZLE_128B tcgen05 shared-memory descriptor accept a start address that is
// row-shifted (start = 1024-aligned base + shift*128 B), and which base_offset value does it need?
// A lives in smem as rows of 64 bf16 (128 B) in the TMA/absolute-address swizzle: 16-B chunk c of row r sits at
// r*128 + ((c ^ (r & 7)) << 4). D[m][n] = sum_k A[m+shift][k] * B[n][k], M=128, N=16, K=64.
#include <cstdio>
#include <cstdlib>
#include <cuda_bf16.h>
#include "../../pdac_pathological_image_segmentation_b200/csrc/ptx.cuh"
using namespace pdac;

__device__ __forceinline__ uint32_t swz(uint32_t a, int bits) { return a ^ (((a >> 7) & ((1u << bits) - 1)) << 4); }
__global__ void probe(const __nv_bfloat16* A, const __nv_bfloat16* B, float* D, int shift, int base_off_mode, int rows_a, int pitch) {
  extern __shared__ uint8_t raw[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t slot;
  uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  uint8_t* sa = smem;                 // rows_a x 128 B
  uint8_t* sb = smem + 32 * 1024;     // 16 x 128 B
  const int cpr = pitch / 16, bits = pitch == 128 ? 3 : (pitch == 64 ? 2 : 1), K = pitch / 2;
  const uint32_t lt = pitch == 128 ? 2 : (pitch == 64 ? 4 : 6);
  for (int i = threadIdx.x; i < rows_a * cpr; i += blockDim.x) {
    int r = i / cpr, c = i % cpr;
    *(uint4*)(sa + swz(r * pitch + c * 16, bits)) = *(const uint4*)(A + r * 64 + c * 8);
  }
  for (int i = threadIdx.x; i < 16 * cpr; i += blockDim.x) {
    int r = i / cpr, c = i % cpr;
    *(uint4*)(sb + swz(r * pitch + c * 16, bits)) = *(const uint4*)(B + r * 64 + c * 8);
  }
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
  if (threadIdx.x < 32) { tmem_alloc(&slot, 32); tmem_relinquish(); }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tb = slot;
  if (threadIdx.x == 0) {
    uint32_t idesc = make_idesc_bf16(128, 16, 0, 0);
    uint32_t a0 = smem_u32(sa) + shift * pitch;
    for (int k = 0; k < K / 16; ++k) {
      uint64_t da = make_smem_desc(a0 + k * 32, 16, 8 * pitch, lt);
      uint64_t bo = base_off_mode == 0 ? 0 : (uint64_t)(((a0) >> 7) & 7);
      da |= bo << 49;
      uint64_t db = make_smem_desc(smem_u32(sb) + k * 32, 16, 8 * pitch, lt);
      umma_bf16(tb, da, db, idesc, k > 0);
    }
    umma_commit(&bar);
  }
  mbar_wait(&bar, 0, nullptr);
  tc_fence_after();
  if (threadIdx.x < 128) {
    uint32_t v[16];
    int q = threadIdx.x >> 5;
    tmem_ld16(tb + ((uint32_t)(q * 32) << 16), v);
    tmem_ld_wait();
    for (int i = 0; i < 16; ++i) D[threadIdx.x * 16 + i] = __uint_as_float(v[i]);
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) { tc_fence_after(); tmem_dealloc(tb, 32); }
}

int main() {
  const int RA = 176;
  __nv_bfloat16 *hA = new __nv_bfloat16[RA * 64], *hB = new __nv_bfloat16[16 * 64];
  float *fa = new float[RA * 64], *fb = new float[16 * 64];
  srand(1);
  for (int i = 0; i < RA * 64; ++i) { fa[i] = (rand() % 17 - 8) / 8.f; hA[i] = __float2bfloat16(fa[i]); }
  for (int i = 0; i < 16 * 64; ++i) { fb[i] = (rand() % 13 - 6) / 4.f; hB[i] = __float2bfloat16(fb[i]); }
  __nv_bfloat16 *dA, *dB; float* dD;
  cudaMalloc(&dA, RA * 64 * 2); cudaMalloc(&dB, 16 * 64 * 2); cudaMalloc(&dD, 128 * 16 * 4);
  cudaMemcpy(dA, hA, RA * 64 * 2, cudaMemcpyHostToDevice); cudaMemcpy(dB, hB, 16 * 64 * 2, cudaMemcpyHostToDevice);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 48 * 1024);
  float hD[128 * 16];
  for (int pitch = 128; pitch >= 32; pitch /= 2)
  for (int mode = 0; mode < 1; ++mode)
    for (int shift = 0; shift <= 35; ++shift) {
      cudaMemset(dD, 0, sizeof(hD));
      probe<<<1, 128, 40 * 1024>>>(dA, dB, dD, shift, mode, RA, pitch);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("mode %d shift %d: CUDA error %s\n", mode, shift, cudaGetErrorString(e)); return 1; }
      cudaMemcpy(hD, dD, sizeof(hD), cudaMemcpyDeviceToHost);
      double maxerr = 0;
      for (int m = 0; m < 128; ++m)
        for (int n = 0; n < 16; ++n) {
          double ref = 0;
          for (int k = 0; k < pitch / 2; ++k) ref += (double)fa[(m + shift) * 64 + k] * fb[n * 64 + k];
          double d = fabs(ref - hD[m * 16 + n]);
          if (d > maxerr) maxerr = d;
        }
      if (maxerr >= 1e-3 || shift == 35) printf("pitch %3d base_offset_mode %d shift %2d : max err %.4f %s\n", pitch, mode, shift, maxerr, maxerr < 1e-3 ? "OK (all shifts up to here)" : "WRONG");
    }
  return 0;
}
