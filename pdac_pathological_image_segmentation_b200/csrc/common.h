// Host-side plumbing shared by every translation unit of libpdac_b200.so: thread-local error string, CUDA error
// mapping, TMA tensor-map encoding through the driver entry point (no link-time dependency on libcuda, so the
// library loads on a GPU-less build box).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/pdac_b200.h"

namespace pdac {

typedef __nv_bfloat16 bf16;

int fail(int code, const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what);

#define PDAC_REQUIRE(cond, ...)                                \
  do {                                                         \
    if (!(cond)) return ::pdac::fail(PDAC_E_INVALID, __VA_ARGS__); \
  } while (0)
#define PDAC_CUDA(expr)                                        \
  do {                                                         \
    cudaError_t e__ = (expr);                                  \
    if (e__ != cudaSuccess) return ::pdac::cuda_fail(e__, #expr); \
  } while (0)
#define PDAC_LAUNCH_CHECK(name) PDAC_CUDA(cudaPeekAtLastError())
#define PDAC_TRY(expr)          \
  do {                          \
    int rc__ = (expr);          \
    if (rc__ != PDAC_OK) return rc__; \
  } while (0)

// Encodes a tiled bf16 tensor map. dims/box are innermost-first; strides_bytes has rank-1 entries (dims 1..).
// swizzle_bytes in {0, 32, 64, 128}.
int encode_tiled_bf16(CUtensorMap* m, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                      const uint32_t* box, int swizzle_bytes);

// Tensor map over an NHWC bf16 view with box (ck, tw, th, tn).
int encode_view_map(CUtensorMap* m, const PdacView* v, int ck, int tw, int th, int tn);

int* device_err_flag();  // device pointer to the pipeline error flag

// Which convolution kernel the last pdac_conv* / pdac_tapconv* call of this thread dispatched to (bench attribution).
enum { KID_NONE = 0, KID_CONV_TC = 1, KID_CONV_HALO = 2, KID_WGRAD_TC = 3, KID_WGRAD_HALO = 4 };
void note_kernel(int id);

// Per-channel epilogue scale of the next tap convolution launched by this thread (set by the *_affine entry points:
// eval-mode BatchNorm folded into the GEMM epilogue as y = act(acc * scale + shift + addend)).
void set_epi_scale(const float* scale);
const float* epi_scale();

inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
inline int pow2_ceil(int x) {
  int p = 1;
  while (p < x) p <<= 1;
  return p;
}
inline int ilog2(int x) {
  int l = 0;
  while ((1 << l) < x) ++l;
  return l;
}
inline int num_sms() {
  static int n = 0;
  if (!n) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

}  // namespace pdac
