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

int* device_err_flag();  // device alias of the current device's pipeline error flag (mapped pinned host memory)

// Which convolution kernel the last pdac_conv* / pdac_tapconv* call of this thread dispatched to (bench attribution).
enum { KID_NONE = 0, KID_CONV_TC = 1, KID_CONV_HALO = 2, KID_WGRAD_TC = 3, KID_WGRAD_HALO = 4 };
void note_kernel(int id);

// Per-channel epilogue scale of the next tap convolution launched by this thread (set by the *_affine entry points:
// eval-mode BatchNorm folded into the GEMM epilogue as y = act(acc * scale + shift + addend)).
void set_epi_scale(const float* scale);
const float* epi_scale();
// Half-resolution destination of the next tap convolution's leading output channels (set by pdac_conv2d_dgrad_pool2:
// the 2x2 sum of a nearest-neighbour x2 upsampling's gradient, formed in the halo kernel's epilogue).
void set_epi_pool(const PdacView* pooled);
const PdacView* epi_pool();
// Full-resolution (2H x 2W) destination of the next tap convolution when its output is to be stored nearest-neighbour
// x2 upsampled (set by pdac_conv2d_fwd_affine_up2; the halo kernel's direct epilogue stores every pixel four times).
void set_epi_up2(const PdacView* full);
const PdacView* epi_up2();

// Deterministic cross-block reductions (no floating-point atomics): out[c] (+)= sum over rows 0..n_rows-1 of
// rows[r][c], added in a fixed order (bn.cu).
int rows_sum(const float* rows, int n_rows, int ncols, float* out, int accumulate, cudaStream_t stream);
// Deterministic split-K of the weight-gradient kernels (conv_tc.cu): partial tensors in a caller scratch buffer, added
// in share order.
int wgrad_plan_partials(float* scratch, long long scratch_bytes, int n_taps, int A, int B, int* splits, float** part,
                        long long* part_stride);
int wgrad_reduce(const float* part, int splits, long long part_stride, int n_taps, int B, int A, float* dw,
                 long long tap_stride, long long a_stride, long long b_stride, void* stream);

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
  static int n[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (!n[dev]) {
    cudaDeviceGetAttribute(&n[dev], cudaDevAttrMultiProcessorCount, dev);
    if (n[dev] <= 0) n[dev] = 148;
  }
  return n[dev];
}

}  // namespace pdac
