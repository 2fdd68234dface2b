#include "common.h"

#include <stdarg.h>
#include <stdio.h>
#include <string.h>

namespace pdac {

static thread_local char g_err[512] = "";

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

int cuda_fail(cudaError_t e, const char* what) {
  return fail(e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver ? PDAC_E_NODEVICE : PDAC_E_CUDA,
              "CUDA error %d (%s) in %s", (int)e, cudaGetErrorString(e), what);
}

static thread_local const float* g_epi_scale = nullptr;
void set_epi_scale(const float* scale) { g_epi_scale = scale; }
const float* epi_scale() { return g_epi_scale; }
static thread_local const PdacView* g_epi_pool = nullptr;
void set_epi_pool(const PdacView* pooled) { g_epi_pool = pooled; }
const PdacView* epi_pool() { return g_epi_pool; }
static thread_local const PdacView* g_epi_up2 = nullptr;
void set_epi_up2(const PdacView* full) { g_epi_up2 = full; }
const PdacView* epi_up2() { return g_epi_up2; }

static thread_local int g_last_kernel = 0;
void note_kernel(int id) { g_last_kernel = id; }
int last_kernel() { return g_last_kernel; }

// Pipeline error flag: one int per device in mapped pinned host memory. A kernel whose mbarrier wait times out stores 1
// through the device alias; the host reads its own alias without synchronising (pdac_peek_device_flag).
static const int kMaxDevices = 64;
static int* g_flag_host[kMaxDevices] = {};
static int* g_flag_dev[kMaxDevices] = {};

static int flag_slot() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return -1;
  if (!g_flag_dev[dev]) {
    int* h = nullptr;
    void* d = nullptr;
    if (cudaHostAlloc((void**)&h, sizeof(int), cudaHostAllocMapped | cudaHostAllocPortable) != cudaSuccess) return -1;
    *h = 0;
    if (cudaHostGetDevicePointer(&d, h, 0) != cudaSuccess) {
      cudaFreeHost(h);
      return -1;
    }
    g_flag_host[dev] = h;
    g_flag_dev[dev] = (int*)d;
  }
  return dev;
}

int* device_err_flag() {
  const int dev = flag_slot();
  return dev < 0 ? nullptr : g_flag_dev[dev];
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  }
  return fn;
}

int encode_tiled_bf16(CUtensorMap* m, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                      const uint32_t* box, int swizzle_bytes) {
  EncodeTiledFn fn = get_encode();
  if (!fn) return fail(PDAC_E_NODEVICE, "cuTensorMapEncodeTiled driver entry point unavailable (no CUDA driver?)");
  cuuint64_t gd[5];
  cuuint64_t gs[4];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; ++i) {
    gd[i] = dims[i];
    bx[i] = box[i];
    es[i] = 1;
  }
  for (int i = 0; i + 1 < rank; ++i) gs[i] = strides_bytes[i];
  CUtensorMapSwizzle sw = swizzle_bytes == 128  ? CU_TENSOR_MAP_SWIZZLE_128B
                          : swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                          : swizzle_bytes == 32 ? CU_TENSOR_MAP_SWIZZLE_32B
                                                : CU_TENSOR_MAP_SWIZZLE_NONE;
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), gd, gs, bx, es,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    char buf[256];
    int n = 0;
    for (int i = 0; i < rank; ++i)
      n += snprintf(buf + n, sizeof(buf) - n, " d%d=%llu/box%u", i, (unsigned long long)dims[i], box[i]);
    for (int i = 0; i + 1 < rank; ++i)
      n += snprintf(buf + n, sizeof(buf) - n, " s%d=%llu", i + 1, (unsigned long long)strides_bytes[i]);
    return fail(PDAC_E_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d (base=%p sw=%d%s)", (int)r, base,
                swizzle_bytes, buf);
  }
  return PDAC_OK;
}

int encode_view_map(CUtensorMap* m, const PdacView* v, int ck, int tw, int th, int tn) {
  if (((uintptr_t)v->ptr & 15) || (v->sW * 2) % 16 || (v->sH * 2) % 16 || (v->sN * 2) % 16)
    return fail(PDAC_E_INVALID, "view not 16-byte aligned for TMA (ptr=%p sW=%lld sH=%lld sN=%lld C=%d)", v->ptr,
                (long long)v->sW, (long long)v->sH, (long long)v->sN, v->C);
  uint64_t dims[4] = {(uint64_t)v->C, (uint64_t)v->W, (uint64_t)v->H, (uint64_t)v->N};
  // a dimension of extent 1 may carry any stride; give TMA a legal one
  uint64_t sW = (uint64_t)v->sW * 2, sH = (uint64_t)v->sH * 2, sN = (uint64_t)v->sN * 2;
  if (sW == 0) sW = 16;
  if (sH == 0) sH = 16;
  if (sN == 0) sN = 16;
  uint64_t strides[3] = {sW, sH, sN};
  uint32_t box[4] = {(uint32_t)ck, (uint32_t)tw, (uint32_t)th, (uint32_t)tn};
  return encode_tiled_bf16(m, v->ptr, 4, dims, strides, box, ck * 2 >= 128 ? 128 : ck * 2);
}

}  // namespace pdac

extern "C" {

int pdac_version(void) { return 100; }

const char* pdac_last_error(void) { return pdac::g_err; }

int pdac_last_kernel_id(void) { return pdac::last_kernel(); }

int pdac_check_device_flag(void* stream) {
  using namespace pdac;
  PDAC_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  return pdac_peek_device_flag();
}

int pdac_peek_device_flag(void) {
  using namespace pdac;
  const int dev = flag_slot();
  if (dev < 0) return fail(PDAC_E_NODEVICE, "no pipeline error flag (no CUDA device?)");
  volatile int* h = g_flag_host[dev];
  const int v = *h;
  if (v) {
    *h = 0;
    return fail(PDAC_E_TIMEOUT, "a tcgen05/TMA pipeline barrier timed out on device %d (flag=%d): results of the "
                                "kernels launched since the last check are invalid", dev, v);
  }
  return PDAC_OK;
}
}
