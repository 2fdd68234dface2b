// Loss / metric / mask reductions and the slide stitcher.
//   DiceLoss.forward            util.py:10-19     1 - (2*sum(p*t) + s) / (sum(p) + sum(t) + s),  p = sigmoid(x), whole batch
//   dice_score                  util.py:22-26     (2*sum(pred*t) + s) / (sum(pred) + sum(t) + s) on hard masks
//   mask                        train_worker.py:24,72; test.py:144   ge(sigmoid(x), 0.5).float()
//   softmax-argmax              BASELINE.json configs[0]/[4] (classes >= 2): argmax_c x_c, ties -> lowest index
//   stitch                      visualize/predict_overlay.ipynb (missing) per configs/visualize_config.yaml:3-15
//   overlay                     README.md:28-35 tumour overlay (notebook missing): integer blend, spec in oracle/util_oracle.py
//   Adam                        torch.optim.Adam(lr) at train_worker.py:184 (betas .9/.999, eps 1e-8, no weight decay)
// Reductions accumulate per-thread in fp32, per-block and globally in fp64 (the reference's fp32 tree sum is only
// approximately exact at 2^24+ elements), counts in int64.
#include "common.h"
#include "vec.cuh"

namespace pdac {

// The reference's sigmoid, as ATen computes it in fp32 on CUDA: 1 / (1 + exp(-x)).
__device__ __forceinline__ float sigmoid_ref(float x) { return 1.0f / (1.0f + expf(-x)); }

__device__ __forceinline__ double block_sum_d(double v, double* sh) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) sh[warp] = v;
  __syncthreads();
  double r = 0.0;
  if (warp == 0) {
    r = lane < (int)(blockDim.x >> 5) ? sh[lane] : 0.0;
    for (int o = 16; o > 0; o >>= 1) r += __shfl_xor_sync(0xffffffffu, r, o);
  }
  __syncthreads();
  return r;
}

template <typename T>
__device__ __forceinline__ float tgt_val(const T* t, long long i) { return (float)t[i]; }

// sums[0] = sum p*t, sums[1] = sum p, sums[2] = sum t
template <typename T>
__global__ void __launch_bounds__(256) dice_fwd_kernel(const float* __restrict__ x, const T* __restrict__ t, long long n,
                                                       double* __restrict__ sums) {
  __shared__ double sh[8];
  float i0 = 0.f, p0 = 0.f, t0 = 0.f;
  double I = 0, P = 0, Tt = 0;
  int cnt = 0;
  // four elements per load when both arrays allow it (the scalar loop ran at 1 TB/s: one 4-byte and one 1-byte load per
  // sigmoid); the element -> thread assignment is a fixed function of n and the launch, so the sums stay reproducible
  const bool vec4 = sizeof(T) == 1 && ((reinterpret_cast<uintptr_t>(x) & 15) == 0) && ((reinterpret_cast<uintptr_t>(t) & 3) == 0);
  const long long n4 = vec4 ? n >> 2 : 0;
  for (long long q = blockIdx.x * (long long)blockDim.x + threadIdx.x; q < n4; q += (long long)gridDim.x * blockDim.x) {
    const float4 xv = __ldg(reinterpret_cast<const float4*>(x) + q);
    const uchar4 tv4 = __ldg(reinterpret_cast<const uchar4*>(t) + q);
    const float xs[4] = {xv.x, xv.y, xv.z, xv.w};
    const float ts[4] = {(float)tv4.x, (float)tv4.y, (float)tv4.z, (float)tv4.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float p = sigmoid_ref(xs[k]);
      i0 = fmaf(p, ts[k], i0);
      p0 += p;
      t0 += ts[k];
    }
    if (++cnt == 16) {  // flush the fp32 partials before they lose bits
      I += i0; P += p0; Tt += t0;
      i0 = p0 = t0 = 0.f;
      cnt = 0;
    }
  }
  for (long long i = n4 * 4 + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float p = sigmoid_ref(__ldg(x + i));
    const float tv = tgt_val(t, i);
    i0 = fmaf(p, tv, i0);
    p0 += p;
    t0 += tv;
    if (++cnt == 64) {
      I += i0; P += p0; Tt += t0;
      i0 = p0 = t0 = 0.f;
      cnt = 0;
    }
  }
  I += i0; P += p0; Tt += t0;
  I = block_sum_d(I, sh);
  P = block_sum_d(P, sh);
  Tt = block_sum_d(Tt, sh);
  if (threadIdx.x == 0) {  // this block's row of the partial table sums[3 + 3 * block]
    double* row = sums + 3 + 3 * (long long)blockIdx.x;
    row[0] = I;
    row[1] = P;
    row[2] = Tt;
  }
}
// One warp adds the block rows in a fixed order (lane-strided, then a fixed shuffle tree): the loss and the sums the
// backward pass reads are bit-reproducible (no floating-point atomics).
__global__ void dice_finalize_kernel(double* __restrict__ sums, int n_rows, float smooth, float* __restrict__ loss) {
  double a[3] = {0.0, 0.0, 0.0};
  for (int r = threadIdx.x; r < n_rows; r += 32)
    for (int k = 0; k < 3; ++k) a[k] += sums[3 + 3 * (long long)r + k];
  for (int k = 0; k < 3; ++k)
    for (int o = 16; o > 0; o >>= 1) a[k] += __shfl_xor_sync(0xffffffffu, a[k], o);
  if (threadIdx.x == 0) {
    sums[0] = a[0];
    sums[1] = a[1];
    sums[2] = a[2];
    const float I = (float)a[0], P = (float)a[1], T = (float)a[2];
    const float dice = (2.0f * I + smooth) / (P + T + smooth);
    *loss = 1.0f - dice;
  }
}
// dL/dx = -gout * [2 t D - (2I + s)] / D^2 * p (1 - p),  D = P + T + s
template <typename T>
__global__ void __launch_bounds__(256) dice_bwd_kernel(const float* __restrict__ x, const T* __restrict__ t, long long n,
                                                       const double* __restrict__ sums, float smooth,
                                                       const float* __restrict__ gout, float* __restrict__ gx) {
  const double D = sums[1] + sums[2] + (double)smooth;
  const double num = 2.0 * sums[0] + (double)smooth;
  const float go = gout ? *gout : 1.f;
  const float k1 = (float)(-2.0 / D) * go, k0 = (float)(num / (D * D)) * go;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float p = sigmoid_ref(__ldg(x + i));
    const float tv = tgt_val(t, i);
    gx[i] = (k1 * tv + k0) * p * (1.f - p);
  }
}

// The same gradient written straight into the layout the head convolution's backward reads: bf16 NHWC with the single
// class in channel 0 of `cpad` (a multiple of 8) zero-padded channels — saves the fp32 round trip through HBM and the
// separate layout pass of the 1-class training step (train_worker.py:150-182 all use classes = 1).
template <typename T>
__global__ void __launch_bounds__(256) dice_bwd_nhwc_kernel(const float* __restrict__ x, const T* __restrict__ t,
                                                            long long n, const double* __restrict__ sums, float smooth,
                                                            __nv_bfloat16* __restrict__ gx, int cpad) {
  const double D = sums[1] + sums[2] + (double)smooth;
  const double num = 2.0 * sums[0] + (double)smooth;
  const float k1 = (float)(-2.0 / D), k0 = (float)(num / (D * D));
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float p = sigmoid_ref(__ldg(x + i));
    const float tv = tgt_val(t, i);
    const float g = (k1 * tv + k0) * p * (1.f - p);
    uint4 v = make_uint4((uint32_t)__bfloat16_as_ushort(__float2bfloat16(g)), 0u, 0u, 0u);
    uint4* o = reinterpret_cast<uint4*>(gx + i * cpad);
    o[0] = v;
    for (int k = 1; k < cpad / 8; ++k) o[k] = make_uint4(0u, 0u, 0u, 0u);
  }
}

// mask = ge(sigmoid(x), 0.5)
__global__ void __launch_bounds__(256) sigmoid_mask_kernel(const float* __restrict__ x, long long n,
                                                           float* __restrict__ mf, uint8_t* __restrict__ mu) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const bool on = sigmoid_ref(__ldg(x + i)) >= 0.5f;
    if (mf) mf[i] = on ? 1.f : 0.f;
    if (mu) mu[i] = on ? 1 : 0;
  }
}

// counts[0] = sum pred*t, counts[1] = sum pred, counts[2] = sum t  (pred from logits via the mask rule, or given)
template <typename P, typename T>
__global__ void __launch_bounds__(256) dice_count_kernel(const P* __restrict__ pred, int pred_is_logits,
                                                         const T* __restrict__ t, long long n,
                                                         unsigned long long* __restrict__ counts) {
  __shared__ double sh[8];
  unsigned int a = 0, b = 0, c = 0;
  double A = 0, B = 0, Cc = 0;
  // torch.ge(x.sigmoid(), 0.5) in fp32: true for every x >= 0 (expf(-x) <= 1 -> 1 / (1 + e) >= 0.5 with correctly rounded
  // operations) and false for every x < -1e-6 (1 + e rounds above 2); only the sliver in between needs the arithmetic
  auto mask_rule = [](float v) { return v >= 0.f ? true : (v > -1e-6f ? sigmoid_ref(v) >= 0.5f : false); };
  const bool vec4 = sizeof(P) == 4 && sizeof(T) == 1 && ((reinterpret_cast<uintptr_t>(pred) & 15) == 0) &&
                    ((reinterpret_cast<uintptr_t>(t) & 3) == 0);
  const long long n4 = vec4 ? n >> 2 : 0;
  for (long long q = blockIdx.x * (long long)blockDim.x + threadIdx.x; q < n4; q += (long long)gridDim.x * blockDim.x) {
    const float4 xv = __ldg(reinterpret_cast<const float4*>(pred) + q);
    const uchar4 tv4 = __ldg(reinterpret_cast<const uchar4*>(t) + q);
    const float xs[4] = {xv.x, xv.y, xv.z, xv.w};
    const unsigned char ts[4] = {tv4.x, tv4.y, tv4.z, tv4.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const bool on = pred_is_logits ? mask_rule(xs[k]) : (xs[k] != 0.f);
      const bool tv = ts[k] != 0;
      a += (on && tv);
      b += on;
      c += tv;
    }
  }
  for (long long i = n4 * 4 + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float pv = (float)pred[i];
    const bool on = pred_is_logits ? mask_rule(pv) : (pv != 0.f);
    const bool tv = (float)t[i] != 0.f;
    a += (on && tv);
    b += on;
    c += tv;
  }
  A = block_sum_d((double)a, sh);
  B = block_sum_d((double)b, sh);
  Cc = block_sum_d((double)c, sh);
  if (threadIdx.x == 0) {
    atomicAdd(counts, (unsigned long long)A);
    atomicAdd(counts + 1, (unsigned long long)B);
    atomicAdd(counts + 2, (unsigned long long)Cc);
  }
}
__global__ void dice_score_finalize_kernel(const unsigned long long* __restrict__ counts, float smooth,
                                           float* __restrict__ score) {
  const float I = (float)counts[0], u = (float)counts[1] + (float)counts[2];
  *score = (2.0f * I + smooth) / (u + smooth);
}

// argmax over the channel dim of NCHW fp32 logits; ties -> lowest index; NaN wins (torch semantics)
__global__ void __launch_bounds__(256) argmax_kernel(const float* __restrict__ x, int C, long long HW, long long NHW,
                                                     uint8_t* __restrict__ out) {
  for (long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x; p < NHW; p += (long long)gridDim.x * blockDim.x) {
    const long long n = p / HW, hw = p - n * HW;
    const float* xp = x + n * C * HW + hw;
    float best = __ldg(xp);
    int bi = 0;
    for (int c = 1; c < C; ++c) {
      const float v = __ldg(xp + (long long)c * HW);
      if (!(best != best) && (v > best || v != v)) {
        best = v;
        bi = c;
      }
    }
    out[p] = (uint8_t)bi;
  }
}

// tile t (th x tw bytes) -> slide[(rows[t]*th - row0) + y][cols[t]*tw + x]; tiles outside the band are skipped
__global__ void __launch_bounds__(256) stitch_kernel(const uint8_t* __restrict__ tiles, const int32_t* __restrict__ rows,
                                                     const int32_t* __restrict__ cols, int T, int th, int tw,
                                                     uint8_t* __restrict__ slide, long long slide_h, long long slide_w,
                                                     long long row0) {
  const int vec = (tw % 16 == 0) ? 16 : 1;
  const int wpr = tw / vec;  // work items per tile row
  const long long per_tile = (long long)th * wpr;
  const long long total = (long long)T * per_tile;
  for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x) {
    const int t = (int)(e / per_tile);
    const long long r = e - (long long)t * per_tile;
    const int y = (int)(r / wpr), xw = (int)(r - (long long)y * wpr);
    const long long sy = (long long)rows[t] * th - row0 + y;
    const long long sx = (long long)cols[t] * tw + (long long)xw * vec;
    if (sy < 0 || sy >= slide_h || sx < 0 || sx + vec > slide_w) continue;
    const uint8_t* src = tiles + ((long long)t * th + y) * tw + (long long)xw * vec;
    uint8_t* dst = slide + sy * slide_w + sx;
    if (vec == 16 && (((uintptr_t)dst | (uintptr_t)src) & 15) == 0)
      *reinterpret_cast<uint4*>(dst) = __ldg(reinterpret_cast<const uint4*>(src));
    else
      for (int k = 0; k < vec; ++k) dst[k] = src[k];
  }
}

// Overlay (oracle/util_oracle.py overlay_blend): where mask != 0, out = (rgb*(256-a) + color*a) >> 8, elsewhere rgb.
// One work item = 16 pixels: one 16-byte mask load, three 16-byte HWC loads and stores. out may alias rgb.
// Two bytes are blended per 32-bit multiply-add (v*(256-a) + c*a <= 65280 never carries out of a 16-bit lane); the
// channel order of a 32-bit word repeats every three words (rgbr gbrg brgb), so the colour constants come in three
// phases: ce[] for bytes 0/2 and co[] for bytes 1/3 of a word, each c*a in a 16-bit lane.
struct OverlayConst {
  uint32_t ce[3], co[3], ca[3], ia;
};

__device__ __forceinline__ uint32_t blend_word(uint32_t w, uint32_t sel, uint32_t ce, uint32_t co, uint32_t ia) {
  const uint32_t e = (w & 0x00ff00ffu) * ia + ce;
  const uint32_t o = ((w >> 8) & 0x00ff00ffu) * ia + co;
  const uint32_t bl = ((e >> 8) & 0x00ff00ffu) | (o & 0xff00ff00u);
  return (bl & sel) | (w & ~sel);
}

// 4 mask bytes -> 3 words of HWC bytes (12 bytes of 4 pixels)
__device__ __forceinline__ void blend_4px(uint32_t m, uint32_t& w0, uint32_t& w1, uint32_t& w2, const OverlayConst& k) {
  if (m == 0u) return;
  const uint32_t nz = __vcmpne4(m, 0u);  // 0xff per non-zero mask byte
  w0 = blend_word(w0, __byte_perm(nz, 0u, 0x1000), k.ce[0], k.co[0], k.ia);  // p0 p0 p0 p1
  w1 = blend_word(w1, __byte_perm(nz, 0u, 0x2211), k.ce[1], k.co[1], k.ia);  // p1 p1 p2 p2
  w2 = blend_word(w2, __byte_perm(nz, 0u, 0x3332), k.ce[2], k.co[2], k.ia);  // p2 p3 p3 p3
}

__global__ void __launch_bounds__(256) overlay_kernel(const uint8_t* rgb, const uint8_t* __restrict__ mask,
                                                      long long npix, OverlayConst k, uint8_t* out, int vec_ok) {
  const long long groups = vec_ok ? npix / 16 : 0;
  for (long long g = blockIdx.x * (long long)blockDim.x + threadIdx.x; g < groups; g += (long long)gridDim.x * blockDim.x) {
    const uint4 m4 = __ldg(reinterpret_cast<const uint4*>(mask) + g);
    const uint4* src = reinterpret_cast<const uint4*>(rgb) + g * 3;
    uint4 a = src[0], b = src[1], c = src[2];
    blend_4px(m4.x, a.x, a.y, a.z, k);
    blend_4px(m4.y, a.w, b.x, b.y, k);
    blend_4px(m4.z, b.z, b.w, c.x, k);
    blend_4px(m4.w, c.y, c.z, c.w, k);
    uint4* dst = reinterpret_cast<uint4*>(out) + g * 3;
    dst[0] = a;
    dst[1] = b;
    dst[2] = c;
  }
  // tail pixels (and the whole image when the pointers are not 16-byte aligned)
  for (long long i = groups * 16 + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < npix;
       i += (long long)gridDim.x * blockDim.x) {
    const bool on = mask[i] != 0;
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) {
      const uint32_t old = rgb[i * 3 + ch];
      out[i * 3 + ch] = (uint8_t)(on ? (old * k.ia + k.ca[ch]) >> 8 : old);
    }
  }
}

// Adam over one flat fp32 arena (parameters, gradients and both moments share the element index)
__global__ void __launch_bounds__(256) adam_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                   float* __restrict__ m, float* __restrict__ v, long long n, float lr,
                                                   float b1, float b2, float eps, float bc1, float bc2_sqrt,
                                                   float grad_scale) {
  const float step = lr / bc1;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float gi = g[i] * grad_scale;
    const float mi = b1 * m[i] + (1.f - b1) * gi;
    const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
    m[i] = mi;
    v[i] = vi;
    p[i] -= step * mi / (sqrtf(vi) / bc2_sqrt + eps);
  }
}

// Device-side step counter and hyper-parameters: the whole train step (incl. Adam) replays from one CUDA graph, and a
// learning-rate scheduler changes lr by writing hyper[0]. hyper: [0] lr [1] beta1 [2] beta2 [3] eps [4] grad_scale
// [5] (out) lr / (1 - beta1^t) [6] (out) sqrt(1 - beta2^t) [7] 1 - beta1 [8] 1 - beta2 (rounded from double by the
// caller, as torch.optim.Adam computes them); bias corrections in double from the fp64 betas hyper64[0..1].
__global__ void adam_tick_kernel(float* __restrict__ hyper, long long* __restrict__ step) {
  const long long t = *step + 1;
  *step = t;
  const double b1 = 1.0 - (double)hyper[7], b2 = 1.0 - (double)hyper[8];
  const double bc1 = 1.0 - pow(b1, (double)t);
  const double bc2 = 1.0 - pow(b2, (double)t);
  hyper[5] = (float)((double)hyper[0] / bc1);
  hyper[6] = (float)sqrt(bc2);
}
__global__ void __launch_bounds__(256) adam_dev_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                       float* __restrict__ m, float* __restrict__ v, long long n,
                                                       const float* __restrict__ hyper) {
  const float b1 = hyper[1], b2 = hyper[2], eps = hyper[3], grad_scale = hyper[4], step = hyper[5], bc2_sqrt = hyper[6];
  const float omb1 = hyper[7], omb2 = hyper[8];
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float gi = g[i] * grad_scale;
    const float mi = b1 * m[i] + omb1 * gi;
    const float vi = b2 * v[i] + omb2 * gi * gi;
    m[i] = mi;
    v[i] = vi;
    p[i] -= step * mi / (sqrtf(vi) / bc2_sqrt + eps);
  }
}

static int rblocks(long long n) {
  long long b = (n + 255) / 256;
  const long long cap = (long long)num_sms() * 8;
  if (b > cap) b = cap;
  return b < 1 ? 1 : (int)b;
}

}  // namespace pdac

using namespace pdac;

// tgt_dtype: 0 = fp32, 1 = uint8
extern "C" int pdac_dice_fwd(const float* logits, const void* targets, int tgt_dtype, long long n, double* sums,
                             float smooth, float* loss, void* stream) {
  PDAC_REQUIRE(logits && targets && sums && loss && n >= 0, "dice_fwd: null argument");
  int blocks = 0;
  if (n > 0) {
    blocks = rblocks(n);
    if (blocks > PDAC_STAT_ROWS) blocks = PDAC_STAT_ROWS;
    if (tgt_dtype == 0)
      dice_fwd_kernel<float><<<blocks, 256, 0, (cudaStream_t)stream>>>(logits, (const float*)targets, n, sums);
    else
      dice_fwd_kernel<uint8_t><<<blocks, 256, 0, (cudaStream_t)stream>>>(logits, (const uint8_t*)targets, n, sums);
    PDAC_LAUNCH_CHECK("dice_fwd");
  }
  dice_finalize_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(sums, blocks, smooth, loss);
  PDAC_LAUNCH_CHECK("dice_finalize");
  return PDAC_OK;
}

extern "C" int pdac_dice_bwd(const float* logits, const void* targets, int tgt_dtype, long long n, const double* sums,
                             float smooth, const float* grad_out, float* grad_logits, void* stream) {
  PDAC_REQUIRE(logits && targets && sums && grad_logits, "dice_bwd: null argument");
  if (n <= 0) return PDAC_OK;
  if (tgt_dtype == 0)
    dice_bwd_kernel<float><<<rblocks(n), 256, 0, (cudaStream_t)stream>>>(logits, (const float*)targets, n, sums, smooth,
                                                                          grad_out, grad_logits);
  else
    dice_bwd_kernel<uint8_t><<<rblocks(n), 256, 0, (cudaStream_t)stream>>>(logits, (const uint8_t*)targets, n, sums,
                                                                            smooth, grad_out, grad_logits);
  PDAC_LAUNCH_CHECK("dice_bwd");
  return PDAC_OK;
}

extern "C" int pdac_dice_bwd_nhwc(const float* logits, const void* targets, int tgt_dtype, long long n, const double* sums,
                                  float smooth, void* grad_nhwc, int cpad, void* stream) {
  PDAC_REQUIRE(logits && targets && sums && grad_nhwc, "dice_bwd_nhwc: null argument");
  PDAC_REQUIRE(cpad >= 8 && cpad % 8 == 0 && ((uintptr_t)grad_nhwc & 15) == 0, "dice_bwd_nhwc: cpad must be a multiple of 8, output 16-byte aligned");
  if (n <= 0) return PDAC_OK;
  if (tgt_dtype == 0)
    dice_bwd_nhwc_kernel<float><<<rblocks(n), 256, 0, (cudaStream_t)stream>>>(logits, (const float*)targets, n, sums, smooth,
                                                                               (__nv_bfloat16*)grad_nhwc, cpad);
  else
    dice_bwd_nhwc_kernel<uint8_t><<<rblocks(n), 256, 0, (cudaStream_t)stream>>>(logits, (const uint8_t*)targets, n, sums,
                                                                                 smooth, (__nv_bfloat16*)grad_nhwc, cpad);
  PDAC_LAUNCH_CHECK("dice_bwd_nhwc");
  return PDAC_OK;
}

extern "C" int pdac_sigmoid_mask(const float* logits, long long n, float* mask_f32, uint8_t* mask_u8, void* stream) {
  PDAC_REQUIRE(logits && (mask_f32 || mask_u8), "sigmoid_mask: null argument");
  if (n <= 0) return PDAC_OK;
  sigmoid_mask_kernel<<<rblocks(n), 256, 0, (cudaStream_t)stream>>>(logits, n, mask_f32, mask_u8);
  PDAC_LAUNCH_CHECK("sigmoid_mask");
  return PDAC_OK;
}

// pred_kind: 0 = fp32 hard mask {0,1}, 1 = fp32 logits (mask rule applied on the fly), 2 = uint8 mask.
// tgt_dtype: 0 = fp32, 1 = uint8. counts: 3 x uint64.
extern "C" int pdac_dice_score(const void* pred, int pred_kind, const void* target, int tgt_dtype, long long n,
                               unsigned long long* counts, float smooth, float* score, void* stream) {
  PDAC_REQUIRE(pred && target && counts && score, "dice_score: null argument");
  PDAC_CUDA(cudaMemsetAsync(counts, 0, 3 * sizeof(unsigned long long), (cudaStream_t)stream));
  if (n > 0) {
    const int b = rblocks(n);
    cudaStream_t s = (cudaStream_t)stream;
    if (pred_kind == 2) {
      if (tgt_dtype == 0)
        dice_count_kernel<uint8_t, float><<<b, 256, 0, s>>>((const uint8_t*)pred, 0, (const float*)target, n, counts);
      else
        dice_count_kernel<uint8_t, uint8_t><<<b, 256, 0, s>>>((const uint8_t*)pred, 0, (const uint8_t*)target, n, counts);
    } else {
      if (tgt_dtype == 0)
        dice_count_kernel<float, float><<<b, 256, 0, s>>>((const float*)pred, pred_kind == 1, (const float*)target, n,
                                                          counts);
      else
        dice_count_kernel<float, uint8_t><<<b, 256, 0, s>>>((const float*)pred, pred_kind == 1, (const uint8_t*)target, n,
                                                            counts);
    }
    PDAC_LAUNCH_CHECK("dice_count");
  }
  dice_score_finalize_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(counts, smooth, score);
  PDAC_LAUNCH_CHECK("dice_score_finalize");
  return PDAC_OK;
}

extern "C" int pdac_argmax_nchw(const float* logits, int N, int C, long long HW, uint8_t* out, void* stream) {
  PDAC_REQUIRE(logits && out && C >= 1 && C <= 255, "argmax: bad arguments");
  const long long nhw = (long long)N * HW;
  if (nhw <= 0) return PDAC_OK;
  argmax_kernel<<<rblocks(nhw), 256, 0, (cudaStream_t)stream>>>(logits, C, HW, nhw, out);
  PDAC_LAUNCH_CHECK("argmax");
  return PDAC_OK;
}

extern "C" int pdac_stitch_tiles(const uint8_t* tiles, const int32_t* rows, const int32_t* cols, int T, int th, int tw,
                                 uint8_t* slide, long long slide_h, long long slide_w, long long row0, void* stream) {
  PDAC_REQUIRE(tiles && rows && cols && slide && th > 0 && tw > 0 && T >= 0, "stitch: bad arguments");
  if (T == 0) return PDAC_OK;
  const long long total = (long long)T * th * (tw % 16 == 0 ? tw / 16 : tw);
  stitch_kernel<<<rblocks(total), 256, 0, (cudaStream_t)stream>>>(tiles, rows, cols, T, th, tw, slide, slide_h, slide_w,
                                                                  row0);
  PDAC_LAUNCH_CHECK("stitch");
  return PDAC_OK;
}

extern "C" int pdac_overlay_blend(const uint8_t* rgb, const uint8_t* mask, long long npix, int color_r, int color_g,
                                  int color_b, int alpha_256, uint8_t* out, void* stream) {
  PDAC_REQUIRE(alpha_256 >= 0 && alpha_256 <= 256, "overlay: alpha_256 must lie in [0, 256]");
  PDAC_REQUIRE(((color_r | color_g | color_b) & ~0xff) == 0, "overlay: colour components must lie in [0, 255]");
  PDAC_REQUIRE(npix >= 0, "overlay: negative pixel count");
  if (npix == 0) return PDAC_OK;  // empty band: the pointers may be null
  PDAC_REQUIRE(rgb && mask && out, "overlay: null buffer");
  const int vec_ok = ((((uintptr_t)rgb | (uintptr_t)mask | (uintptr_t)out) & 15) == 0) ? 1 : 0;
  const uint32_t a = (uint32_t)alpha_256;
  const uint32_t cr = (uint32_t)color_r * a, cg = (uint32_t)color_g * a, cb = (uint32_t)color_b * a;
  OverlayConst k;
  k.ca[0] = cr, k.ca[1] = cg, k.ca[2] = cb, k.ia = 256u - a;
  k.ce[0] = cr | (cb << 16), k.co[0] = cg | (cr << 16);  // word bytes r g b r
  k.ce[1] = cg | (cr << 16), k.co[1] = cb | (cg << 16);  //            g b r g
  k.ce[2] = cb | (cg << 16), k.co[2] = cr | (cb << 16);  //            b r g b
  overlay_kernel<<<rblocks(vec_ok ? (npix + 15) / 16 : npix), 256, 0, (cudaStream_t)stream>>>(rgb, mask, npix, k, out,
                                                                                             vec_ok);
  PDAC_LAUNCH_CHECK("overlay");
  return PDAC_OK;
}

extern "C" int pdac_adam_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, long long n, float lr,
                              float beta1, float beta2, float eps, int step, float grad_scale, void* stream) {
  PDAC_REQUIRE(params && grads && exp_avg && exp_avg_sq && step >= 1, "adam: bad arguments");
  if (n <= 0) return PDAC_OK;
  const float bc1 = 1.f - powf(beta1, (float)step);
  const float bc2 = 1.f - powf(beta2, (float)step);
  adam_kernel<<<rblocks(n), 256, 0, (cudaStream_t)stream>>>(params, grads, exp_avg, exp_avg_sq, n, lr, beta1, beta2, eps,
                                                            bc1, sqrtf(bc2), grad_scale);
  PDAC_LAUNCH_CHECK("adam");
  return PDAC_OK;
}

extern "C" int pdac_adam_step_dev(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, long long n,
                                  float* hyper, long long* step, void* stream) {
  PDAC_REQUIRE(params && grads && exp_avg && exp_avg_sq && hyper && step, "adam: null argument");
  adam_tick_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(hyper, step);
  PDAC_LAUNCH_CHECK("adam_tick");
  if (n <= 0) return PDAC_OK;
  adam_dev_kernel<<<rblocks(n), 256, 0, (cudaStream_t)stream>>>(params, grads, exp_avg, exp_avg_sq, n, hyper);
  PDAC_LAUNCH_CHECK("adam_dev");
  return PDAC_OK;
}
