// On-GPU training augmentation of the reference's data path (train_main.py:70-80), fused into the stem's input
// preparation: uint8 HWC patch -> Normalize(mean, std) -> ColorJitter(0.3, 0.3, 0.3, 0.3) -> OneOf([HorizontalFlip,
// RandomRotate90, VerticalFlip], p = 0.3) -> space-to-depth bf16 stem input (+ the label mask with the same geometry).
// The random draws come from a counter-based hash of (seed, step counter in DEVICE memory, patch index, draw index), so a
// CUDA-graph replay draws fresh parameters every step with no host involvement.
//
// albumentations is absent from the reference tree and unpinned: the semantics below restate its published
// transforms (PARITY UNPINNED), pinned here by an executable torch restatement (oracle/augment_oracle.py):
//   ColorJitter: applied with probability p_color (library default 0.5); brightness / contrast / saturation factors
//     ~ U(1 - r, 1 + r), hue shift ~ U(-r, r), the four operations in a uniformly random order;
//       brightness  x <- clip01(x * f)
//       contrast    x <- clip01(x * f + m * (1 - f)),  m = mean over the patch of gray(x) = .299 R + .587 G + .114 B
//                   of the patch as it is at that point of the chain (a per-patch reduction: pre-pass kernel)
//       saturation  x <- clip01(x * f + gray(x) * (1 - f))
//       hue         RGB -> HSV (H in degrees, S = (V - min) / V, 0 for V <= 0), H <- (H + 360 f) mod 360, HSV -> RGB
//     clip01 is the library's float-image path; the reference applies ColorJitter AFTER Normalize (train_main.py:73-74),
//     i.e. to zero-mean data that the first clipping operation cuts to [0, 1] — a reference quirk, reproduced as is.
//   OneOf(p = 0.3): with that probability exactly one of the three (equal weights) is applied; RandomRotate90 rotates by
//     k * 90 degrees counter-clockwise, k ~ U{0, 1, 2, 3}.
#include "common.h"
#include "vec.cuh"

namespace pdac {

struct AugNorm {
  float scale[3], shift[3];
};

// per-patch parameter record (also returned to the caller: tests, logging)
enum { AP_APPLY = 0, AP_ORDER = 1, AP_BRIGHT = 2, AP_CONTRAST = 3, AP_SAT = 4, AP_HUE = 5, AP_GEOM = 6, AP_MEAN = 7 };

__host__ __device__ inline unsigned long long splitmix64(unsigned long long z) {
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
__device__ __forceinline__ float draw_uniform(unsigned long long key, unsigned long long idx) {
  return (float)(splitmix64(key + idx * 0xD1342543DE82EF95ull) >> 40) * (1.0f / 16777216.0f);  // [0, 1), 24 bits
}

// order index 0..23 -> permutation of {0 brightness, 1 contrast, 2 saturation, 3 hue} (factorial number system)
__host__ __device__ inline void decode_order(int idx, int (&ops)[4]) {
  int pool[4] = {0, 1, 2, 3};
  int n = 4, div = 6;
  for (int k = 0; k < 4; ++k) {
    const int sel = idx / div;
    idx -= sel * div;
    ops[k] = pool[sel];
    for (int j = sel; j + 1 < n; ++j) pool[j] = pool[j + 1];
    --n;
    if (k < 2) div /= (3 - k);  // 6 -> 2 -> 1
    else div = 1;
  }
}

__global__ void aug_draw_kernel(int N, PdacAugment cfg, const long long* __restrict__ step, float* __restrict__ params) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  const unsigned long long key = splitmix64(cfg.seed ^ splitmix64((unsigned long long)(step ? *step : 0)));
  const unsigned long long base = (unsigned long long)n * 16;
  float* p = params + (size_t)n * 8;
  const bool color = draw_uniform(key, base + 0) < cfg.p_color;
  p[AP_APPLY] = color ? 1.f : 0.f;
  int order = (int)(draw_uniform(key, base + 1) * 24.f);
  p[AP_ORDER] = (float)(order > 23 ? 23 : order);
  p[AP_BRIGHT] = 1.f + (2.f * draw_uniform(key, base + 2) - 1.f) * cfg.brightness;
  p[AP_CONTRAST] = 1.f + (2.f * draw_uniform(key, base + 3) - 1.f) * cfg.contrast;
  p[AP_SAT] = 1.f + (2.f * draw_uniform(key, base + 4) - 1.f) * cfg.saturation;
  p[AP_HUE] = (2.f * draw_uniform(key, base + 5) - 1.f) * cfg.hue;
  int geom = 0;  // 0 none, 1 hflip, 2 vflip, 3 rot90, 4 rot180, 5 rot270
  if (draw_uniform(key, base + 6) < cfg.p_geom) {
    int which = (int)(draw_uniform(key, base + 7) * 3.f);
    if (which > 2) which = 2;
    if (which == 0) geom = 1;
    else if (which == 2) geom = 2;
    else {
      int k = (int)(draw_uniform(key, base + 8) * 4.f);
      if (k > 3) k = 3;
      geom = k == 0 ? 0 : 2 + k;
    }
  }
  p[AP_GEOM] = (float)geom;
  p[AP_MEAN] = 0.f;
}

__device__ __forceinline__ float clip01(float v) { return fminf(fmaxf(v, 0.f), 1.f); }
__device__ __forceinline__ float gray_of(const float (&c)[3]) {
  return fmaf(0.114f, c[2], fmaf(0.587f, c[1], 0.299f * c[0]));
}

__device__ __forceinline__ void hue_shift(float (&c)[3], float f) {
  const float r = c[0], g = c[1], b = c[2];
  const float v = fmaxf(r, fmaxf(g, b)), mn = fminf(r, fminf(g, b));
  const float d = v - mn;
  const float s = v > 0.f ? d / v : 0.f;
  float h = 0.f;
  if (d > 0.f) {
    if (v == r) h = 60.f * (g - b) / d;
    else if (v == g) h = 120.f + 60.f * (b - r) / d;
    else h = 240.f + 60.f * (r - g) / d;
    if (h < 0.f) h += 360.f;
  }
  h = h + 360.f * f;
  h = h - 360.f * floorf(h * (1.f / 360.f));
  if (h >= 360.f) h -= 360.f;
  const float sect = h * (1.f / 60.f);
  const int i = (int)floorf(sect);
  const float fr = sect - (float)i;
  const float p = v * (1.f - s), q = v * (1.f - s * fr), t = v * (1.f - s * (1.f - fr));
  switch (i) {
    case 0: c[0] = v; c[1] = t; c[2] = p; break;
    case 1: c[0] = q; c[1] = v; c[2] = p; break;
    case 2: c[0] = p; c[1] = v; c[2] = t; break;
    case 3: c[0] = p; c[1] = q; c[2] = v; break;
    case 4: c[0] = t; c[1] = p; c[2] = v; break;
    default: c[0] = v; c[1] = p; c[2] = q; break;
  }
}

// applies the colour operations ops[0 .. n_ops-1] to one pixel
__device__ __forceinline__ void color_chain(float (&c)[3], const int (&ops)[4], int n_ops, const float* __restrict__ p,
                                            float mean) {
  for (int k = 0; k < n_ops; ++k) {
    switch (ops[k]) {
      case 0: {
        const float f = p[AP_BRIGHT];
#pragma unroll
        for (int i = 0; i < 3; ++i) c[i] = clip01(c[i] * f);
        break;
      }
      case 1: {
        const float f = p[AP_CONTRAST], a = mean * (1.f - f);
#pragma unroll
        for (int i = 0; i < 3; ++i) c[i] = clip01(fmaf(c[i], f, a));
        break;
      }
      case 2: {
        const float f = p[AP_SAT], a = gray_of(c) * (1.f - f);
#pragma unroll
        for (int i = 0; i < 3; ++i) c[i] = clip01(fmaf(c[i], f, a));
        break;
      }
      default: hue_shift(c, p[AP_HUE]); break;
    }
  }
}

// contrast's per-patch mean: gray of the patch after the operations that precede contrast in this patch's order. One
// block per patch, thread-strided fp32 partials, fixed-order double tree: bit-reproducible.
__global__ void __launch_bounds__(1024) aug_mean_kernel(const uint8_t* __restrict__ img, int HW, AugNorm nm,
                                                        float* __restrict__ params) {
  __shared__ double sh[1024];
  const int n = blockIdx.x;
  float* p = params + (size_t)n * 8;
  if (p[AP_APPLY] == 0.f) return;  // uniform over the block
  int ops[4];
  decode_order((int)p[AP_ORDER], ops);
  int n_before = 0;
  while (ops[n_before] != 1) ++n_before;
  const uint8_t* ip = img + (size_t)n * HW * 3;
  double acc = 0.0;
  float part = 0.f;
  int cnt = 0;
  for (int i = threadIdx.x; i < HW; i += blockDim.x) {
    float c[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) c[k] = fmaf((float)ip[(size_t)i * 3 + k], nm.scale[k], nm.shift[k]);
    color_chain(c, ops, n_before, p, 0.f);
    part += gray_of(c);
    if (++cnt == 32) {
      acc += (double)part;
      part = 0.f;
      cnt = 0;
    }
  }
  acc += (double)part;
  sh[threadIdx.x] = acc;
  __syncthreads();
  for (int s = 512; s > 0; s >>= 1) {
    if ((int)threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s];
    __syncthreads();
  }
  if (threadIdx.x == 0) p[AP_MEAN] = (float)(sh[0] / (double)HW);
}

__device__ __forceinline__ void geom_src(int code, int H, int W, int y, int x, int& sy, int& sx) {
  switch (code) {
    case 1: sy = y; sx = W - 1 - x; break;          // horizontal flip
    case 2: sy = H - 1 - y; sx = x; break;          // vertical flip
    case 3: sy = x; sx = W - 1 - y; break;          // np.rot90(k=1): out[y][x] = in[x][W-1-y]
    case 4: sy = H - 1 - y; sx = W - 1 - x; break;  // k=2
    case 5: sy = H - 1 - x; sx = y; break;          // k=3: out[y][x] = in[H-1-x][y]
    default: sy = y; sx = x; break;
  }
}

__global__ void __launch_bounds__(256) stem_prep_aug_kernel(const uint8_t* __restrict__ img,
                                                            const uint8_t* __restrict__ mask_in, int N, int H, int W,
                                                            AugNorm nm, const float* __restrict__ params,
                                                            bf16* __restrict__ out, uint8_t* __restrict__ mask_out) {
  const int Ho = H / 2, Wo = W / 2;
  const long long total = (long long)N * Ho * Wo;
  for (long long o = blockIdx.x * (long long)blockDim.x + threadIdx.x; o < total; o += (long long)gridDim.x * blockDim.x) {
    const int j = (int)(o % Wo);
    const long long r = o / Wo;
    const int i = (int)(r % Ho);
    const long long n = r / Ho;
    const float* p = params + (size_t)n * 8;
    const int code = (int)p[AP_GEOM];
    const bool color = p[AP_APPLY] != 0.f;
    int ops[4] = {0, 1, 2, 3};
    if (color) decode_order((int)p[AP_ORDER], ops);
    const float mean = p[AP_MEAN];
    const uint8_t* ip = img + n * (long long)H * W * 3;
    float v[16];
#pragma unroll
    for (int k = 12; k < 16; ++k) v[k] = 0.f;
#pragma unroll
    for (int pp = 0; pp < 2; ++pp)
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        int sy, sx;
        geom_src(code, H, W, 2 * i + pp, 2 * j + q, sy, sx);
        const uint8_t* px = ip + ((long long)sy * W + sx) * 3;
        float c[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) c[k] = fmaf((float)px[k], nm.scale[k], nm.shift[k]);
        if (color) color_chain(c, ops, 4, p, mean);
#pragma unroll
        for (int k = 0; k < 3; ++k) v[(pp * 2 + q) * 3 + k] = c[k];
        if (mask_in)
          mask_out[n * (long long)H * W + (long long)(2 * i + pp) * W + 2 * j + q] =
              mask_in[n * (long long)H * W + (long long)sy * W + sx];
      }
    F8 lo, hi;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      lo.v[k] = v[k];
      hi.v[k] = v[8 + k];
    }
    st16(out + o * 16, pack8(lo));
    st16(out + o * 16 + 8, pack8(hi));
  }
}

}  // namespace pdac

using namespace pdac;

extern "C" int pdac_stem_prep_u8_augment(const uint8_t* img, const uint8_t* mask_in, int N, int H, int W,
                                         const float* mean, const float* stdv, const PdacAugment* cfg,
                                         const long long* step, float* params, void* out, uint8_t* mask_out,
                                         void* stream) {
  PDAC_REQUIRE(img && out && mean && stdv && cfg && params, "null argument");
  PDAC_REQUIRE(H % 2 == 0 && W % 2 == 0, "stem_prep: H, W must be even");
  PDAC_REQUIRE(!mask_in || mask_out, "mask_in without mask_out");
  PDAC_REQUIRE(cfg->p_geom == 0.f || H == W, "rot90 augmentation needs square patches");
  PDAC_REQUIRE(cfg->brightness >= 0.f && cfg->brightness <= 1.f && cfg->contrast >= 0.f && cfg->contrast <= 1.f &&
                   cfg->saturation >= 0.f && cfg->saturation <= 1.f && cfg->hue >= 0.f && cfg->hue <= 0.5f,
               "ColorJitter ranges out of bounds");
  AugNorm nm;
  for (int c = 0; c < 3; ++c) {
    nm.scale[c] = 1.f / (255.f * stdv[c]);
    nm.shift[c] = -mean[c] / stdv[c];
  }
  const long long total = (long long)N * (H / 2) * (W / 2);
  if (!total) return PDAC_OK;
  cudaStream_t s = (cudaStream_t)stream;
  aug_draw_kernel<<<ceil_div(N, 128), 128, 0, s>>>(N, *cfg, step, params);
  PDAC_LAUNCH_CHECK("aug_draw");
  if (cfg->p_color > 0.f) {
    aug_mean_kernel<<<N, 1024, 0, s>>>(img, H * W, nm, params);
    PDAC_LAUNCH_CHECK("aug_mean");
  }
  long long blocks = (total + 255) / 256;
  const long long cap = (long long)num_sms() * 16;
  if (blocks > cap) blocks = cap;
  stem_prep_aug_kernel<<<(int)blocks, 256, 0, s>>>(img, mask_in, N, H, W, nm, params, (bf16*)out, mask_out);
  PDAC_LAUNCH_CHECK("stem_prep_aug");
  return PDAC_OK;
}
