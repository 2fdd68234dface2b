// Pooling / resampling passes on NHWC bf16 (all HBM-bound, 16-byte vectorised, gather-form backward = deterministic):
//   MaxPool2d(3, stride 2, pad 1)                 torchvision resnet.py:200 (run at models/resunet.py:28)
//   nearest x2 up-sampling                         smp UnetDecoder / FPNBlock
//   bilinear resize, align_corners False / True    models/resunet.py:51-56; smp SegmentationHead / FPN / PSP / ASPP
//   AdaptiveAvgPool2d                              smp PSPModule / ASPPPooling (App. B.4/B.5)
#include "common.h"
#include "vec.cuh"

namespace pdac {

struct V4 {  // device-side view
  bf16* ptr;
  int N, H, W, C;
  long long sN, sH, sW;
  __device__ __forceinline__ bf16* at(int n, int h, int w) const { return ptr + n * sN + h * sH + w * sW; }
};
static int to_v4(const PdacView* v, V4* o, const char* what) {
  PDAC_REQUIRE(v && v->ptr, "%s: null view", what);
  PDAC_REQUIRE(v->C % 8 == 0, "%s: channels (%d) must be a multiple of 8", what, v->C);
  PDAC_REQUIRE(((uintptr_t)v->ptr & 15) == 0 && v->sW % 8 == 0 && v->sH % 8 == 0 && v->sN % 8 == 0,
               "%s: view must be 16-byte aligned", what);
  o->ptr = (bf16*)v->ptr;
  o->N = v->N;
  o->H = v->H;
  o->W = v->W;
  o->C = v->C;
  o->sN = v->sN;
  o->sH = v->sH;
  o->sW = v->sW;
  return PDAC_OK;
}
static int blocks_for(long long n) {
  long long b = (n + 255) / 256;
  const long long cap = (long long)num_sms() * 16;
  if (b > cap) b = cap;
  return b < 1 ? 1 : (int)b;
}
#define PIX_LOOP(total)                                                                               \
  for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < (total);                    \
       e += (long long)gridDim.x * blockDim.x)
#define DECODE(e, CG, Wd, Hd)            \
  const int g = (int)((e) % (CG));       \
  long long r_ = (e) / (CG);             \
  const int w = (int)(r_ % (Wd));        \
  r_ /= (Wd);                            \
  const int h = (int)(r_ % (Hd));        \
  const int n = (int)(r_ / (Hd));

// ------------------------------------------------------------------------------------------------ maxpool 3x3 s2 p1
// Packed bf16x2 arithmetic throughout: the first version unpacked every tap to fp32 and tracked value and index with a
// compare + two selects per element (~400 instructions per thread, instruction-bound at 2.5 TB/s). Here the maximum is
// 8 HMNMX2 per 32-bit word and the arg-max is "the first tap equal to the maximum" (one compare-to-mask and one LOP3 per
// word and tap, walked backwards so the first match wins) — torch's `val > maxval` scan keeps the first maximum too.
// NaNs (torch: the LAST NaN tap wins) take a scalar path.
__global__ void __launch_bounds__(256) maxpool_fwd_kernel(V4 x, V4 y, uint8_t* __restrict__ idx) {
  const int cg = x.C / 8;
  const long long total = (long long)y.N * y.H * y.W * cg;
  PIX_LOOP(total) {
    DECODE(e, cg, y.W, y.H)
    uint4 v[9];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int ih = 2 * h - 1 + r;
#pragma unroll
      for (int s = 0; s < 3; ++s) {
        const int iw = 2 * w - 1 + s;
        const bool ok = ih >= 0 && ih < x.H && iw >= 0 && iw < x.W;
        v[r * 3 + s] = ok ? ld16(x.at(n, ih, iw) + g * 8) : make_uint4(0xFF80FF80u, 0xFF80FF80u, 0xFF80FF80u, 0xFF80FF80u);
      }
    }
    uint32_t m[4], ip[4];
    bool any_nan = false;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      __nv_bfloat162 mm = *reinterpret_cast<const __nv_bfloat162*>(&(&v[0].x)[k]);
#pragma unroll
      for (int t = 1; t < 9; ++t) mm = __hmax2_nan(mm, *reinterpret_cast<const __nv_bfloat162*>(&(&v[t].x)[k]));
      m[k] = *reinterpret_cast<const uint32_t*>(&mm);
      any_nan = any_nan || __hisnan(mm.x) || __hisnan(mm.y);
      uint32_t id = 0x00040004u;
#pragma unroll
      for (int t = 8; t >= 0; --t) {
        const uint32_t eq = __heq2_mask(*reinterpret_cast<const __nv_bfloat162*>(&(&v[t].x)[k]), mm);
        id = (id & ~eq) | ((uint32_t)t * 0x00010001u & eq);
      }
      ip[k] = id;
    }
    if (any_nan) {  // rare: redo this vector the way torch scans it
      float best[8];
      int bi[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        best[i] = -INFINITY;
        bi[i] = 4;
      }
      for (int t = 0; t < 9; ++t) {
        const int ih = 2 * h - 1 + t / 3, iw = 2 * w - 1 + t % 3;
        if (ih < 0 || ih >= x.H || iw < 0 || iw >= x.W) continue;
        const F8 f = unpack8(v[t]);
#pragma unroll
        for (int i = 0; i < 8; ++i)
          if (f.v[i] > best[i] || f.v[i] != f.v[i]) {
            best[i] = f.v[i];
            bi[i] = t;
          }
      }
      F8 o;
#pragma unroll
      for (int i = 0; i < 8; ++i) o.v[i] = best[i];
      const uint4 pk = pack8(o);
      m[0] = pk.x, m[1] = pk.y, m[2] = pk.z, m[3] = pk.w;
#pragma unroll
      for (int k = 0; k < 4; ++k) ip[k] = (uint32_t)bi[2 * k] | ((uint32_t)bi[2 * k + 1] << 16);
    }
    st16(y.at(n, h, w) + g * 8, make_uint4(m[0], m[1], m[2], m[3]));
    if (idx) {
      uint2 p;  // byte 2k = low half of ip[k], byte 2k + 1 = high half
      p.x = __byte_perm(ip[0], ip[1], 0x6420);
      p.y = __byte_perm(ip[2], ip[3], 0x6420);
      *reinterpret_cast<uint2*>(idx + (((long long)n * y.H + h) * y.W + w) * x.C + g * 8) = p;
    }
  }
}

// One thread produces the 2x2 block (2i+a, 2j+b) of dx for 8 channels: those four pixels are covered by the windows
// (i,j), (i,j+1), (i+1,j), (i+1,j+1) only, so each window's arg-max bytes and gradient are loaded once for four outputs.
__global__ void __launch_bounds__(256) maxpool_bwd_kernel(V4 dy, const uint8_t* __restrict__ idx, V4 add, V4 dx) {
  const int cg = dx.C / 8;
  const int bh = (dx.H + 1) >> 1, bw = (dx.W + 1) >> 1;
  const long long total = (long long)dx.N * bh * bw * cg;
  PIX_LOOP(total) {
    DECODE(e, cg, bw, bh)
    float acc[4][8];
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[q][i] = 0.f;
#pragma unroll
    for (int a = 0; a < 2; ++a) {
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        const int oh = h + a, ow = w + b;  // window (oh, ow) covers rows 2*oh-1 .. 2*oh+1
        if (oh >= dy.H || ow >= dy.W) continue;
        const uint2 p = __ldg(reinterpret_cast<const uint2*>(idx + (((long long)n * dy.H + oh) * dy.W + ow) * dx.C + g * 8));
        const F8 d = unpack8(ld16(dy.at(n, oh, ow) + g * 8));
#pragma unroll
        for (int pa = 0; pa < 2; ++pa) {
#pragma unroll
          for (int pb = 0; pb < 2; ++pb) {
            // pixel (2h+pa, 2w+pb) sits at tap (2h+pa - (2oh-1), 2w+pb - (2ow-1)) of this window, if inside 0..2
            const int tr = pa + 1 - 2 * a, tc = pb + 1 - 2 * b;
            if (tr < 0 || tr > 2 || tc < 0 || tc > 2) continue;
            const int tap = tr * 3 + tc;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const int bt = (int)(((i < 4 ? p.x : p.y) >> ((i & 3) * 8)) & 0xff);
              if (bt == tap) acc[pa * 2 + pb][i] += d.v[i];
            }
          }
        }
      }
    }
#pragma unroll
    for (int pa = 0; pa < 2; ++pa)
#pragma unroll
      for (int pb = 0; pb < 2; ++pb) {
        const int hh = 2 * h + pa, ww = 2 * w + pb;
        if (hh >= dx.H || ww >= dx.W) continue;
        F8 o;
#pragma unroll
        for (int i = 0; i < 8; ++i) o.v[i] = acc[pa * 2 + pb][i];
        if (add.ptr) {  // gradient already accumulated at this tensor by another consumer (skip connection)
          const F8 a = unpack8(ld16(add.at(n, hh, ww) + g * 8));
#pragma unroll
          for (int i = 0; i < 8; ++i) o.v[i] += a.v[i];
        }
        st16(dx.at(n, hh, ww) + g * 8, pack8(o));
      }
  }
}

// ------------------------------------------------------------------------------------------------ nearest x2
__global__ void __launch_bounds__(256) up2_fwd_kernel(V4 x, V4 y) {
  const int cg = x.C / 8;
  const long long total = (long long)x.N * x.H * x.W * cg;  // one thread per INPUT vector: 1 load, 4 stores
  PIX_LOOP(total) {
    DECODE(e, cg, x.W, x.H)
    const uint4 v = ld16_stream(x.at(n, h, w) + g * 8);
    bf16* o = y.at(n, 2 * h, 2 * w) + g * 8;
    st16(o, v);
    st16(o + y.sW, v);
    st16(o + y.sH, v);
    st16(o + y.sH + y.sW, v);
  }
}
// dx = sum of the 2x2 block of dy (+ addend)
__global__ void __launch_bounds__(256) up2_bwd_kernel(V4 dy, V4 add, V4 dx) {
  const int cg = dx.C / 8;
  const long long total = (long long)dx.N * dx.H * dx.W * cg;
  PIX_LOOP(total) {
    DECODE(e, cg, dx.W, dx.H)
    const F8 a = unpack8(ld16_stream(dy.at(n, 2 * h, 2 * w) + g * 8));
    const F8 b = unpack8(ld16_stream(dy.at(n, 2 * h, 2 * w + 1) + g * 8));
    const F8 c = unpack8(ld16_stream(dy.at(n, 2 * h + 1, 2 * w) + g * 8));
    const F8 d = unpack8(ld16_stream(dy.at(n, 2 * h + 1, 2 * w + 1) + g * 8));
    F8 o;
#pragma unroll
    for (int i = 0; i < 8; ++i) o.v[i] = (a.v[i] + b.v[i]) + (c.v[i] + d.v[i]);
    if (add.ptr) {
      const F8 z = unpack8(ld16_stream(add.at(n, h, w) + g * 8));
#pragma unroll
      for (int i = 0; i < 8; ++i) o.v[i] += z.v[i];
    }
    st16(dx.at(n, h, w) + g * 8, pack8(o));
  }
}

// ------------------------------------------------------------------------------------------------ bilinear
// Same index arithmetic as ATen's upsample_bilinear2d (area_pixel_compute_source_index).
struct Lin {
  float scale;  // align_corners ? (in-1)/(out-1) : in/out
  int align;
  int in;
  __device__ __forceinline__ void src(int o, int& i0, int& i1, float& l1) const {
    float s = align ? scale * o : fmaxf(scale * (o + 0.5f) - 0.5f, 0.f);
    i0 = (int)s;
    if (i0 > in - 1) i0 = in - 1;
    i1 = i0 + (i0 < in - 1 ? 1 : 0);
    l1 = s - (float)i0;
  }
};
static Lin make_lin(int in, int out, int align) {
  Lin l;
  l.align = align;
  l.in = in;
  l.scale = align ? (out > 1 ? (float)(in - 1) / (float)(out - 1) : 0.f) : (float)in / (float)out;
  return l;
}

__global__ void __launch_bounds__(256) bilinear_fwd_kernel(V4 x, V4 y, Lin lh, Lin lw) {
  const int cg = x.C / 8;
  const long long total = (long long)y.N * y.H * y.W * cg;
  PIX_LOOP(total) {
    DECODE(e, cg, y.W, y.H)
    int h0, h1, w0, w1;
    float a, b;
    lh.src(h, h0, h1, a);
    lw.src(w, w0, w1, b);
    const F8 v00 = unpack8(ld16(x.at(n, h0, w0) + g * 8)), v01 = unpack8(ld16(x.at(n, h0, w1) + g * 8));
    const F8 v10 = unpack8(ld16(x.at(n, h1, w0) + g * 8)), v11 = unpack8(ld16(x.at(n, h1, w1) + g * 8));
    F8 o;
#pragma unroll
    for (int i = 0; i < 8; ++i)
      o.v[i] = (1.f - a) * ((1.f - b) * v00.v[i] + b * v01.v[i]) + a * ((1.f - b) * v10.v[i] + b * v11.v[i]);
    st16(y.at(n, h, w) + g * 8, pack8(o));
  }
}

// candidate output range whose 2-tap footprint can touch input index i (checked exactly by the caller)
__device__ __forceinline__ void out_range(const Lin& l, int i, int out, int& lo, int& hi) {
  if (l.scale <= 0.f) {
    lo = 0;
    hi = out - 1;
    return;
  }
  const float inv = 1.f / l.scale;
  float flo, fhi;
  if (l.align) {
    flo = (i - 1) * inv;
    fhi = (i + 1) * inv;
  } else {
    flo = (i - 1 + 0.5f) * inv - 0.5f;
    fhi = (i + 1 + 0.5f) * inv - 0.5f;
  }
  lo = (int)floorf(flo) - 1;
  hi = (int)ceilf(fhi) + 1;
  if (lo < 0) lo = 0;
  if (hi > out - 1) hi = out - 1;
}

__global__ void __launch_bounds__(256) bilinear_bwd_kernel(V4 dy, V4 dx, Lin lh, Lin lw) {
  const int cg = dx.C / 8;
  const long long total = (long long)dx.N * dx.H * dx.W * cg;
  PIX_LOOP(total) {
    DECODE(e, cg, dx.W, dx.H)
    int ohl, ohh, owl, owh;
    out_range(lh, h, dy.H, ohl, ohh);
    out_range(lw, w, dy.W, owl, owh);
    float acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = 0.f;
    for (int oh = ohl; oh <= ohh; ++oh) {
      int h0, h1;
      float a;
      lh.src(oh, h0, h1, a);
      const float wh = (h0 == h ? 1.f - a : 0.f) + (h1 == h ? a : 0.f);
      if (wh == 0.f) continue;
      for (int ow = owl; ow <= owh; ++ow) {
        int w0, w1;
        float b;
        lw.src(ow, w0, w1, b);
        const float ww = (w0 == w ? 1.f - b : 0.f) + (w1 == w ? b : 0.f);
        if (ww == 0.f) continue;
        const F8 d = unpack8(ld16(dy.at(n, oh, ow) + g * 8));
        const float k = wh * ww;
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i] = fmaf(k, d.v[i], acc[i]);
      }
    }
    F8 o;
#pragma unroll
    for (int i = 0; i < 8; ++i) o.v[i] = acc[i];
    st16(dx.at(n, h, w) + g * 8, pack8(o));
  }
}

// Sum of the pixel lanes' partial 8-channel vectors of a block (thread = (pixel lane, channel group g), cg groups):
// returns true on the lane-0 threads, whose acc[] then holds the block total.
__device__ __forceinline__ bool lanes_reduce(float (&acc)[8], int cg) {
  __shared__ float red[256][9];
#pragma unroll
  for (int i = 0; i < 8; ++i) red[threadIdx.x][i] = acc[i];
  __syncthreads();
  const int lanes = blockDim.x / cg;
  if ((int)threadIdx.x >= cg) return false;
  for (int q = 1; q < lanes; ++q)
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] += red[q * cg + threadIdx.x][i];
  return true;
}

// Small dx (PSPNet's 1x1 .. 6x6 pooling bins upsampled to 64x64): one thread per dx element would gather thousands of
// dy pixels serially from a few thousand threads (1.7 ms per call); here one block owns a dx pixel and its pixel lanes
// split the contributing dy window.
__global__ void __launch_bounds__(256) bilinear_bwd_coop_kernel(V4 dy, V4 dx, Lin lh, Lin lw) {
  const int cg = dx.C / 8;
  const int g = threadIdx.x % cg, pl = threadIdx.x / cg, lanes = blockDim.x / cg;
  int r = blockIdx.x;
  const int w = r % dx.W;
  r /= dx.W;
  const int h = r % dx.H, n = r / dx.H;
  int ohl, ohh, owl, owh;
  out_range(lh, h, dy.H, ohl, ohh);
  out_range(lw, w, dy.W, owl, owh);
  const int ww_n = owh - owl + 1, cnt = (ohh - ohl + 1) * ww_n;
  float acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = 0.f;
  for (int k = pl; k < cnt && pl < lanes; k += lanes) {
    const int oh = ohl + k / ww_n, ow = owl + k % ww_n;
    int h0, h1, w0, w1;
    float a, b;
    lh.src(oh, h0, h1, a);
    lw.src(ow, w0, w1, b);
    const float kk = ((h0 == h ? 1.f - a : 0.f) + (h1 == h ? a : 0.f)) * ((w0 == w ? 1.f - b : 0.f) + (w1 == w ? b : 0.f));
    if (kk == 0.f) continue;
    const F8 d = unpack8(ld16(dy.at(n, oh, ow) + g * 8));
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = fmaf(kk, d.v[i], acc[i]);
  }
  if (lanes_reduce(acc, cg)) {
    F8 o;
#pragma unroll
    for (int i = 0; i < 8; ++i) o.v[i] = acc[i];
    st16(dx.at(n, h, w) + g * 8, pack8(o));
  }
}

// head variants: NHWC bf16 (first Cu channels of a padded buffer) -> NCHW fp32 logits, and the gradient back
__global__ void __launch_bounds__(256) bilinear_head_fwd_kernel(V4 x, int Cu, float* __restrict__ out, int Ho, int Wo,
                                                                Lin lh, Lin lw) {
  const long long total = (long long)x.N * Cu * Ho * Wo;
  PIX_LOOP(total) {
    const int w = (int)(e % Wo);
    long long r = e / Wo;
    const int h = (int)(r % Ho);
    r /= Ho;
    const int c = (int)(r % Cu);
    const int n = (int)(r / Cu);
    int h0, h1, w0, w1;
    float a, b;
    lh.src(h, h0, h1, a);
    lw.src(w, w0, w1, b);
    const float v00 = __bfloat162float(x.at(n, h0, w0)[c]), v01 = __bfloat162float(x.at(n, h0, w1)[c]);
    const float v10 = __bfloat162float(x.at(n, h1, w0)[c]), v11 = __bfloat162float(x.at(n, h1, w1)[c]);
    out[e] = (1.f - a) * ((1.f - b) * v00 + b * v01) + a * ((1.f - b) * v10 + b * v11);
  }
}
__global__ void __launch_bounds__(256) bilinear_head_bwd_kernel(const float* __restrict__ g, int Cu, int Ho, int Wo,
                                                                V4 dx, Lin lh, Lin lw) {
  const long long total = (long long)dx.N * dx.H * dx.W;
  PIX_LOOP(total) {
    const int w = (int)(e % dx.W);
    long long r = e / dx.W;
    const int h = (int)(r % dx.H);
    const int n = (int)(r / dx.H);
    int ohl, ohh, owl, owh;
    out_range(lh, h, Ho, ohl, ohh);
    out_range(lw, w, Wo, owl, owh);
    bf16* op = dx.at(n, h, w);
    for (int c = 0; c < dx.C; ++c) {
      float acc = 0.f;
      if (c < Cu) {
        const float* gp = g + ((long long)n * Cu + c) * Ho * Wo;
        for (int oh = ohl; oh <= ohh; ++oh) {
          int h0, h1;
          float a;
          lh.src(oh, h0, h1, a);
          const float wh = (h0 == h ? 1.f - a : 0.f) + (h1 == h ? a : 0.f);
          if (wh == 0.f) continue;
          for (int ow = owl; ow <= owh; ++ow) {
            int w0, w1;
            float b;
            lw.src(ow, w0, w1, b);
            const float ww = (w0 == w ? 1.f - b : 0.f) + (w1 == w ? b : 0.f);
            if (ww != 0.f) acc = fmaf(wh * ww, __ldg(gp + (long long)oh * Wo + ow), acc);
          }
        }
      }
      op[c] = __float2bfloat16(acc);
    }
  }
}

// ------------------------------------------------------------------------------------------------ adaptive avg pool
__device__ __forceinline__ int bin_lo(int i, int in, int out) { return (i * in) / out; }
__device__ __forceinline__ int bin_hi(int i, int in, int out) { return ((i + 1) * in + out - 1) / out; }

__global__ void __launch_bounds__(256) adaptive_avgpool_fwd_kernel(V4 x, V4 y) {
  const int cg = x.C / 8;
  const long long total = (long long)y.N * y.H * y.W * cg;
  PIX_LOOP(total) {
    DECODE(e, cg, y.W, y.H)
    const int h0 = bin_lo(h, x.H, y.H), h1 = bin_hi(h, x.H, y.H), w0 = bin_lo(w, x.W, y.W), w1 = bin_hi(w, x.W, y.W);
    float acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = 0.f;
    for (int ih = h0; ih < h1; ++ih)
      for (int iw = w0; iw < w1; ++iw) {
        const F8 v = unpack8(ld16(x.at(n, ih, iw) + g * 8));
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i] += v.v[i];
      }
    const float inv = 1.f / (float)((h1 - h0) * (w1 - w0));
    F8 o;
#pragma unroll
    for (int i = 0; i < 8; ++i) o.v[i] = acc[i] * inv;
    st16(y.at(n, h, w) + g * 8, pack8(o));
  }
}
// large bins (PSPNet: 64x64 -> 1x1 .. 6x6): one block per output pixel, its pixel lanes split the bin
__global__ void __launch_bounds__(256) adaptive_avgpool_fwd_coop_kernel(V4 x, V4 y) {
  const int cg = x.C / 8;
  const int g = threadIdx.x % cg, pl = threadIdx.x / cg, lanes = blockDim.x / cg;
  int r = blockIdx.x;
  const int w = r % y.W;
  r /= y.W;
  const int h = r % y.H, n = r / y.H;
  const int h0 = bin_lo(h, x.H, y.H), h1 = bin_hi(h, x.H, y.H), w0 = bin_lo(w, x.W, y.W), w1 = bin_hi(w, x.W, y.W);
  const int ww_n = w1 - w0, cnt = (h1 - h0) * ww_n;
  float acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = 0.f;
  for (int k = pl; k < cnt && pl < lanes; k += lanes) {
    const F8 v = unpack8(ld16(x.at(n, h0 + k / ww_n, w0 + k % ww_n) + g * 8));
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] += v.v[i];
  }
  if (lanes_reduce(acc, cg)) {
    const float inv = 1.f / (float)cnt;
    F8 o;
#pragma unroll
    for (int i = 0; i < 8; ++i) o.v[i] = acc[i] * inv;
    st16(y.at(n, h, w) + g * 8, pack8(o));
  }
}
__global__ void __launch_bounds__(256) adaptive_avgpool_bwd_kernel(V4 dy, V4 dx) {
  const int cg = dx.C / 8;
  const long long total = (long long)dx.N * dx.H * dx.W * cg;
  PIX_LOOP(total) {
    DECODE(e, cg, dx.W, dx.H)
    float acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = 0.f;
    for (int oh = 0; oh < dy.H; ++oh) {
      const int h0 = bin_lo(oh, dx.H, dy.H), h1 = bin_hi(oh, dx.H, dy.H);
      if (h < h0 || h >= h1) continue;
      for (int ow = 0; ow < dy.W; ++ow) {
        const int w0 = bin_lo(ow, dx.W, dy.W), w1 = bin_hi(ow, dx.W, dy.W);
        if (w < w0 || w >= w1) continue;
        const F8 d = unpack8(ld16(dy.at(n, oh, ow) + g * 8));
        const float inv = 1.f / (float)((h1 - h0) * (w1 - w0));
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i] = fmaf(d.v[i], inv, acc[i]);
      }
    }
    F8 o;
#pragma unroll
    for (int i = 0; i < 8; ++i) o.v[i] = acc[i];
    st16(dx.at(n, h, w) + g * 8, pack8(o));
  }
}

// y = a + b (+ relu)  on arbitrary (same-shape) views
__global__ void __launch_bounds__(256) add_kernel(V4 a, V4 b, V4 y, int relu) {
  const int cg = y.C / 8;
  const long long total = (long long)y.N * y.H * y.W * cg;
  PIX_LOOP(total) {
    DECODE(e, cg, y.W, y.H)
    F8 u = unpack8(ld16_stream(a.at(n, h, w) + g * 8));
    const F8 v = unpack8(ld16_stream(b.at(n, h, w) + g * 8));
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      u.v[i] += v.v[i];
      if (relu) u.v[i] = fmaxf(u.v[i], 0.f);
    }
    st16(y.at(n, h, w) + g * 8, pack8(u));
  }
}

}  // namespace pdac

using namespace pdac;

extern "C" int pdac_maxpool3x3s2_fwd(const PdacView* x, const PdacView* y, uint8_t* idx, void* stream) {
  V4 vx, vy;
  PDAC_TRY(to_v4(x, &vx, "x"));
  PDAC_TRY(to_v4(y, &vy, "y"));
  PDAC_REQUIRE(vy.N == vx.N && vy.C == vx.C && vy.H == (vx.H - 1) / 2 + 1 && vy.W == (vx.W - 1) / 2 + 1,
               "maxpool: output shape mismatch");
  PDAC_REQUIRE(!idx || ((uintptr_t)idx & 7) == 0, "maxpool: idx must be 8-byte aligned");
  const long long total = (long long)vy.N * vy.H * vy.W * (vx.C / 8);
  if (!total) return PDAC_OK;
  maxpool_fwd_kernel<<<blocks_for(total), 256, 0, (cudaStream_t)stream>>>(vx, vy, idx);
  PDAC_LAUNCH_CHECK("maxpool_fwd");
  return PDAC_OK;
}
extern "C" int pdac_maxpool3x3s2_bwd(const PdacView* dy, const uint8_t* idx, const PdacView* addend, const PdacView* dx,
                                     void* stream) {
  V4 vdy, vdx, va = {nullptr, 0, 0, 0, 0, 0, 0, 0};
  PDAC_TRY(to_v4(dy, &vdy, "dy"));
  PDAC_TRY(to_v4(dx, &vdx, "dx"));
  if (addend) {
    PDAC_TRY(to_v4(addend, &va, "addend"));
    PDAC_REQUIRE(va.N == vdx.N && va.H == vdx.H && va.W == vdx.W && va.C == vdx.C, "maxpool bwd: addend shape mismatch");
  }
  PDAC_REQUIRE(idx, "maxpool bwd: null idx");
  PDAC_REQUIRE(vdy.N == vdx.N && vdy.C == vdx.C && vdy.H == (vdx.H - 1) / 2 + 1 && vdy.W == (vdx.W - 1) / 2 + 1,
               "maxpool bwd: shape mismatch");
  const long long total = (long long)vdx.N * ((vdx.H + 1) / 2) * ((vdx.W + 1) / 2) * (vdx.C / 8);
  if (!total) return PDAC_OK;
  maxpool_bwd_kernel<<<blocks_for(total), 256, 0, (cudaStream_t)stream>>>(vdy, idx, va, vdx);
  PDAC_LAUNCH_CHECK("maxpool_bwd");
  return PDAC_OK;
}

extern "C" int pdac_upsample_nearest2x_fwd(const PdacView* x, const PdacView* y, void* stream) {
  V4 vx, vy;
  PDAC_TRY(to_v4(x, &vx, "x"));
  PDAC_TRY(to_v4(y, &vy, "y"));
  PDAC_REQUIRE(vy.N == vx.N && vy.C == vx.C && vy.H == 2 * vx.H && vy.W == 2 * vx.W, "nearest2x: shape mismatch");
  const long long total = (long long)vx.N * vx.H * vx.W * (vx.C / 8);
  if (!total) return PDAC_OK;
  up2_fwd_kernel<<<blocks_for(total), 256, 0, (cudaStream_t)stream>>>(vx, vy);
  PDAC_LAUNCH_CHECK("up2_fwd");
  return PDAC_OK;
}
extern "C" int pdac_upsample_nearest2x_bwd(const PdacView* dy, const PdacView* addend, const PdacView* dx,
                                           void* stream) {
  V4 vdy, vdx, va = {nullptr, 0, 0, 0, 0, 0, 0, 0};
  PDAC_TRY(to_v4(dy, &vdy, "dy"));
  PDAC_TRY(to_v4(dx, &vdx, "dx"));
  if (addend) PDAC_TRY(to_v4(addend, &va, "addend"));
  PDAC_REQUIRE(vdy.N == vdx.N && vdy.C == vdx.C && vdy.H == 2 * vdx.H && vdy.W == 2 * vdx.W,
               "nearest2x bwd: shape mismatch");
  const long long total = (long long)vdx.N * vdx.H * vdx.W * (vdx.C / 8);
  if (!total) return PDAC_OK;
  up2_bwd_kernel<<<blocks_for(total), 256, 0, (cudaStream_t)stream>>>(vdy, va, vdx);
  PDAC_LAUNCH_CHECK("up2_bwd");
  return PDAC_OK;
}

extern "C" int pdac_bilinear_fwd(const PdacView* x, int align_corners, const PdacView* y, void* stream) {
  V4 vx, vy;
  PDAC_TRY(to_v4(x, &vx, "x"));
  PDAC_TRY(to_v4(y, &vy, "y"));
  PDAC_REQUIRE(vy.N == vx.N && vy.C == vx.C, "bilinear: shape mismatch");
  const long long total = (long long)vy.N * vy.H * vy.W * (vx.C / 8);
  if (!total) return PDAC_OK;
  bilinear_fwd_kernel<<<blocks_for(total), 256, 0, (cudaStream_t)stream>>>(vx, vy, make_lin(vx.H, vy.H, align_corners),
                                                                          make_lin(vx.W, vy.W, align_corners));
  PDAC_LAUNCH_CHECK("bilinear_fwd");
  return PDAC_OK;
}
extern "C" int pdac_bilinear_bwd(const PdacView* dy, int align_corners, const PdacView* dx, void* stream) {
  V4 vdy, vdx;
  PDAC_TRY(to_v4(dy, &vdy, "dy"));
  PDAC_TRY(to_v4(dx, &vdx, "dx"));
  PDAC_REQUIRE(vdy.N == vdx.N && vdy.C == vdx.C, "bilinear bwd: shape mismatch");
  const long long total = (long long)vdx.N * vdx.H * vdx.W * (vdx.C / 8);
  if (!total) return PDAC_OK;
  if ((long long)vdy.H * vdy.W >= 64LL * vdx.H * vdx.W && vdx.C / 8 <= 128) {
    bilinear_bwd_coop_kernel<<<vdx.N * vdx.H * vdx.W, 256, 0, (cudaStream_t)stream>>>(
        vdy, vdx, make_lin(vdx.H, vdy.H, align_corners), make_lin(vdx.W, vdy.W, align_corners));
    PDAC_LAUNCH_CHECK("bilinear_bwd_coop");
    return PDAC_OK;
  }
  bilinear_bwd_kernel<<<blocks_for(total), 256, 0, (cudaStream_t)stream>>>(
      vdy, vdx, make_lin(vdx.H, vdy.H, align_corners), make_lin(vdx.W, vdy.W, align_corners));
  PDAC_LAUNCH_CHECK("bilinear_bwd");
  return PDAC_OK;
}
extern "C" int pdac_bilinear_head_fwd(const PdacView* x, int C_used, int align_corners, float* out, int Ho, int Wo,
                                      void* stream) {
  V4 vx;
  PDAC_TRY(to_v4(x, &vx, "x"));
  PDAC_REQUIRE(out && C_used >= 1 && C_used <= vx.C && Ho > 0 && Wo > 0, "bilinear head: bad arguments");
  const long long total = (long long)vx.N * C_used * Ho * Wo;
  if (!total) return PDAC_OK;
  bilinear_head_fwd_kernel<<<blocks_for(total), 256, 0, (cudaStream_t)stream>>>(
      vx, C_used, out, Ho, Wo, make_lin(vx.H, Ho, align_corners), make_lin(vx.W, Wo, align_corners));
  PDAC_LAUNCH_CHECK("bilinear_head_fwd");
  return PDAC_OK;
}
extern "C" int pdac_bilinear_head_bwd(const float* g, int C_used, int Ho, int Wo, int align_corners, const PdacView* dx,
                                      void* stream) {
  V4 vdx;
  PDAC_TRY(to_v4(dx, &vdx, "dx"));
  PDAC_REQUIRE(g && C_used >= 1 && C_used <= vdx.C, "bilinear head bwd: bad arguments");
  const long long total = (long long)vdx.N * vdx.H * vdx.W;
  if (!total) return PDAC_OK;
  bilinear_head_bwd_kernel<<<blocks_for(total), 256, 0, (cudaStream_t)stream>>>(
      g, C_used, Ho, Wo, vdx, make_lin(vdx.H, Ho, align_corners), make_lin(vdx.W, Wo, align_corners));
  PDAC_LAUNCH_CHECK("bilinear_head_bwd");
  return PDAC_OK;
}

extern "C" int pdac_adaptive_avgpool_fwd(const PdacView* x, const PdacView* y, void* stream) {
  V4 vx, vy;
  PDAC_TRY(to_v4(x, &vx, "x"));
  PDAC_TRY(to_v4(y, &vy, "y"));
  PDAC_REQUIRE(vy.N == vx.N && vy.C == vx.C, "adaptive_avgpool: shape mismatch");
  const long long total = (long long)vy.N * vy.H * vy.W * (vx.C / 8);
  if (!total) return PDAC_OK;
  if ((long long)vx.H * vx.W >= 64LL * vy.H * vy.W && vx.C / 8 <= 128) {
    adaptive_avgpool_fwd_coop_kernel<<<vy.N * vy.H * vy.W, 256, 0, (cudaStream_t)stream>>>(vx, vy);
    PDAC_LAUNCH_CHECK("adaptive_avgpool_fwd_coop");
    return PDAC_OK;
  }
  adaptive_avgpool_fwd_kernel<<<blocks_for(total), 256, 0, (cudaStream_t)stream>>>(vx, vy);
  PDAC_LAUNCH_CHECK("adaptive_avgpool_fwd");
  return PDAC_OK;
}
extern "C" int pdac_adaptive_avgpool_bwd(const PdacView* dy, const PdacView* dx, void* stream) {
  V4 vdy, vdx;
  PDAC_TRY(to_v4(dy, &vdy, "dy"));
  PDAC_TRY(to_v4(dx, &vdx, "dx"));
  PDAC_REQUIRE(vdy.N == vdx.N && vdy.C == vdx.C, "adaptive_avgpool bwd: shape mismatch");
  const long long total = (long long)vdx.N * vdx.H * vdx.W * (vdx.C / 8);
  if (!total) return PDAC_OK;
  adaptive_avgpool_bwd_kernel<<<blocks_for(total), 256, 0, (cudaStream_t)stream>>>(vdy, vdx);
  PDAC_LAUNCH_CHECK("adaptive_avgpool_bwd");
  return PDAC_OK;
}

extern "C" int pdac_add(const PdacView* a, const PdacView* b, int relu, const PdacView* y, void* stream) {
  V4 va, vb, vy;
  PDAC_TRY(to_v4(a, &va, "a"));
  PDAC_TRY(to_v4(b, &vb, "b"));
  PDAC_TRY(to_v4(y, &vy, "y"));
  PDAC_REQUIRE(va.N == vy.N && va.H == vy.H && va.W == vy.W && va.C == vy.C && vb.N == vy.N && vb.H == vy.H &&
                   vb.W == vy.W && vb.C == vy.C,
               "add: shape mismatch");
  const long long total = (long long)vy.N * vy.H * vy.W * (vy.C / 8);
  if (!total) return PDAC_OK;
  add_kernel<<<blocks_for(total), 256, 0, (cudaStream_t)stream>>>(va, vb, vy, relu);
  PDAC_LAUNCH_CHECK("add");
  return PDAC_OK;
}
