// Layout conversion at the module boundary (NCHW fp32 <-> NHWC bf16), the stem's space-to-depth input preparation
// (fp32 NCHW tensors as the reference feeds them, train_worker.py:70, or uint8 HWC patches as dataset.py:24-30 decodes
// them, with the Normalize / flip / rot90 of train_main.py:70-88 fused in), and weight packing
// (state_dict fp32 OIHW -> bf16 K-major GEMM matrices, and the stem's 7x7 s2 -> 4x4 s1 x 16ch rewrite).
#include "common.h"
#include "vec.cuh"
#include <string.h>

namespace pdac {

// ---------------------------------------------------------------------------------------------- NCHW f32 -> NHWC bf16
__global__ void nchw_to_nhwc_kernel(const float* __restrict__ x, int C, long long HW, long long NHW, bf16* out,
                                    int Cout, long long ps) {
  for (long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x; p < NHW; p += (long long)gridDim.x * blockDim.x) {
    const long long n = p / HW, hw = p - n * HW;
    const float* xp = x + n * C * HW + hw;
    bf16* op = out + p * ps;
    for (int c0 = 0; c0 < Cout; c0 += 8) {
      F8 f;
#pragma unroll
      for (int i = 0; i < 8; ++i) f.v[i] = (c0 + i < C) ? __ldg(xp + (long long)(c0 + i) * HW) : 0.f;
      if (Cout - c0 >= 8) {
        st16(op + c0, pack8(f));
      } else {
        for (int i = 0; i < Cout - c0; ++i) op[c0 + i] = __float2bfloat16(f.v[i]);
      }
    }
  }
}

__global__ void nhwc_to_nchw_kernel(const bf16* __restrict__ x, long long ps, int C, long long HW, long long NHW,
                                    float* __restrict__ out) {
  for (long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x; p < NHW; p += (long long)gridDim.x * blockDim.x) {
    const long long n = p / HW, hw = p - n * HW;
    const bf16* xp = x + p * ps;
    float* op = out + n * C * HW + hw;
    for (int c = 0; c < C; ++c) op[(long long)c * HW] = __bfloat162float(xp[c]);
  }
}

// ---------------------------------------------------------------------------------------------- stem input prep
// out[n, i, j, (p*2+q)*3 + c] = x[n, c, 2i+p, 2j+q]; channels 12..15 are zero.
__global__ void stem_prep_f32_kernel(const float* __restrict__ x, int N, int H, int W, bf16* __restrict__ out) {
  const int Ho = H / 2, Wo = W / 2;
  const long long total = (long long)N * Ho * Wo;
  const long long HW = (long long)H * W;
  for (long long o = blockIdx.x * (long long)blockDim.x + threadIdx.x; o < total; o += (long long)gridDim.x * blockDim.x) {
    const int j = (int)(o % Wo);
    const long long r = o / Wo;
    const int i = (int)(r % Ho);
    const long long n = r / Ho;
    float v[16];
#pragma unroll
    for (int k = 12; k < 16; ++k) v[k] = 0.f;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float* xp = x + (n * 3 + c) * HW + (long long)(2 * i) * W + 2 * j;
      const float2 a = __ldg(reinterpret_cast<const float2*>(xp));
      const float2 b = __ldg(reinterpret_cast<const float2*>(xp + W));
      v[0 * 3 + c] = a.x;
      v[1 * 3 + c] = a.y;
      v[2 * 3 + c] = b.x;
      v[3 * 3 + c] = b.y;
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

// uint8 HWC patches -> normalised, augmented, space-to-depth bf16. aug[n]: 0 none, 1 horizontal flip, 2 vertical flip,
// 3 rot90 (counter-clockwise, np.rot90 semantics; square patches). The label mask (optional) gets the same geometry.
__device__ __forceinline__ void aug_src(int code, int H, int W, int y, int x, int& sy, int& sx) {
  switch (code) {
    case 1: sy = y; sx = W - 1 - x; break;
    case 2: sy = H - 1 - y; sx = x; break;
    case 3: sy = x; sx = W - 1 - y; break;  // out[y][x] = in[x][W-1-y]
    default: sy = y; sx = x; break;
  }
}
struct Norm3 {
  float scale[3], shift[3];
};
__global__ void stem_prep_u8_kernel(const uint8_t* __restrict__ img, const uint8_t* __restrict__ mask_in,
                                    const int32_t* __restrict__ aug, int N, int H, int W, Norm3 nm,
                                    bf16* __restrict__ out, uint8_t* __restrict__ mask_out) {
  const int Ho = H / 2, Wo = W / 2;
  const long long total = (long long)N * Ho * Wo;
  for (long long o = blockIdx.x * (long long)blockDim.x + threadIdx.x; o < total; o += (long long)gridDim.x * blockDim.x) {
    const int j = (int)(o % Wo);
    const long long r = o / Wo;
    const int i = (int)(r % Ho);
    const long long n = r / Ho;
    const int code = aug ? aug[n] : 0;
    const uint8_t* ip = img + n * (long long)H * W * 3;
    float v[16];
#pragma unroll
    for (int k = 12; k < 16; ++k) v[k] = 0.f;
#pragma unroll
    for (int p = 0; p < 2; ++p)
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        int sy, sx;
        aug_src(code, H, W, 2 * i + p, 2 * j + q, sy, sx);
        const uint8_t* px = ip + ((long long)sy * W + sx) * 3;
#pragma unroll
        for (int c = 0; c < 3; ++c) v[(p * 2 + q) * 3 + c] = fmaf((float)px[c], nm.scale[c], nm.shift[c]);
        if (mask_in) mask_out[n * (long long)H * W + (long long)(2 * i + p) * W + 2 * j + q] =
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

// ---------------------------------------------------------------------------------------------- weight packing
struct TapPerm {
  int32_t pos[PDAC_MAX_TAPS];
};
// w: fp32 [O][I][RS]. wf[(t*wrows_f + o)][i] (K = I, kf columns); wt[(perm[t]*wrows_t + i)][o] (K = O padded to kt).
__global__ void pack_conv_weight_kernel(const float* __restrict__ w, int O, int I, int RS, int wrows_f, int kf,
                                        bf16* __restrict__ wf, int wrows_t, int kt, bf16* __restrict__ wt, TapPerm perm) {
  const long long nf = wf ? (long long)RS * wrows_f * kf : 0;
  const long long nt = wt ? (long long)RS * wrows_t * kt : 0;
  for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < nf + nt; e += (long long)gridDim.x * blockDim.x) {
    if (e < nf) {
      const int i = (int)(e % kf);
      const long long r = e / kf;
      const int o = (int)(r % wrows_f);
      const int t = (int)(r / wrows_f);
      const float v = (o < O && i < I) ? w[((long long)o * I + i) * RS + t] : 0.f;
      wf[e] = __float2bfloat16(v);
    } else {
      const long long e2 = e - nf;
      const int o = (int)(e2 % kt);
      const long long r = e2 / kt;
      const int i = (int)(r % wrows_t);
      const int tp = (int)(r / wrows_t);
      int t = 0;
      for (int k = 0; k < RS; ++k)
        if (perm.pos[k] == tp) t = k;
      const float v = (o < O && i < I) ? w[((long long)o * I + i) * RS + t] : 0.f;
      wt[e2] = __float2bfloat16(v);
    }
  }
}

// Batched form: one launch re-packs up to kPackBatch weight tensors (blockIdx.y = tensor): after an optimizer step every
// conv weight of the network is re-packed at once instead of by ~30 tiny launches inside the next forward.
static constexpr int kPackBatch = 24;
struct PackItem {
  const float* w;
  bf16* wf;
  bf16* wt;
  int O, I, RS, wrows_f, kf, wrows_t, kt;
  int8_t pos[PDAC_MAX_TAPS];
};
struct PackBatch {
  PackItem it[kPackBatch];
};
// One thread per (row, k) pair of a packed matrix: it reads the R*S taps of its (o, i) weight, which are contiguous in
// the state_dict layout, and writes one element into each tap plane. (The first version walked the packed elements and
// gathered every tap plane separately: nine passes over the fp32 weights through L2 with 36-byte strides, 0.2 ms per
// step for 57 MB of parameters.)
__global__ void __launch_bounds__(256) pack_conv_weight_batched_kernel(const __grid_constant__ PackBatch b) {
  const PackItem& d = b.it[blockIdx.y];
  const long long nf = d.wf ? (long long)d.wrows_f * d.kf : 0;
  const long long nt = d.wt ? (long long)d.wrows_t * d.kt : 0;
  const long long plane_f = (long long)d.wrows_f * d.kf, plane_t = (long long)d.wrows_t * d.kt;
  for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < nf + nt; e += (long long)gridDim.x * blockDim.x) {
    if (e < nf) {  // forward matrix [tap][o][i]
      const int i = (int)(e % d.kf), o = (int)(e / d.kf);
      const bool ok = o < d.O && i < d.I;
      const float* src = d.w + ((long long)o * d.I + i) * d.RS;
      for (int t = 0; t < d.RS; ++t) d.wf[t * plane_f + e] = __float2bfloat16(ok ? src[t] : 0.f);
    } else {  // data-gradient matrix [tap position][i][o]
      const long long e2 = e - nf;
      const int o = (int)(e2 % d.kt), i = (int)(e2 / d.kt);
      const bool ok = o < d.O && i < d.I;
      const float* src = d.w + ((long long)o * d.I + i) * d.RS;
      for (int t = 0; t < d.RS; ++t) d.wt[d.pos[t] * plane_t + e2] = __float2bfloat16(ok ? src[t] : 0.f);
    }
  }
}

// stem: w fp32 [64][3][7][7] -> [16 taps (a,b)][O][16] with channel (p*2+q)*3+c <- w[o][c][2a+p-1][2b+q-1]
__global__ void pack_stem_weight_kernel(const float* __restrict__ w, int O, bf16* __restrict__ wp) {
  const int total = 16 * O * 16;
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < total; e += gridDim.x * blockDim.x) {
    const int ch = e % 16, o = (e / 16) % O, t = e / (16 * O);
    float v = 0.f;
    if (ch < 12) {
      const int c = ch % 3, pq = ch / 3, p = pq >> 1, q = pq & 1;
      const int a = t >> 2, b = t & 3;
      const int r = 2 * a + p - 1, s = 2 * b + q - 1;
      if (r >= 0 && s >= 0) v = w[((o * 3 + c) * 7 + r) * 7 + s];
    }
    wp[e] = __float2bfloat16(v);
  }
}
// inverse mapping for the stem weight gradient: dwp fp32 [16][O][16] -> dw fp32 [O][3][7][7]
__global__ void unpack_stem_wgrad_kernel(const float* __restrict__ dwp, int O, float* __restrict__ dw) {
  const int total = O * 3 * 49;
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < total; e += gridDim.x * blockDim.x) {
    const int s = e % 7, r = (e / 7) % 7, c = (e / 49) % 3, o = e / 147;
    const int rp = r + 1, sp = s + 1;
    const int a = rp >> 1, p = rp & 1, b = sp >> 1, q = sp & 1;
    dw[e] = dwp[((a * 4 + b) * O + o) * 16 + (p * 2 + q) * 3 + c];
  }
}

static int ew_blocks(long long n, int threads) {
  long long b = (n + threads - 1) / threads;
  long long cap = (long long)num_sms() * 16;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}

}  // namespace pdac

using namespace pdac;

static int check_flat(const PdacView* v, const char* what) {
  PDAC_REQUIRE(v && v->ptr, "%s: null view", what);
  PDAC_REQUIRE((v->H == 1 || v->sH == (int64_t)v->W * v->sW) && (v->N == 1 || v->sN == (int64_t)v->H * v->W * v->sW),
               "%s: view must be pixel-contiguous", what);
  return PDAC_OK;
}

extern "C" int pdac_nchw_f32_to_nhwc_bf16(const float* x, int N, int C, int H, int W, const PdacView* out,
                                          void* stream) {
  PDAC_REQUIRE(x, "null input");
  PDAC_TRY(check_flat(out, "out"));
  PDAC_REQUIRE(out->N == N && out->H == H && out->W == W && out->C >= C, "nchw->nhwc: shape mismatch");
  if (out->C >= 8) PDAC_REQUIRE(((uintptr_t)out->ptr & 15) == 0 && out->sW % 8 == 0, "out must be 16-byte aligned");
  const long long nhw = (long long)N * H * W;
  if (!nhw) return PDAC_OK;
  nchw_to_nhwc_kernel<<<ew_blocks(nhw, 256), 256, 0, (cudaStream_t)stream>>>(x, C, (long long)H * W, nhw,
                                                                             (bf16*)out->ptr, out->C, out->sW);
  PDAC_LAUNCH_CHECK("nchw_to_nhwc");
  return PDAC_OK;
}

extern "C" int pdac_nhwc_bf16_to_nchw_f32(const PdacView* x, int C_used, float* out, void* stream) {
  PDAC_TRY(check_flat(x, "x"));
  PDAC_REQUIRE(out && C_used >= 1 && C_used <= x->C, "nhwc->nchw: bad arguments");
  const long long nhw = (long long)x->N * x->H * x->W;
  if (!nhw) return PDAC_OK;
  nhwc_to_nchw_kernel<<<ew_blocks(nhw, 256), 256, 0, (cudaStream_t)stream>>>((const bf16*)x->ptr, x->sW, C_used,
                                                                             (long long)x->H * x->W, nhw, out);
  PDAC_LAUNCH_CHECK("nhwc_to_nchw");
  return PDAC_OK;
}

extern "C" int pdac_stem_prep_f32(const float* x, int N, int H, int W, void* out, void* stream) {
  PDAC_REQUIRE(x && out, "null argument");
  PDAC_REQUIRE(H % 2 == 0 && W % 2 == 0 && ((uintptr_t)x & 7) == 0, "stem_prep: H, W must be even, x 8-byte aligned");
  const long long total = (long long)N * (H / 2) * (W / 2);
  if (!total) return PDAC_OK;
  stem_prep_f32_kernel<<<ew_blocks(total, 256), 256, 0, (cudaStream_t)stream>>>(x, N, H, W, (bf16*)out);
  PDAC_LAUNCH_CHECK("stem_prep_f32");
  return PDAC_OK;
}

extern "C" int pdac_stem_prep_u8(const uint8_t* img, const uint8_t* mask_in, const int32_t* aug, int N, int H, int W,
                                 const float* mean, const float* stdv, void* out, uint8_t* mask_out, void* stream) {
  PDAC_REQUIRE(img && out && mean && stdv, "null argument");
  PDAC_REQUIRE(H % 2 == 0 && W % 2 == 0, "stem_prep: H, W must be even");
  PDAC_REQUIRE(!mask_in || mask_out, "mask_in without mask_out");
  PDAC_REQUIRE(!aug || H == W, "rot90 augmentation needs square patches");
  Norm3 nm;
  for (int c = 0; c < 3; ++c) {
    nm.scale[c] = 1.f / (255.f * stdv[c]);
    nm.shift[c] = -mean[c] / stdv[c];
  }
  const long long total = (long long)N * (H / 2) * (W / 2);
  if (!total) return PDAC_OK;
  stem_prep_u8_kernel<<<ew_blocks(total, 256), 256, 0, (cudaStream_t)stream>>>(img, mask_in, aug, N, H, W, nm,
                                                                              (bf16*)out, mask_out);
  PDAC_LAUNCH_CHECK("stem_prep_u8");
  return PDAC_OK;
}

extern "C" int pdac_pack_conv_weight(const float* w, int O, int I, int R, int S, int wrows_f, int kf, void* wp_f,
                                     int wrows_t, int kt, void* wp_t, int s2_pad, void* stream) {
  PDAC_REQUIRE(w && (wp_f || wp_t), "null argument");
  const int RS = R * S;
  PDAC_REQUIRE(RS >= 1 && RS <= PDAC_MAX_TAPS, "kernel %dx%d unsupported", R, S);
  PDAC_REQUIRE(!wp_f || (wrows_f >= O && kf >= I), "forward packing: wrows/k too small");
  PDAC_REQUIRE(!wp_t || (wrows_t >= I && kt >= O), "dgrad packing: wrows/k too small");
  TapPerm perm;
  for (int t = 0; t < PDAC_MAX_TAPS; ++t) perm.pos[t] = t;
  if (s2_pad >= 0) {
    // stride-2 data gradient: taps grouped by output parity class (ph,pw), see pdac_conv2d_dgrad
    int pos = 0;
    for (int ph = 0; ph < 2; ++ph)
      for (int pw = 0; pw < 2; ++pw)
        for (int r = 0; r < R; ++r) {
          if ((r - s2_pad - ph) & 1) continue;
          for (int s = 0; s < S; ++s) {
            if ((s - s2_pad - pw) & 1) continue;
            perm.pos[r * S + s] = pos++;
          }
        }
  }
  const long long n = (wp_f ? (long long)RS * wrows_f * kf : 0) + (wp_t ? (long long)RS * wrows_t * kt : 0);
  pack_conv_weight_kernel<<<ew_blocks(n, 256), 256, 0, (cudaStream_t)stream>>>(w, O, I, RS, wrows_f, kf, (bf16*)wp_f,
                                                                               wrows_t, kt, (bf16*)wp_t, perm);
  PDAC_LAUNCH_CHECK("pack_conv_weight");
  return PDAC_OK;
}

static void s2_tap_positions(int R, int S, int s2_pad, int8_t* pos) {
  for (int t = 0; t < PDAC_MAX_TAPS; ++t) pos[t] = (int8_t)t;
  if (s2_pad < 0) return;
  int p = 0;
  for (int ph = 0; ph < 2; ++ph)
    for (int pw = 0; pw < 2; ++pw)
      for (int r = 0; r < R; ++r) {
        if ((r - s2_pad - ph) & 1) continue;
        for (int s = 0; s < S; ++s) {
          if ((s - s2_pad - pw) & 1) continue;
          pos[r * S + s] = (int8_t)p++;
        }
      }
}

extern "C" int pdac_pack_conv_weights_batched(const PdacPackDesc* descs, int n, void* stream) {
  PDAC_REQUIRE(descs && n >= 0, "pack batch: null argument");
  for (int base = 0; base < n; base += kPackBatch) {
    PackBatch b;
    memset(&b, 0, sizeof(b));
    const int m = n - base < kPackBatch ? n - base : kPackBatch;
    for (int j = 0; j < m; ++j) {
      const PdacPackDesc& d = descs[base + j];
      PDAC_REQUIRE(d.w && (d.wp_f || d.wp_t), "pack batch %d: null pointer", base + j);
      const int RS = d.R * d.S;
      PDAC_REQUIRE(RS >= 1 && RS <= PDAC_MAX_TAPS, "pack batch %d: kernel %dx%d unsupported", base + j, d.R, d.S);
      PDAC_REQUIRE(!d.wp_f || (d.wrows_f >= d.O && d.kf >= d.I), "pack batch %d: forward wrows/k too small", base + j);
      PDAC_REQUIRE(!d.wp_t || (d.wrows_t >= d.I && d.kt >= d.O), "pack batch %d: dgrad wrows/k too small", base + j);
      PackItem& it = b.it[j];
      it.w = d.w;
      it.wf = (bf16*)d.wp_f;
      it.wt = (bf16*)d.wp_t;
      it.O = d.O;
      it.I = d.I;
      it.RS = RS;
      it.wrows_f = d.wrows_f;
      it.kf = d.kf;
      it.wrows_t = d.wrows_t;
      it.kt = d.kt;
      s2_tap_positions(d.R, d.S, d.s2_pad, it.pos);
    }
    pack_conv_weight_batched_kernel<<<dim3(2 * num_sms(), m), 256, 0, (cudaStream_t)stream>>>(b);
    PDAC_LAUNCH_CHECK("pack_conv_weight_batched");
  }
  return PDAC_OK;
}

extern "C" int pdac_pack_stem_weight(const float* w, int O, void* wp, void* stream) {
  PDAC_REQUIRE(w && wp && O > 0, "null argument");
  pack_stem_weight_kernel<<<ew_blocks(16 * O * 16, 256), 256, 0, (cudaStream_t)stream>>>(w, O, (bf16*)wp);
  PDAC_LAUNCH_CHECK("pack_stem_weight");
  return PDAC_OK;
}

extern "C" int pdac_unpack_stem_wgrad(const float* dwp, int O, float* dw, void* stream) {
  PDAC_REQUIRE(dwp && dw && O > 0, "null argument");
  unpack_stem_wgrad_kernel<<<ew_blocks(O * 147, 256), 256, 0, (cudaStream_t)stream>>>(dwp, O, dw);
  PDAC_LAUNCH_CHECK("unpack_stem_wgrad");
  return PDAC_OK;
}
