// GroupNorm (forward / backward), Dropout / Dropout2d and depthwise 3x3 (dilated) convolution on NHWC bf16: the layers
// smp.FPN (Conv3x3GNReLU, Dropout2d), smp.PSPNet (Dropout2d) and smp.DeepLabV3Plus (SeparableConv2d, Dropout) add to
// the ResNet18 path (train_worker.py:152-182). All HBM-bound: one thread owns 8 consecutive
// channels (16 bytes) of a pixel, fixed across its grid-stride loop over the pixels of ONE image (blockIdx.y = image),
// so per-channel / per-(image, group) constants live in registers.
#include "common.h"
#include "vec.cuh"
#include <string.h>

namespace pdac {

struct ImgView {
  bf16* ptr;
  long long sN, ps;  // image stride, pixel stride (elements)
  int npix, C;       // pixels per image
};

static int img_view(const PdacView* v, ImgView* f, const char* what) {
  PDAC_REQUIRE(v && v->ptr, "%s: null view", what);
  PDAC_REQUIRE(v->C % 8 == 0, "%s: channels (%d) must be a multiple of 8", what, v->C);
  PDAC_REQUIRE(v->H == 1 || v->sH == (int64_t)v->W * v->sW, "%s: view must be pixel-contiguous inside an image", what);
  PDAC_REQUIRE(((uintptr_t)v->ptr & 15) == 0 && v->sW % 8 == 0 && v->sN % 8 == 0, "%s: view must be 16-byte aligned", what);
  f->ptr = (bf16*)v->ptr;
  f->sN = v->sN;
  f->ps = v->sW;
  f->npix = v->H * v->W;
  f->C = v->C;
  return PDAC_OK;
}

struct ImgGrid {
  dim3 grid;
  int cg, ppb;
};
static ImgGrid img_grid(int npix, int C, int N) {
  ImgGrid g;
  g.cg = C / 8;
  g.ppb = 256 / g.cg;
  if (g.ppb < 1) g.ppb = 1;
  int need = (npix + g.ppb - 1) / g.ppb;
  int cap = (num_sms() * 8 + N - 1) / N;
  if (cap < 1) cap = 1;
  g.grid = dim3(need < cap ? need : cap, N, 1);
  return g;
}
#define IMG_SETUP(cg_, ppb_)                                   \
  const int g = threadIdx.x % (cg_), pl = threadIdx.x / (cg_); \
  const bool active = pl < (ppb_);                             \
  const int n = blockIdx.y;                                    \
  const int pstep = gridDim.x * (ppb_);                        \
  const int pfirst = blockIdx.x * (ppb_) + pl;

// ------------------------------------------------------------------------------------------------ GroupNorm
// stats[n][grp][0..1] += sum, sum of squares over the group's channels and the image's pixels
__global__ void __launch_bounds__(256) gn_stats_kernel(ImgView x, int cpg, float* __restrict__ stats,
                                                       float* __restrict__ rows, int G, int cg, int ppb) {
  extern __shared__ float s_grp[];  // [G][2] (deterministic path)
  IMG_SETUP(cg, ppb)
  float s[8], q[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) s[i] = q[i] = 0.f;
  if (active && g < cg) {
    const bf16* base = x.ptr + (long long)n * x.sN + g * 8;
    for (int p = pfirst; p < x.npix; p += pstep) {
      const F8 a = unpack8(ld16(base + (long long)p * x.ps));
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        s[i] += a.v[i];
        q[i] = fmaf(a.v[i], a.v[i], q[i]);
      }
    }
    float* st = stats + (long long)n * G * 2;
    if (rows) {
      // handled below (all threads of the block take part)
    } else if (cpg >= 8) {
      float ss = 0.f, qq = 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        ss += s[i];
        qq += q[i];
      }
      const int grp = (g * 8) / cpg;
      atomicAdd(st + grp * 2, ss);
      atomicAdd(st + grp * 2 + 1, qq);
    } else {
      for (int i0 = 0; i0 < 8; i0 += cpg) {
        float ss = 0.f, qq = 0.f;
        for (int i = i0; i < i0 + cpg; ++i) {
          ss += s[i];
          qq += q[i];
        }
        const int grp = (g * 8 + i0) / cpg;
        atomicAdd(st + grp * 2, ss);
        atomicAdd(st + grp * 2 + 1, qq);
      }
    }
  }
  if (rows) {
    // Reproducible form: the block's per-group sums are built in shared memory in a fixed order (one (pixel lane,
    // thread-within-group) pair adds per round, so no two threads touch a slot in the same round) and stored as this
    // block's row of the [blocks][N][G][2] table, which rows_sum adds over the blocks in index order.
    for (int i = threadIdx.x; i < G * 2; i += blockDim.x) s_grp[i] = 0.f;
    __syncthreads();
    const int tpg = cpg >= 8 ? cpg / 8 : 1;
    for (int round = 0; round < ppb * tpg; ++round) {
      const int qq_ = round / tpg, jj = round - qq_ * tpg;
      if (active && g < cg && pl == qq_ && (g % tpg) == jj) {
        if (cpg >= 8) {
          float ss = 0.f, qq = 0.f;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            ss += s[i];
            qq += q[i];
          }
          const int grp = (g * 8) / cpg;
          s_grp[grp * 2] += ss;
          s_grp[grp * 2 + 1] += qq;
        } else {
          for (int i0 = 0; i0 < 8; i0 += cpg) {
            float ss = 0.f, qq = 0.f;
            for (int i = i0; i < i0 + cpg; ++i) {
              ss += s[i];
              qq += q[i];
            }
            const int grp = (g * 8 + i0) / cpg;
            s_grp[grp * 2] += ss;
            s_grp[grp * 2 + 1] += qq;
          }
        }
      }
      __syncthreads();
    }
    float* row = rows + ((size_t)blockIdx.x * gridDim.y + n) * G * 2;
    for (int i = threadIdx.x; i < G * 2; i += blockDim.x) row[i] = s_grp[i];
  }
}

__device__ __forceinline__ void gn_consts(const float* stats, int n, int G, int cpg, int g, float count, float eps,
                                          float (&mean)[8], float (&rstd)[8]) {
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int grp = (g * 8 + i) / cpg;
    const float m = stats[((long long)n * G + grp) * 2] / count;
    float v = stats[((long long)n * G + grp) * 2 + 1] / count - m * m;
    v = v < 0.f ? 0.f : v;
    mean[i] = m;
    rstd[i] = rsqrtf(v + eps);
  }
}

__global__ void __launch_bounds__(256) gn_fwd_kernel(ImgView x, int cpg, const float* __restrict__ stats, int G,
                                                     const float* __restrict__ gamma, const float* __restrict__ beta,
                                                     float eps, int relu, ImgView y, int cg, int ppb) {
  IMG_SETUP(cg, ppb)
  if (!active) return;
  float mean[8], rstd[8], sc[8], sh[8];
  gn_consts(stats, n, G, cpg, g, (float)x.npix * cpg, eps, mean, rstd);
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    sc[i] = rstd[i] * gamma[g * 8 + i];
    sh[i] = beta[g * 8 + i] - mean[i] * sc[i];
  }
  const bf16* xb = x.ptr + (long long)n * x.sN + g * 8;
  bf16* yb = y.ptr + (long long)n * y.sN + g * 8;
  for (int p = pfirst; p < x.npix; p += pstep) {
    F8 a = unpack8(ld16_stream(xb + (long long)p * x.ps));
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      a.v[i] = fmaf(a.v[i], sc[i], sh[i]);
      if (relu) a.v[i] = fmaxf(a.v[i], 0.f);
    }
    st16(yb + (long long)p * y.ps, pack8(a));
  }
}

// gsums[n][grp][0] += sum dz*gamma, [1] += sum dz*gamma*xhat; dgamma[c] += sum dz*xhat, dbeta[c] += sum dz
__global__ void __launch_bounds__(256) gn_bwd_reduce_kernel(ImgView dy, ImgView y, ImgView x, int cpg,
                                                            const float* __restrict__ stats, int G,
                                                            const float* __restrict__ gamma, float eps,
                                                            float* __restrict__ gsums, float* __restrict__ dgamma,
                                                            float* __restrict__ dbeta, float* __restrict__ rows, int cg,
                                                            int ppb) {
  __shared__ float red[256][17];
  IMG_SETUP(cg, ppb)
  float mean[8], rstd[8], a0[8], a1[8];
  if (active) gn_consts(stats, n, G, cpg, g, (float)x.npix * cpg, eps, mean, rstd);
#pragma unroll
  for (int i = 0; i < 8; ++i) a0[i] = a1[i] = 0.f;
  const bf16* db = dy.ptr + (long long)n * dy.sN + g * 8;
  const bf16* yb = y.ptr ? y.ptr + (long long)n * y.sN + g * 8 : nullptr;
  const bf16* xb = x.ptr + (long long)n * x.sN + g * 8;
  for (int p = pfirst; active && p < x.npix; p += pstep) {
    F8 d = unpack8(ld16(db + (long long)p * dy.ps));
    const F8 xv = unpack8(ld16(xb + (long long)p * x.ps));
    if (yb) {
      const F8 yv = unpack8(ld16(yb + (long long)p * y.ps));
#pragma unroll
      for (int i = 0; i < 8; ++i) d.v[i] = yv.v[i] > 0.f ? d.v[i] : 0.f;
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      a0[i] += d.v[i];
      a1[i] = fmaf(d.v[i], (xv.v[i] - mean[i]) * rstd[i], a1[i]);
    }
  }
  // block-level reduction over the pixel lanes first (32 global atomics per THREAD on a few hundred addresses were
  // 1.5 ms per call on FPN's 128-channel maps)
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    red[threadIdx.x][i] = a0[i];
    red[threadIdx.x][8 + i] = a1[i];
  }
  __syncthreads();
  float* gs = gsums + (long long)n * G * 2;
  for (int t = threadIdx.x; t < cg * 16; t += blockDim.x) {
    const int gg = t / 16, slot = t % 16;
    float acc = 0.f;
    for (int q = 0; q < ppb; ++q) acc += red[q * cg + gg][slot];
    const int c = gg * 8 + (slot & 7);
    if (rows) {  // reproducible form: raw per-channel sums of this block; gn_bwd_finalize adds the blocks in index order
      rows[((size_t)blockIdx.x * gridDim.y + n) * 2 * dy.C + (slot < 8 ? 0 : dy.C) + c] = acc;
      continue;
    }
    const float ga = gamma[c];
    if (slot < 8) {
      atomicAdd(gs + (c / cpg) * 2, acc * ga);
      if (dbeta) atomicAdd(dbeta + c, acc);
    } else {
      atomicAdd(gs + (c / cpg) * 2 + 1, acc * ga);
      if (dgamma) atomicAdd(dgamma + c, acc);
    }
  }
}

// rows: [nbx][N][2][C] per-block raw sums (sum dz, sum dz*xhat per channel). One thread per (image, group) builds
// gsums, one per channel accumulates dgamma / dbeta; blocks, images and channels are walked in index order.
__global__ void gn_bwd_finalize_kernel(const float* __restrict__ rows, int nbx, int N, int C, int G,
                                       const float* __restrict__ gamma, float* __restrict__ gsums,
                                       float* __restrict__ dgamma, float* __restrict__ dbeta) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int cpg = C / G;
  if (t < N * G) {
    const int n = t / G, grp = t - n * G;
    float a0 = 0.f, a1 = 0.f;
    for (int bx = 0; bx < nbx; ++bx) {
      const float* r = rows + ((size_t)bx * N + n) * 2 * C;
      for (int c = grp * cpg; c < (grp + 1) * cpg; ++c) {
        a0 += r[c] * gamma[c];
        a1 += r[C + c] * gamma[c];
      }
    }
    gsums[(size_t)t * 2] = a0;
    gsums[(size_t)t * 2 + 1] = a1;
  } else if (t < N * G + C) {
    const int c = t - N * G;
    float b = 0.f, gm = 0.f;
    for (int n = 0; n < N; ++n)
      for (int bx = 0; bx < nbx; ++bx) {
        const float* r = rows + ((size_t)bx * N + n) * 2 * C;
        b += r[c];
        gm += r[C + c];
      }
    if (dbeta) dbeta[c] += b;
    if (dgamma) dgamma[c] += gm;
  }
}

__global__ void __launch_bounds__(256) gn_bwd_apply_kernel(ImgView dy, ImgView y, ImgView x, int cpg,
                                                           const float* __restrict__ stats, int G,
                                                           const float* __restrict__ gamma, float eps,
                                                           const float* __restrict__ gsums, ImgView dx, int cg, int ppb) {
  IMG_SETUP(cg, ppb)
  if (!active) return;
  float mean[8], rstd[8], ga[8], k1[8], k2[8];
  const float count = (float)x.npix * cpg;
  gn_consts(stats, n, G, cpg, g, count, eps, mean, rstd);
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int c = g * 8 + i;
    ga[i] = gamma[c];
    k1[i] = gsums[((long long)n * G + c / cpg) * 2] / count;
    k2[i] = gsums[((long long)n * G + c / cpg) * 2 + 1] / count;
  }
  const bf16* db = dy.ptr + (long long)n * dy.sN + g * 8;
  const bf16* yb = y.ptr ? y.ptr + (long long)n * y.sN + g * 8 : nullptr;
  const bf16* xb = x.ptr + (long long)n * x.sN + g * 8;
  bf16* ob = dx.ptr + (long long)n * dx.sN + g * 8;
  for (int p = pfirst; p < x.npix; p += pstep) {
    F8 d = unpack8(ld16_stream(db + (long long)p * dy.ps));
    const F8 xv = unpack8(ld16_stream(xb + (long long)p * x.ps));
    if (yb) {
      const F8 yv = unpack8(ld16_stream(yb + (long long)p * y.ps));
#pragma unroll
      for (int i = 0; i < 8; ++i) d.v[i] = yv.v[i] > 0.f ? d.v[i] : 0.f;
    }
    F8 o;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float xhat = (xv.v[i] - mean[i]) * rstd[i];
      o.v[i] = rstd[i] * (d.v[i] * ga[i] - k1[i] - xhat * k2[i]);
    }
    st16(ob + (long long)p * dx.ps, pack8(o));
  }
}

// ------------------------------------------------------------------------------------------------ dropout
__device__ __forceinline__ float hash_uniform(unsigned long long seed, unsigned long long idx) {
  unsigned long long z = idx * 0x9E3779B97F4A7C15ull + seed;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z ^= z >> 31;
  return (float)(z >> 40) * (1.0f / 16777216.0f);
}

// y = x * keep / (1 - p); keep is a function of (seed, n, c) for Dropout2d and of (seed, element) for Dropout, so the
// backward pass is the same call on the gradient.
__global__ void __launch_bounds__(256) dropout_kernel(ImgView x, ImgView y, float p, unsigned long long seed,
                                                      const long long* __restrict__ step, int per_channel, int cg,
                                                      int ppb) {
  IMG_SETUP(cg, ppb)
  if (!active) return;
  if (step) seed += (unsigned long long)(*step) * 0xD6E8FEB86659FD93ull;  // a fresh mask per optimizer step, no host value
  const float inv = 1.f / (1.f - p);
  float f[8];
  if (per_channel) {
#pragma unroll
    for (int i = 0; i < 8; ++i)
      f[i] = hash_uniform(seed, (unsigned long long)n * x.C + g * 8 + i) >= p ? inv : 0.f;
  }
  const bf16* xb = x.ptr + (long long)n * x.sN + g * 8;
  bf16* yb = y.ptr + (long long)n * y.sN + g * 8;
  for (int pp = pfirst; pp < x.npix; pp += pstep) {
    F8 a = unpack8(ld16_stream(xb + (long long)pp * x.ps));
    if (!per_channel) {
      const unsigned long long e0 = ((unsigned long long)n * x.npix + pp) * x.C + g * 8;
#pragma unroll
      for (int i = 0; i < 8; ++i) f[i] = hash_uniform(seed, e0 + i) >= p ? inv : 0.f;
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) a.v[i] *= f[i];
    st16(yb + (long long)pp * y.ps, pack8(a));
  }
}

// ------------------------------------------------------------------------------------------------ depthwise 3x3
// y[n,h,w,c] = sum_{r,s} wgt[c][r*3+s] * x[n, h+(r-1)*dil, w+(s-1)*dil, c]   (stride 1, zero padding = dil);
// flip = 1 uses wgt[c][8 - (r*3+s)]: the data gradient.
__global__ void __launch_bounds__(256) dw3x3_kernel(ImgView x, int H, int W, const float* __restrict__ wgt, int dil,
                                                    int flip, ImgView y, int cg, int ppb) {
  IMG_SETUP(cg, ppb)
  if (!active) return;
  float wv[9][8];
#pragma unroll
  for (int t = 0; t < 9; ++t)
#pragma unroll
    for (int i = 0; i < 8; ++i) wv[t][i] = wgt[(long long)(g * 8 + i) * 9 + (flip ? 8 - t : t)];
  const bf16* xb = x.ptr + (long long)n * x.sN + g * 8;
  bf16* yb = y.ptr + (long long)n * y.sN + g * 8;
  for (int p = pfirst; p < x.npix; p += pstep) {
    const int h = p / W, w = p - h * W;
    F8 acc;
#pragma unroll
    for (int i = 0; i < 8; ++i) acc.v[i] = 0.f;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int hh = h + (r - 1) * dil;
      if (hh < 0 || hh >= H) continue;
#pragma unroll
      for (int s = 0; s < 3; ++s) {
        const int ww = w + (s - 1) * dil;
        if (ww < 0 || ww >= W) continue;
        const F8 a = unpack8(ld16(xb + (long long)(hh * W + ww) * x.ps));
#pragma unroll
        for (int i = 0; i < 8; ++i) acc.v[i] = fmaf(a.v[i], wv[r * 3 + s][i], acc.v[i]);
      }
    }
    st16(yb + (long long)p * y.ps, pack8(acc));
  }
}

// Block-level reduction of the per-thread 9 x 8 partial weight gradients in (dynamic) shared memory, then one global
// atomic per (channel, tap) and block: the pixel lanes of a block share their channels, and 72 global atomics per
// THREAD (5 M per call on 2.7 k addresses) were what the 304-channel layer's 3.3 ms consisted of.
__device__ __forceinline__ void dw_flush(float (&acc)[9][8], bool active, int g, int pl, int ppb, int cg,
                                         float* __restrict__ dw, float* __restrict__ rows) {
  extern __shared__ float s_dw[];  // [cg * 8 channels][9 taps]
  for (int i = threadIdx.x; i < cg * 72; i += blockDim.x) s_dw[i] = 0.f;
  __syncthreads();
  if (rows) {
    // reproducible form: one pixel lane adds per round (distinct channel groups -> distinct slots), the block's row of
    // the [blocks][C*9] table is added over the blocks in index order by rows_sum
    for (int q = 0; q < ppb; ++q) {
      if (active && pl == q) {
#pragma unroll
        for (int t = 0; t < 9; ++t)
#pragma unroll
          for (int i = 0; i < 8; ++i) s_dw[(g * 8 + i) * 9 + t] += acc[t][i];
      }
      __syncthreads();
    }
    float* row = rows + ((size_t)blockIdx.y * gridDim.x + blockIdx.x) * cg * 72;
    for (int i = threadIdx.x; i < cg * 72; i += blockDim.x) row[i] = s_dw[i];
    return;
  }
  if (active) {
#pragma unroll
    for (int t = 0; t < 9; ++t)
#pragma unroll
      for (int i = 0; i < 8; ++i) atomicAdd(&s_dw[(g * 8 + i) * 9 + t], acc[t][i]);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < cg * 72; i += blockDim.x) atomicAdd(dw + i, s_dw[i]);
}

// dw[c][r*3+s] += sum_{n,h,w} dy[n,h,w,c] * x[n, h+(r-1)*dil, w+(s-1)*dil, c]
__global__ void __launch_bounds__(256) dw3x3_wgrad_kernel(ImgView x, ImgView dy, int H, int W, int dil,
                                                          float* __restrict__ dw, float* __restrict__ rows, int cg,
                                                          int ppb) {
  IMG_SETUP(cg, ppb)
  float acc[9][8];
#pragma unroll
  for (int t = 0; t < 9; ++t)
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[t][i] = 0.f;
  const bf16* xb = x.ptr + (long long)n * x.sN + g * 8;
  const bf16* db = dy.ptr + (long long)n * dy.sN + g * 8;
  for (int p = pfirst; active && p < x.npix; p += pstep) {
    const int h = p / W, w = p - h * W;
    const F8 d = unpack8(ld16(db + (long long)p * dy.ps));
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int hh = h + (r - 1) * dil;
      if (hh < 0 || hh >= H) continue;
#pragma unroll
      for (int s = 0; s < 3; ++s) {
        const int ww = w + (s - 1) * dil;
        if (ww < 0 || ww >= W) continue;
        const F8 a = unpack8(ld16(xb + (long long)(hh * W + ww) * x.ps));
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[r * 3 + s][i] = fmaf(a.v[i], d.v[i], acc[r * 3 + s][i]);
      }
    }
  }
  dw_flush(acc, active, g, pl, ppb, cg, dw, rows);
}

// dilation 1 (DeepLabV3+ decoder: 304 channels at 128x128, the expensive one): a thread owns one image row and slides a
// 3x3 window of packed columns along it, so a pixel costs 3 new 16-byte loads instead of 9; the pixel lanes of a block
// are consecutive rows, whose halo rows hit L1.
__global__ void __launch_bounds__(256, 2) dw3x3_d1_kernel(ImgView x, int H, int W, const float* __restrict__ wgt, int flip,
                                                          ImgView y, int cg, int ppb) {
  const int g = threadIdx.x % cg, pl = threadIdx.x / cg;
  const int n = blockIdx.y, h = blockIdx.x * ppb + pl;
  if (pl >= ppb || h >= H) return;
  // weights as packed bf16 pairs (36 registers instead of 72: room for two pixels per iteration at 2 blocks / SM);
  // the tensor-core convolutions round their weights to bf16 as well
  uint32_t wp[9][4];
#pragma unroll
  for (int t = 0; t < 9; ++t)
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float* wr = wgt + (long long)(g * 8 + 2 * i) * 9 + (flip ? 8 - t : t);
      const __nv_bfloat162 hh = __floats2bfloat162_rn(wr[0], wr[9]);
      wp[t][i] = *reinterpret_cast<const uint32_t*>(&hh);
    }
  const bf16* xb = x.ptr + (long long)n * x.sN + g * 8;
  bf16* yb = y.ptr + (long long)n * y.sN + g * 8;
  const uint4 zero = make_uint4(0, 0, 0, 0);
  bool rok[3];
  const bf16* rp[3];
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    const int hh = h + r - 1;
    rok[r] = hh >= 0 && hh < H;
    rp[r] = xb + (long long)((rok[r] ? hh : h) * W) * x.ps;
  }
  uint4 c0[3], c1[3], c2[3], c3[3];  // columns w-1, w, w+1, w+2 of rows h-1, h, h+1
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    c0[r] = zero;
    c1[r] = rok[r] ? ld16(rp[r]) : zero;
  }
  auto tap = [&](F8& acc, const uint4& v, int t) {
    const uint32_t u[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      acc.v[2 * i] = fmaf(__uint_as_float(u[i] << 16), __uint_as_float(wp[t][i] << 16), acc.v[2 * i]);
      acc.v[2 * i + 1] =
          fmaf(__uint_as_float(u[i] & 0xffff0000u), __uint_as_float(wp[t][i] & 0xffff0000u), acc.v[2 * i + 1]);
    }
  };
  for (int w = 0; w < W; w += 2) {  // W is even (checked by the launcher)
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      c2[r] = rok[r] ? ld16(rp[r] + (long long)(w + 1) * x.ps) : zero;
      c3[r] = (rok[r] && w + 2 < W) ? ld16(rp[r] + (long long)(w + 2) * x.ps) : zero;
    }
    F8 a0, a1;
#pragma unroll
    for (int i = 0; i < 8; ++i) a0.v[i] = a1.v[i] = 0.f;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      tap(a0, c0[r], r * 3);
      tap(a0, c1[r], r * 3 + 1);
      tap(a0, c2[r], r * 3 + 2);
      tap(a1, c1[r], r * 3);
      tap(a1, c2[r], r * 3 + 1);
      tap(a1, c3[r], r * 3 + 2);
    }
    st16(yb + (long long)(h * W + w) * y.ps, pack8(a0));
    st16(yb + (long long)(h * W + w + 1) * y.ps, pack8(a1));
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      c0[r] = c2[r];
      c1[r] = c3[r];
    }
  }
}

// Weight gradient, dilation 1: a thread owns (channel group, image row, kernel row r) and slides three packed columns of
// input row h + r - 1 against dy row h: 24 accumulators per thread (3 blocks / SM) instead of 72 (1 block / SM).
__global__ void __launch_bounds__(256, 3) dw3x3_d1_wgrad_kernel(ImgView x, ImgView dy, int H, int W,
                                                                float* __restrict__ dw, float* __restrict__ out_rows,
                                                                int cg, int ppb) {
  extern __shared__ float s_dw[];  // [cg * 8 channels][9 taps]
  for (int i = threadIdx.x; i < cg * 72; i += blockDim.x) s_dw[i] = 0.f;
  __syncthreads();
  const int g = threadIdx.x % cg, pl = threadIdx.x / cg;
  const int rows = ppb / 3, rl = pl / 3, r = pl - rl * 3;
  const int n = blockIdx.y;
  const bool lane_ok = rl < rows;
  float acc[3][8];
#pragma unroll
  for (int t = 0; t < 3; ++t)
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[t][i] = 0.f;
  const bf16* xb = x.ptr + (long long)n * x.sN + g * 8;
  const bf16* db = dy.ptr + (long long)n * dy.sN + g * 8;
  const uint4 zero = make_uint4(0, 0, 0, 0);
  for (int h = blockIdx.x * rows + rl; lane_ok && h < H; h += gridDim.x * rows) {
    const int hh = h + r - 1;
    if (hh < 0 || hh >= H) continue;  // zero padding row: contributes nothing
    const bf16* xr = xb + (long long)(hh * W) * x.ps;
    const bf16* dr = db + (long long)(h * W) * dy.ps;
    uint4 c0 = zero, c1 = ld16(xr), c2;
    for (int w = 0; w < W; ++w) {
      const F8 d = unpack8(ld16(dr + (long long)w * dy.ps));
      c2 = (w + 1 < W) ? ld16(xr + (long long)(w + 1) * x.ps) : zero;
      const F8 a0 = unpack8(c0), a1 = unpack8(c1), a2 = unpack8(c2);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        acc[0][i] = fmaf(a0.v[i], d.v[i], acc[0][i]);
        acc[1][i] = fmaf(a1.v[i], d.v[i], acc[1][i]);
        acc[2][i] = fmaf(a2.v[i], d.v[i], acc[2][i]);
      }
      c0 = c1;
      c1 = c2;
    }
  }
  if (out_rows) {  // reproducible form, see dw_flush: one image-row lane adds per round ((g, r) slots are distinct)
    for (int q = 0; q < rows; ++q) {
      if (lane_ok && rl == q) {
#pragma unroll
        for (int t = 0; t < 3; ++t)
#pragma unroll
          for (int i = 0; i < 8; ++i) s_dw[(g * 8 + i) * 9 + r * 3 + t] += acc[t][i];
      }
      __syncthreads();
    }
    float* row = out_rows + ((size_t)blockIdx.y * gridDim.x + blockIdx.x) * cg * 72;
    for (int i = threadIdx.x; i < cg * 72; i += blockDim.x) row[i] = s_dw[i];
    return;
  }
  if (lane_ok) {
#pragma unroll
    for (int t = 0; t < 3; ++t)
#pragma unroll
      for (int i = 0; i < 8; ++i) atomicAdd(&s_dw[(g * 8 + i) * 9 + r * 3 + t], acc[t][i]);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < cg * 72; i += blockDim.x) atomicAdd(dw + i, s_dw[i]);
}

}  // namespace pdac

using namespace pdac;

static int gn_check(const PdacView* x, int G, int* cpg) {
  PDAC_REQUIRE(G > 0 && x->C % G == 0, "GroupNorm: %d channels do not split into %d groups", x->C, G);
  *cpg = x->C / G;
  PDAC_REQUIRE(*cpg <= 8 ? 8 % *cpg == 0 : *cpg % 8 == 0, "GroupNorm: %d channels per group unsupported", *cpg);
  return PDAC_OK;
}

extern "C" int pdac_groupnorm_stats(const PdacView* x, int G, float* stats, float* scratch, void* stream) {
  ImgView xv;
  int cpg = 0;
  PDAC_TRY(img_view(x, &xv, "groupnorm x"));
  PDAC_TRY(gn_check(x, G, &cpg));
  const ImgGrid g = img_grid(xv.npix, x->C, x->N);
  gn_stats_kernel<<<g.grid, 256, scratch ? (size_t)G * 2 * sizeof(float) : 0, (cudaStream_t)stream>>>(
      xv, cpg, stats, scratch, G, g.cg, g.ppb);
  PDAC_LAUNCH_CHECK("gn_stats_kernel");
  if (scratch) PDAC_TRY(rows_sum(scratch, (int)g.grid.x, x->N * G * 2, stats, 0, (cudaStream_t)stream));
  return PDAC_OK;
}

extern "C" int pdac_groupnorm_fwd(const PdacView* x, int G, const float* stats, const float* gamma, const float* beta,
                                  float eps, int relu, const PdacView* y, void* stream) {
  ImgView xv, yv;
  int cpg = 0;
  PDAC_TRY(img_view(x, &xv, "groupnorm x"));
  PDAC_TRY(img_view(y, &yv, "groupnorm y"));
  PDAC_TRY(gn_check(x, G, &cpg));
  const ImgGrid g = img_grid(xv.npix, x->C, x->N);
  gn_fwd_kernel<<<g.grid, 256, 0, (cudaStream_t)stream>>>(xv, cpg, stats, G, gamma, beta, eps, relu, yv, g.cg, g.ppb);
  PDAC_LAUNCH_CHECK("gn_fwd_kernel");
  return PDAC_OK;
}

extern "C" int pdac_groupnorm_bwd_reduce(const PdacView* dy, const PdacView* y, const PdacView* x, int G,
                                         const float* stats, const float* gamma, float eps, float* gsums, float* dgamma,
                                         float* dbeta, float* scratch, void* stream) {
  ImgView dv, yv, xv;
  int cpg = 0;
  PDAC_TRY(img_view(dy, &dv, "groupnorm dy"));
  PDAC_TRY(img_view(x, &xv, "groupnorm x"));
  memset(&yv, 0, sizeof(yv));
  if (y) PDAC_TRY(img_view(y, &yv, "groupnorm y"));
  PDAC_TRY(gn_check(x, G, &cpg));
  const ImgGrid g = img_grid(xv.npix, x->C, x->N);
  gn_bwd_reduce_kernel<<<g.grid, 256, 0, (cudaStream_t)stream>>>(dv, yv, xv, cpg, stats, G, gamma, eps, gsums, dgamma,
                                                                 dbeta, scratch, g.cg, g.ppb);
  PDAC_LAUNCH_CHECK("gn_bwd_reduce_kernel");
  if (scratch) {
    const int work = x->N * G + x->C;
    gn_bwd_finalize_kernel<<<ceil_div(work, 128), 128, 0, (cudaStream_t)stream>>>(scratch, (int)g.grid.x, x->N, x->C, G,
                                                                                gamma, gsums, dgamma, dbeta);
    PDAC_LAUNCH_CHECK("gn_bwd_finalize_kernel");
  }
  return PDAC_OK;
}

extern "C" int pdac_groupnorm_bwd_apply(const PdacView* dy, const PdacView* y, const PdacView* x, int G,
                                        const float* stats, const float* gamma, float eps, const float* gsums,
                                        const PdacView* dx, void* stream) {
  ImgView dv, yv, xv, ov;
  int cpg = 0;
  PDAC_TRY(img_view(dy, &dv, "groupnorm dy"));
  PDAC_TRY(img_view(x, &xv, "groupnorm x"));
  PDAC_TRY(img_view(dx, &ov, "groupnorm dx"));
  memset(&yv, 0, sizeof(yv));
  if (y) PDAC_TRY(img_view(y, &yv, "groupnorm y"));
  PDAC_TRY(gn_check(x, G, &cpg));
  const ImgGrid g = img_grid(xv.npix, x->C, x->N);
  gn_bwd_apply_kernel<<<g.grid, 256, 0, (cudaStream_t)stream>>>(dv, yv, xv, cpg, stats, G, gamma, eps, gsums, ov, g.cg,
                                                                g.ppb);
  PDAC_LAUNCH_CHECK("gn_bwd_apply_kernel");
  return PDAC_OK;
}

extern "C" int pdac_dropout(const PdacView* x, const PdacView* y, float p, unsigned long long seed,
                            const long long* step, int per_channel, void* stream) {
  ImgView xv, yv;
  PDAC_TRY(img_view(x, &xv, "dropout x"));
  PDAC_TRY(img_view(y, &yv, "dropout y"));
  PDAC_REQUIRE(p >= 0.f && p < 1.f, "dropout probability %g out of range [0, 1)", (double)p);
  const ImgGrid g = img_grid(xv.npix, x->C, x->N);
  dropout_kernel<<<g.grid, 256, 0, (cudaStream_t)stream>>>(xv, yv, p, seed, step, per_channel, g.cg, g.ppb);
  PDAC_LAUNCH_CHECK("dropout_kernel");
  return PDAC_OK;
}

extern "C" int pdac_dwconv3x3(const PdacView* x, const float* w, int dil, int flip, const PdacView* y, void* stream) {
  ImgView xv, yv;
  PDAC_TRY(img_view(x, &xv, "dwconv x"));
  PDAC_TRY(img_view(y, &yv, "dwconv y"));
  PDAC_REQUIRE(x->H == y->H && x->W == y->W && x->C == y->C && x->N == y->N && dil >= 1, "dwconv: shape mismatch");
  const ImgGrid g = img_grid(xv.npix, x->C, x->N);
  if (dil == 1 && x->W >= 16 && x->W % 2 == 0) {
    dw3x3_d1_kernel<<<dim3(ceil_div(x->H, g.ppb), x->N), 256, 0, (cudaStream_t)stream>>>(xv, x->H, x->W, w, flip, yv,
                                                                                         g.cg, g.ppb);
    PDAC_LAUNCH_CHECK("dw3x3_d1_kernel");
    return PDAC_OK;
  }
  dw3x3_kernel<<<g.grid, 256, 0, (cudaStream_t)stream>>>(xv, x->H, x->W, w, dil, flip, yv, g.cg, g.ppb);
  PDAC_LAUNCH_CHECK("dw3x3_kernel");
  return PDAC_OK;
}

extern "C" int pdac_dwconv3x3_wgrad(const PdacView* x, const PdacView* dy, int dil, float* dw, float* scratch,
                                    void* stream) {
  ImgView xv, dv;
  PDAC_TRY(img_view(x, &xv, "dwconv x"));
  PDAC_TRY(img_view(dy, &dv, "dwconv dy"));
  PDAC_REQUIRE(x->H == dy->H && x->W == dy->W && x->C == dy->C && x->N == dy->N && dil >= 1, "dwconv: shape mismatch");
  ImgGrid g = img_grid(xv.npix, x->C, x->N);
  // fewer, longer-running blocks: every thread ends with 72 accumulators to hand over
  int cap = (num_sms() * 2 + x->N - 1) / x->N;
  if (cap < 1) cap = 1;
  if ((int)g.grid.x > cap) g.grid.x = cap;
  const size_t smem = (size_t)g.cg * 72 * sizeof(float);
  PDAC_REQUIRE(smem <= 48 * 1024, "dwconv wgrad: too many channels (%d)", x->C);
  const int ncols = g.cg * 72;  // == C * 9
  if (dil == 1 && x->W >= 16 && g.ppb >= 3) {
    int bands = ceil_div(x->H, g.ppb / 3);
    int per_img = (num_sms() * 6 + x->N - 1) / x->N;  // ~6 blocks per SM in total, each walking whole rows
    if (per_img > bands) per_img = bands;
    dw3x3_d1_wgrad_kernel<<<dim3(per_img, x->N), 256, smem, (cudaStream_t)stream>>>(xv, dv, x->H, x->W, dw, scratch,
                                                                                    g.cg, g.ppb);
    PDAC_LAUNCH_CHECK("dw3x3_d1_wgrad_kernel");
    if (scratch) PDAC_TRY(rows_sum(scratch, per_img * x->N, ncols, dw, 1, (cudaStream_t)stream));
    return PDAC_OK;
  }
  dw3x3_wgrad_kernel<<<g.grid, 256, smem, (cudaStream_t)stream>>>(xv, dv, x->H, x->W, dil, dw, scratch, g.cg, g.ppb);
  PDAC_LAUNCH_CHECK("dw3x3_wgrad_kernel");
  if (scratch) PDAC_TRY(rows_sum(scratch, (int)g.grid.x * x->N, ncols, dw, 1, (cudaStream_t)stream));
  return PDAC_OK;
}
