// BatchNorm2d (train + eval, forward + backward), ReLU / residual-add / bias-gradient passes on NHWC bf16.
// All kernels are HBM-bound: one thread owns 8 consecutive channels (16 bytes) of a pixel and keeps its per-channel
// constants in registers (the thread->channel-group mapping is fixed across the grid-stride loop).
// Reference: torchvision resnet.py:82-101 (bn1/relu/bn2/+=identity/relu) and nn.BatchNorm2d semantics (eps 1e-5,
// momentum 0.1, biased variance to normalise, unbiased for running_var).
#include "common.h"
#include "vec.cuh"

namespace pdac {

static int flatten(const PdacView* v, Flat* f, const char* what) {
  PDAC_REQUIRE(v && v->ptr, "%s: null view", what);
  PDAC_REQUIRE(v->C % 8 == 0, "%s: channels (%d) must be a multiple of 8", what, v->C);
  PDAC_REQUIRE((v->H == 1 || v->sH == (int64_t)v->W * v->sW) && (v->N == 1 || v->sN == (int64_t)v->H * v->W * v->sW),
               "%s: view must be pixel-contiguous", what);
  PDAC_REQUIRE(((uintptr_t)v->ptr & 15) == 0 && v->sW % 8 == 0, "%s: view must be 16-byte aligned", what);
  f->ptr = (bf16*)v->ptr;
  f->npix = (long long)v->N * v->H * v->W;
  f->C = v->C;
  f->ps = v->sW;
  return PDAC_OK;
}

struct EwGrid {
  int blocks, cg, ppb;
};
// The grid is persistent: exactly the number of 256-thread blocks that are resident at once (occupancy query per
// kernel), so every thread loads its per-channel constants once and there is a single wave (with 8 blocks per SM
// requested and 2 resident, bn_bwd_apply ran four waves, each paying the constant loads and the DRAM latency chain:
// 25 us for a 17 MB tensor).
template <typename K>
static int resident_blocks(K kernel) {
  int per_sm = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, 256, 0) != cudaSuccess || per_sm < 1) per_sm = 1;
  return per_sm * num_sms();
}
static EwGrid ew_grid(long long npix, int C, int resident) {
  EwGrid g;
  g.cg = C / 8;
  g.ppb = 256 / g.cg;
  if (g.ppb < 1) g.ppb = 1;
  long long need = (npix + g.ppb - 1) / g.ppb;
  long long cap = resident;
  g.blocks = (int)(need < cap ? need : cap);
  if (g.blocks < 1) g.blocks = 1;
  return g;
}
#define EW_SETUP(cg_, ppb_)                                   \
  const int g = threadIdx.x % (cg_), pl = threadIdx.x / (cg_); \
  const bool active = pl < (ppb_) && g < (cg_);                \
  const long long pstep = (long long)gridDim.x * (ppb_);       \
  const long long pfirst = (long long)blockIdx.x * (ppb_) + pl;

// ------------------------------------------------------------------------------------------------ bn_prepare
// stats: [n_rows][2][C] partial sums (one row per producing CTA / warpgroup, unused rows zero). The rows are added in
// a fixed order (32 interleaved row groups per channel, four independent loads in flight per thread, groups combined in
// index order), in double: bit-reproducible batch statistics whatever order the producers ran in.
__global__ void __launch_bounds__(1024) bn_prepare_kernel(const float* __restrict__ stats, int n_rows, double count,
                                  const float* __restrict__ gamma,
                                  const float* __restrict__ beta, float eps, float momentum, int train,
                                  float* running_mean, float* running_var, long long* num_batches_tracked,
                                  float* __restrict__ scale_shift, float* __restrict__ mean_invstd, int C) {
  __shared__ double part[32][32][2];
  const int cl = threadIdx.x & 31, rg = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + cl;
  if (train) {
    double a0 = 0.0, a1 = 0.0;
    if (c < C) {
      int r = rg;
      for (; r + 96 < n_rows; r += 128) {  // four rows (r, r+32, r+64, r+96) per trip: eight loads in flight
        const float* p0 = stats + ((size_t)r * 2) * C + c;
        const size_t st = (size_t)64 * C;
        const float x0 = p0[0], y0 = p0[C], x1 = p0[st], y1 = p0[st + C];
        const float x2 = p0[2 * st], y2 = p0[2 * st + C], x3 = p0[3 * st], y3 = p0[3 * st + C];
        a0 += (double)x0;
        a0 += (double)x1;
        a0 += (double)x2;
        a0 += (double)x3;
        a1 += (double)y0;
        a1 += (double)y1;
        a1 += (double)y2;
        a1 += (double)y3;
      }
      for (; r < n_rows; r += 32) {
        a0 += (double)stats[((size_t)r * 2) * C + c];
        a1 += (double)stats[((size_t)r * 2 + 1) * C + c];
      }
    }
    part[rg][cl][0] = a0;
    part[rg][cl][1] = a1;
  }
  __syncthreads();
  if (rg != 0 || c >= C) return;
  if (c == 0 && train && num_batches_tracked) *num_batches_tracked += 1;
  float mean, var;
  if (train) {
    double t0 = 0.0, t1 = 0.0;
    for (int g = 0; g < 32; ++g) {
      t0 += part[g][cl][0];
      t1 += part[g][cl][1];
    }
    const double m = t0 / count;
    double v = t1 / count - m * m;
    if (v < 0) v = 0;
    mean = (float)m;
    var = (float)v;
    if (running_mean) {
      const double unb = count > 1 ? v * count / (count - 1) : v;
      running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
      running_var[c] = (1.f - momentum) * running_var[c] + momentum * (float)unb;
    }
  } else {
    mean = running_mean[c];
    var = running_var[c];
  }
  const float invstd = rsqrtf(var + eps);
  const float ga = gamma ? gamma[c] : 1.f, be = beta ? beta[c] : 0.f;
  const float sc = ga * invstd;
  scale_shift[c] = sc;
  scale_shift[C + c] = be - mean * sc;
  if (mean_invstd) {
    mean_invstd[c] = mean;
    mean_invstd[C + c] = invstd;
  }
}

// out[c] (+)= sum_r rows[r][c], r = 0 .. n_rows-1 in a fixed order: the deterministic second stage of every
// cross-block reduction (each block stores its partial row with plain stores; no floating-point atomics).
__global__ void __launch_bounds__(1024) rows_sum_kernel(const float* __restrict__ rows, int n_rows, int ncols,
                                                        float* __restrict__ out, int accumulate) {
  __shared__ float part[32][33];
  const int cl = threadIdx.x & 31, rg = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + cl;
  float a = 0.f;
  if (c < ncols) {
    int r = rg;
    for (; r + 96 < n_rows; r += 128) {
      const float* p0 = rows + (size_t)r * ncols + c;
      const size_t st = (size_t)32 * ncols;
      const float x0 = p0[0], x1 = p0[st], x2 = p0[2 * st], x3 = p0[3 * st];
      a += x0;
      a += x1;
      a += x2;
      a += x3;
    }
    for (; r < n_rows; r += 32) a += rows[(size_t)r * ncols + c];
  }
  part[rg][cl] = a;
  __syncthreads();
  if (rg == 0 && c < ncols) {
    float t = 0.f;
    for (int g = 0; g < 32; ++g) t += part[g][cl];
    out[c] = accumulate ? out[c] + t : t;
  }
}
int rows_sum(const float* rows, int n_rows, int ncols, float* out, int accumulate, cudaStream_t stream) {
  rows_sum_kernel<<<ceil_div(ncols, 32), 1024, 0, stream>>>(rows, n_rows, ncols, out, accumulate);
  PDAC_LAUNCH_CHECK("rows_sum");
  return PDAC_OK;
}

// ------------------------------------------------------------------------------------------------ y = act(x*s + b + r)
// UP2: y is the nearest-neighbour x2 upsampling of the result (smp UnetDecoder's F.interpolate(scale_factor=2) fused
// into the producer's normalise pass: every output pixel is stored four times, the low-resolution tensor never exists)
template <bool UP2>
__global__ void __launch_bounds__(256) scale_shift_act_kernel(Flat x, const float* __restrict__ ss, Flat res, int relu,
                                                              Flat y, int cg, int ppb, int H, int W) {
  EW_SETUP(cg, ppb)
  if (!active) return;
  float sc[8], sh[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    sc[i] = ss ? ss[g * 8 + i] : 1.f;
    sh[i] = ss ? ss[x.C + g * 8 + i] : 0.f;
  }
  for (long long p = pfirst; p < x.npix; p += 2 * pstep) {
    const long long p2 = p + pstep;
    const bool two = p2 < x.npix;
    uint4 rx[2], rr[2];
    rx[0] = ld16_stream(x.ptr + p * x.ps + g * 8);
    if (two) rx[1] = ld16_stream(x.ptr + p2 * x.ps + g * 8);
    if (res.ptr) {
      rr[0] = ld16_stream(res.ptr + p * res.ps + g * 8);
      if (two) rr[1] = ld16_stream(res.ptr + p2 * res.ps + g * 8);
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      if (u == 1 && !two) break;
      F8 a = unpack8(rx[u]);
#pragma unroll
      for (int i = 0; i < 8; ++i) a.v[i] = fmaf(a.v[i], sc[i], sh[i]);
      if (res.ptr) {
        const F8 r = unpack8(rr[u]);
#pragma unroll
        for (int i = 0; i < 8; ++i) a.v[i] += r.v[i];
      }
      if (relu) {
#pragma unroll
        for (int i = 0; i < 8; ++i) a.v[i] = fmaxf(a.v[i], 0.f);
      }
      const long long pp = u ? p2 : p;
      if (UP2) {
        const long long hw = (long long)H * W;
        const long long n = pp / hw;
        const int rem = (int)(pp - n * hw);
        const int h = rem / W, w = rem - h * W;
        const uint4 v = pack8(a);
        bf16* o = y.ptr + ((n * 2 * H + 2 * h) * (2LL * W) + 2 * w) * y.ps + g * 8;
        st16(o, v);
        st16(o + y.ps, v);
        st16(o + 2LL * W * y.ps, v);
        st16(o + (2LL * W + 1) * y.ps, v);
      } else {
        st16(y.ptr + pp * y.ps + g * 8, pack8(a));
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------ BN backward
// pass 1: sums[c] = sum dz, sums[C+c] = sum dz * xhat, dz = dy * (y > 0). The ReLU mask comes from y or, when y is
// NULL and `ss` (the forward's scale / shift) is given, is recomputed as x*scale + shift > 0 — the same fma the forward
// rounded to bf16, so the decision is identical and y is not read at all. Two pixels per iteration keep more loads
// in flight.
// XMASK: compile-time copy of the hot case (mask from x, y never read) so that its loads and constants fit 3 blocks / SM
template <bool XMASK>
__global__ void __launch_bounds__(256, XMASK ? 3 : 2) bn_bwd_reduce_kernel(Flat dy, Flat y, Flat x, const float* __restrict__ mi,
                                                            const float* __restrict__ ss, float* __restrict__ sums,
                                                            int cg, int ppb) {
  __shared__ float red[256][17];
  EW_SETUP(cg, ppb)
  float s0[8], s1[8], mean[8], sc[8], sh[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    s0[i] = 0.f;
    s1[i] = 0.f;
    mean[i] = (active && mi) ? mi[g * 8 + i] : 0.f;
    sc[i] = (active && ss) ? ss[g * 8 + i] : 0.f;
    sh[i] = (active && ss) ? ss[dy.C + g * 8 + i] : 0.f;
  }
  const bool xmask = XMASK || (ss != nullptr && y.ptr == nullptr && x.ptr != nullptr);
  if (XMASK) y.ptr = nullptr;
  if (active) {
    for (long long p = pfirst; p < dy.npix; p += 2 * pstep) {
      const long long p2 = p + pstep;
      const bool two = p2 < dy.npix;
      uint4 rd[2], ry[2], rx[2];
      rd[0] = ld16_stream(dy.ptr + p * dy.ps + g * 8);
      if (two) rd[1] = ld16_stream(dy.ptr + p2 * dy.ps + g * 8);
      if (!XMASK && y.ptr) {
        ry[0] = ld16_stream(y.ptr + p * y.ps + g * 8);
        if (two) ry[1] = ld16_stream(y.ptr + p2 * y.ps + g * 8);
      }
      if (x.ptr) {
        rx[0] = ld16_stream(x.ptr + p * x.ps + g * 8);
        if (two) rx[1] = ld16_stream(x.ptr + p2 * x.ps + g * 8);
      }
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        if (u == 1 && !two) break;
        F8 d = unpack8(rd[u]);
        F8 xx;
        if (x.ptr) xx = unpack8(rx[u]);
        if (!XMASK && y.ptr) {
          const F8 yy = unpack8(ry[u]);
#pragma unroll
          for (int i = 0; i < 8; ++i) d.v[i] = yy.v[i] > 0.f ? d.v[i] : 0.f;
        } else if (xmask) {
#pragma unroll
          for (int i = 0; i < 8; ++i) d.v[i] = fmaf(xx.v[i], sc[i], sh[i]) > 0.f ? d.v[i] : 0.f;
        }
        if (x.ptr) {
#pragma unroll
          for (int i = 0; i < 8; ++i) s1[i] = fmaf(d.v[i], xx.v[i] - mean[i], s1[i]);
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) s0[i] += d.v[i];
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    red[threadIdx.x][i] = s0[i];
    red[threadIdx.x][8 + i] = s1[i];
  }
  __syncthreads();
  // thread t < cg*16 reduces one (channel-group, slot) column over the ppb pixel lanes
  for (int t = threadIdx.x; t < cg * 16; t += blockDim.x) {
    const int gg = t / 16, slot = t % 16;
    float acc = 0.f;
    for (int q = 0; q < ppb; ++q) acc += red[q * cg + gg][slot];
    const int c = gg * 8 + (slot & 7);
    // this block's row of the [blocks][2][C] partial table (rows_sum adds the rows in a fixed order afterwards)
    float* row = sums + (size_t)blockIdx.x * 2 * dy.C;
    if (slot < 8)
      row[c] = acc;
    else
      row[dy.C + c] = x.ptr ? acc * (mi ? mi[dy.C + c] : 1.f) : 0.f;
  }
}

// pass 2: dx = gamma*invstd*(dz - sum_dz/M - xhat * sum_dz_xhat/M)   (train)   |   gamma*invstd*dz   (eval)
template <bool XMASK>  // XMASK: train mode, mask from x, no dz output (the hot case)
__global__ void __launch_bounds__(256, XMASK ? 3 : 2) bn_bwd_apply_kernel(Flat dy, Flat y, Flat x, const float* __restrict__ mi,
                                                           const float* __restrict__ gamma, const float* __restrict__ ss,
                                                           const float* __restrict__ sums, float inv_count, int train,
                                                           Flat dx, Flat dz, float* __restrict__ dgamma,
                                                           float* __restrict__ dbeta, int cg, int ppb) {
  EW_SETUP(cg, ppb)
  if (!active) return;
  // dx = a * (dz - m1 - (x - mean) * invstd * m2) = a * dz + k1 * (x - mean) + k0 (x - mean is formed first: folding the
  // mean into k0 cancels catastrophically on the few-sample BatchNorms of the PSP pooling branches)
  float a[8], k1[8], k0[8], mean[8], sc[8], sh[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int c = g * 8 + i;
    const float invstd = mi[dy.C + c];
    mean[i] = mi[c];
    a[i] = (gamma ? gamma[c] : 1.f) * invstd;
    const float m1 = train ? sums[c] * inv_count : 0.f;
    const float m2 = train ? sums[dy.C + c] * inv_count : 0.f;
    k1[i] = -a[i] * invstd * m2;
    k0[i] = -a[i] * m1;
    sc[i] = ss ? ss[c] : 0.f;
    sh[i] = ss ? ss[dy.C + c] : 0.f;
    if (blockIdx.x == 0 && pl == 0) {
      if (dgamma) dgamma[c] = sums[dy.C + c];
      if (dbeta) dbeta[c] = sums[c];
    }
  }
  const bool xmask = XMASK || (ss != nullptr && y.ptr == nullptr && x.ptr != nullptr);
  const bool need_x = XMASK || train || xmask;
  if (XMASK) {
    y.ptr = nullptr;
    dz.ptr = nullptr;
    train = 1;
  }
  for (long long p = pfirst; p < dy.npix; p += 2 * pstep) {
    const long long p2 = p + pstep;
    const bool two = p2 < dy.npix;
    uint4 rd[2], ry[2], rx[2];
    rd[0] = ld16_stream(dy.ptr + p * dy.ps + g * 8);
    if (two) rd[1] = ld16_stream(dy.ptr + p2 * dy.ps + g * 8);
    if (!XMASK && y.ptr) {
      ry[0] = ld16_stream(y.ptr + p * y.ps + g * 8);
      if (two) ry[1] = ld16_stream(y.ptr + p2 * y.ps + g * 8);
    }
    if (need_x) {
      rx[0] = ld16_stream(x.ptr + p * x.ps + g * 8);
      if (two) rx[1] = ld16_stream(x.ptr + p2 * x.ps + g * 8);
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      if (u == 1 && !two) break;
      const long long pp = u ? p2 : p;
      F8 d = unpack8(rd[u]);
      F8 xx;
      if (need_x) xx = unpack8(rx[u]);
      if (!XMASK && y.ptr) {
        const F8 yy = unpack8(ry[u]);
#pragma unroll
        for (int i = 0; i < 8; ++i) d.v[i] = yy.v[i] > 0.f ? d.v[i] : 0.f;
      } else if (xmask) {
#pragma unroll
        for (int i = 0; i < 8; ++i) d.v[i] = fmaf(xx.v[i], sc[i], sh[i]) > 0.f ? d.v[i] : 0.f;
      }
      if (dz.ptr) st16(dz.ptr + pp * dz.ps + g * 8, pack8(d));
      F8 o;
      if (train) {
#pragma unroll
        for (int i = 0; i < 8; ++i) o.v[i] = fmaf(a[i], d.v[i], fmaf(k1[i], xx.v[i] - mean[i], k0[i]));
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) o.v[i] = a[i] * d.v[i];
      }
      st16(dx.ptr + pp * dx.ps + g * 8, pack8(o));
    }
  }
}

// dz = dy * (y > 0); dbias[c] += sum dz  (conv+bias+ReLU layers without BatchNorm: models/resunet.py:36-46)
__global__ void __launch_bounds__(256) relu_bwd_bias_kernel(Flat dy, Flat y, Flat dz, float* __restrict__ dbias,
                                                            float* __restrict__ rows, int cg, int ppb) {
  __shared__ float red[256][9];
  EW_SETUP(cg, ppb)
  float s0[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) s0[i] = 0.f;
  if (active) {
    for (long long p = pfirst; p < dy.npix; p += pstep) {
      F8 d = unpack8(ld16_stream(dy.ptr + p * dy.ps + g * 8));
      if (y.ptr) {
        const F8 yy = unpack8(ld16_stream(y.ptr + p * y.ps + g * 8));
#pragma unroll
        for (int i = 0; i < 8; ++i) d.v[i] = yy.v[i] > 0.f ? d.v[i] : 0.f;
      }
      if (dz.ptr) st16(dz.ptr + p * dz.ps + g * 8, pack8(d));
#pragma unroll
      for (int i = 0; i < 8; ++i) s0[i] += d.v[i];
    }
  }
  if (!dbias) return;
#pragma unroll
  for (int i = 0; i < 8; ++i) red[threadIdx.x][i] = s0[i];
  __syncthreads();
  for (int t = threadIdx.x; t < cg * 8; t += blockDim.x) {
    const int gg = t / 8, slot = t % 8;
    float acc = 0.f;
    for (int q = 0; q < ppb; ++q) acc += red[q * cg + gg][slot];
    if (rows)
      rows[(size_t)blockIdx.x * dy.C + gg * 8 + slot] = acc;  // deterministic: fixed-order row sum afterwards
    else
      atomicAdd(dbias + gg * 8 + slot, acc);
  }
}

}  // namespace pdac

using namespace pdac;

extern "C" int pdac_bn_prepare(const float* stats, int n_rows, double count, const float* gamma, const float* beta,
                               float eps, float momentum, int train, float* running_mean, float* running_var,
                               long long* num_batches_tracked, float* scale_shift, float* mean_invstd, int C,
                               void* stream) {
  PDAC_REQUIRE(scale_shift && C > 0, "bn_prepare: null/empty");
  PDAC_REQUIRE(train ? (stats != nullptr && n_rows >= 1) : (running_mean && running_var),
               "bn_prepare: train needs stats, eval needs running stats");
  if (train) PDAC_REQUIRE(count > 1.0, "Expected more than 1 value per channel when training (count=%g)", count);
  bn_prepare_kernel<<<ceil_div(C, 32), 1024, 0, (cudaStream_t)stream>>>(
      stats, n_rows, count, gamma, beta, eps, momentum, train, running_mean, running_var, num_batches_tracked,
      scale_shift, mean_invstd, C);
  PDAC_LAUNCH_CHECK("bn_prepare");
  return PDAC_OK;
}

extern "C" int pdac_scale_shift_act(const PdacView* x, const float* scale_shift, const PdacView* residual, int relu,
                                    const PdacView* y, void* stream) {
  Flat fx, fy, fr = {nullptr, 0, 0, 0};
  PDAC_TRY(flatten(x, &fx, "x"));
  PDAC_TRY(flatten(y, &fy, "y"));
  if (residual) PDAC_TRY(flatten(residual, &fr, "residual"));
  // y either has x's shape or is its nearest-neighbour x2 upsampling (2H x 2W)
  const bool up2 = y->H == 2 * x->H && y->W == 2 * x->W && y->N == x->N && x->H > 0;
  PDAC_REQUIRE((up2 ? fy.npix == 4 * fx.npix : fy.npix == fx.npix) && fy.C == fx.C &&
                   (!residual || (fr.npix == fx.npix && fr.C == fx.C)),
               "scale_shift_act: shape mismatch");
  PDAC_REQUIRE(fx.C <= 2048, "scale_shift_act: too many channels");
  if (fx.npix == 0) return PDAC_OK;
  if (up2) {
    static const int resident = resident_blocks(scale_shift_act_kernel<true>);
    const EwGrid g = ew_grid(fx.npix, fx.C, resident);
    scale_shift_act_kernel<true><<<g.blocks, 256, 0, (cudaStream_t)stream>>>(fx, scale_shift, fr, relu, fy, g.cg, g.ppb,
                                                                            x->H, x->W);
  } else {
    static const int resident = resident_blocks(scale_shift_act_kernel<false>);
    const EwGrid g = ew_grid(fx.npix, fx.C, resident);
    scale_shift_act_kernel<false><<<g.blocks, 256, 0, (cudaStream_t)stream>>>(fx, scale_shift, fr, relu, fy, g.cg,
                                                                             g.ppb, x->H, x->W);
  }
  PDAC_LAUNCH_CHECK("scale_shift_act");
  return PDAC_OK;
}

extern "C" int pdac_bn_bwd_reduce(const PdacView* dy, const PdacView* y, const PdacView* x, const float* mean_invstd,
                                  const float* scale_shift, float* sums, void* stream) {
  Flat fd, fy = {nullptr, 0, 0, 0}, fx = {nullptr, 0, 0, 0};
  PDAC_TRY(flatten(dy, &fd, "dy"));
  if (y) PDAC_TRY(flatten(y, &fy, "y"));
  if (x) PDAC_TRY(flatten(x, &fx, "x"));
  PDAC_REQUIRE(sums, "bn_bwd_reduce: null sums");
  PDAC_REQUIRE(fd.C <= 2048, "bn_bwd_reduce: too many channels");
  // sums: [1 + PDAC_STAT_ROWS][2][C]; rows 1.. take the blocks' partials, row 0 the fixed-order total
  float* rows = sums + 2 * (size_t)fd.C;
  int blocks = 0;
  if (scale_shift && !y && x && mean_invstd) {
    static const int resident = resident_blocks(bn_bwd_reduce_kernel<true>);
    const EwGrid g = ew_grid(fd.npix, fd.C, resident < PDAC_STAT_ROWS ? resident : PDAC_STAT_ROWS);
    blocks = g.blocks;
    bn_bwd_reduce_kernel<true><<<g.blocks, 256, 0, (cudaStream_t)stream>>>(fd, fy, fx, mean_invstd, scale_shift, rows,
                                                                          g.cg, g.ppb);
  } else {
    static const int resident = resident_blocks(bn_bwd_reduce_kernel<false>);
    const EwGrid g = ew_grid(fd.npix, fd.C, resident < PDAC_STAT_ROWS ? resident : PDAC_STAT_ROWS);
    blocks = g.blocks;
    bn_bwd_reduce_kernel<false><<<g.blocks, 256, 0, (cudaStream_t)stream>>>(fd, fy, fx, mean_invstd, scale_shift, rows,
                                                                           g.cg, g.ppb);
  }
  PDAC_LAUNCH_CHECK("bn_bwd_reduce");
  PDAC_TRY(rows_sum(rows, blocks, 2 * fd.C, sums, 0, (cudaStream_t)stream));
  return PDAC_OK;
}

extern "C" int pdac_bn_bwd_apply(const PdacView* dy, const PdacView* y, const PdacView* x, const float* mean_invstd,
                                 const float* gamma, const float* scale_shift, const float* sums, double count,
                                 int train, const PdacView* dx, const PdacView* dz_out, float* dgamma, float* dbeta,
                                 void* stream) {
  Flat fd, fy = {nullptr, 0, 0, 0}, fx = {nullptr, 0, 0, 0}, fdx, fdz = {nullptr, 0, 0, 0};
  PDAC_TRY(flatten(dy, &fd, "dy"));
  if (y) PDAC_TRY(flatten(y, &fy, "y"));
  if (x) PDAC_TRY(flatten(x, &fx, "x"));
  PDAC_TRY(flatten(dx, &fdx, "dx"));
  if (dz_out) PDAC_TRY(flatten(dz_out, &fdz, "dz"));
  PDAC_REQUIRE(mean_invstd && sums, "bn_bwd_apply: null stats");
  PDAC_REQUIRE(!train || x, "bn_bwd_apply: training mode needs the conv output x");
  PDAC_REQUIRE(!(scale_shift && !y) || x, "bn_bwd_apply: recomputing the ReLU mask needs the conv output x");
  if (scale_shift && !y && x && train && !dz_out) {
    static const int resident = resident_blocks(bn_bwd_apply_kernel<true>);
    const EwGrid g = ew_grid(fd.npix, fd.C, resident);
    bn_bwd_apply_kernel<true><<<g.blocks, 256, 0, (cudaStream_t)stream>>>(fd, fy, fx, mean_invstd, gamma, scale_shift,
                                                                         sums, (float)(1.0 / count), train, fdx, fdz,
                                                                         dgamma, dbeta, g.cg, g.ppb);
  } else {
    static const int resident = resident_blocks(bn_bwd_apply_kernel<false>);
    const EwGrid g = ew_grid(fd.npix, fd.C, resident);
    bn_bwd_apply_kernel<false><<<g.blocks, 256, 0, (cudaStream_t)stream>>>(fd, fy, fx, mean_invstd, gamma, scale_shift,
                                                                          sums, (float)(1.0 / count), train, fdx, fdz,
                                                                          dgamma, dbeta, g.cg, g.ppb);
  }
  PDAC_LAUNCH_CHECK("bn_bwd_apply");
  return PDAC_OK;
}

extern "C" int pdac_relu_bwd_bias(const PdacView* dy, const PdacView* y, const PdacView* dz, float* dbias,
                                  float* scratch, void* stream) {
  Flat fd, fy = {nullptr, 0, 0, 0}, fdz = {nullptr, 0, 0, 0};
  PDAC_TRY(flatten(dy, &fd, "dy"));
  if (y) PDAC_TRY(flatten(y, &fy, "y"));
  if (dz) PDAC_TRY(flatten(dz, &fdz, "dz"));
  static const int resident = resident_blocks(relu_bwd_bias_kernel);
  const bool det = dbias && scratch;
  const EwGrid g = ew_grid(fd.npix, fd.C, det && resident > PDAC_STAT_ROWS ? PDAC_STAT_ROWS : resident);
  relu_bwd_bias_kernel<<<g.blocks, 256, 0, (cudaStream_t)stream>>>(fd, fy, fdz, dbias, det ? scratch : nullptr, g.cg,
                                                                   g.ppb);
  PDAC_LAUNCH_CHECK("relu_bwd_bias");
  if (det) PDAC_TRY(rows_sum(scratch, g.blocks, fd.C, dbias, 1, (cudaStream_t)stream));
  return PDAC_OK;
}

// ---- allocation-free variant (no counterpart in the reference) -----------
// pdac_relu_bwd_bias with its [PDAC_STAT_ROWS][C] partial-row table in a file-scope device array instead of caller
// scratch, for callers that cannot cudaMalloc (graph capture, serving loops). Limit: C <= PDAC_NOALLOC_MAX_C (512), i.e.
// 592 * 512 * 4 = 1.2 MB held by every GPU that loads the library, until exit. All calls share the array: calls that
// overlap on one GPU (different streams) race. Nothing has to start at zero — every launched block overwrites its whole
// row and rows_sum reads exactly the rows of this launch — so the previous call's contents never leak into the next.
static __device__ __align__(256) float g_noalloc_rows[PDAC_STAT_ROWS * PDAC_NOALLOC_MAX_C];

// The array's name is only a host-side stand-in here: the device address comes from cudaGetSymbolAddress, looked up once
// per device with that device current.
static int noalloc_rows_ptr(float** out) {
  static float* ptr[64] = {};
  int dev = 0;
  PDAC_CUDA(cudaGetDevice(&dev));
  PDAC_REQUIRE(dev >= 0 && dev < 64, "relu_bwd_bias_noalloc: device ordinal %d out of range", dev);
  float* p = __atomic_load_n(&ptr[dev], __ATOMIC_ACQUIRE);
  if (!p) {
    PDAC_CUDA(cudaGetSymbolAddress((void**)&p, g_noalloc_rows));
    PDAC_REQUIRE(p != nullptr, "relu_bwd_bias_noalloc: cudaGetSymbolAddress returned a null address");
    __atomic_store_n(&ptr[dev], p, __ATOMIC_RELEASE);  // racing threads look up the same address
  }
  *out = p;
  return PDAC_OK;
}

extern "C" int pdac_relu_bwd_bias_noalloc(const PdacView* dy, const PdacView* y, const PdacView* dz, float* dbias,
                                          void* stream) {
  PDAC_REQUIRE(dy && dbias, "relu_bwd_bias_noalloc: dy and dbias are required");
  Flat fd, fy = {nullptr, 0, 0, 0}, fdz = {nullptr, 0, 0, 0};
  PDAC_TRY(flatten(dy, &fd, "dy"));
  if (y) PDAC_TRY(flatten(y, &fy, "y"));
  if (dz) PDAC_TRY(flatten(dz, &fdz, "dz"));
  PDAC_REQUIRE((!y || (fy.npix == fd.npix && fy.C == fd.C)) && (!dz || (fdz.npix == fd.npix && fdz.C == fd.C)),
               "relu_bwd_bias_noalloc: y / dz must have dy's shape");
  PDAC_REQUIRE(fd.C <= PDAC_NOALLOC_MAX_C,
               "relu_bwd_bias_noalloc: C = %d exceeds the static scratch limit of %d channels (use pdac_relu_bwd_bias "
               "with caller scratch)",
               fd.C, PDAC_NOALLOC_MAX_C);
  float* rows = nullptr;
  PDAC_TRY(noalloc_rows_ptr(&rows));
  static const int resident = resident_blocks(relu_bwd_bias_kernel);
  const EwGrid g = ew_grid(fd.npix, fd.C, resident > PDAC_STAT_ROWS ? PDAC_STAT_ROWS : resident);
  relu_bwd_bias_kernel<<<g.blocks, 256, 0, (cudaStream_t)stream>>>(fd, fy, fdz, dbias, rows, g.cg, g.ppb);
  PDAC_LAUNCH_CHECK("relu_bwd_bias_noalloc");
  PDAC_TRY(rows_sum(rows, g.blocks, fd.C, dbias, 1, (cudaStream_t)stream));
  return PDAC_OK;
}
