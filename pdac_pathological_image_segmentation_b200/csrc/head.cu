// The segmentation head at full resolution: Conv2d(C, classes, 3, padding=1) with C = 16 / 32 input channels and 1 or 2
// classes (smp SegmentationHead of Unet, train_worker.py / BASELINE configs 2 and 5: classes = 1 at 512 x 512).
// As a tensor-core GEMM this layer has N = 1: the implicit-GEMM kernels pad it to 16 output channels, write a 16-channel
// bf16 tensor of which one channel is real (537 MB per step at batch 64), convert it to fp32 NCHW in another pass, and do
// the same on the way back (a 16-channel dY, its own bias-gradient pass). It is 144 MACs per pixel — CUDA-core work that is
// purely HBM-bound — so it gets three direct kernels that touch each tensor once:
//   forward   x (bf16 NHWC) -> logits fp32 NCHW (what the callers see; NHWC == NCHW for one class), bias added
//   dgrad     dL/dlogits fp32 NCHW -> dx bf16 NHWC
//   wgrad     dw [classes][C][3][3] and dbias [classes] in the state_dict layout, per-block partial rows added in a
//             fixed order by rows_sum (reproducible, no atomics)
// fp32 master weights are used directly (no bf16 copy): the arithmetic is fp32 on bf16 activations.
#include "common.h"
#include "vec.cuh"
#include <stdlib.h>

namespace pdac {

static constexpr int kHTW = 32, kHTH = 8;  // output tile (one pixel per thread)

__device__ __forceinline__ uint32_t smem_u32_h(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

struct HeadView {
  const bf16* ptr;
  long long sN, sH, sW;
  int N, H, W;
};

// shared-memory tile of the bf16 NHWC source with a one-pixel halo, zero-filled outside the image
template <int CIN>
__device__ __forceinline__ void load_x_tile(const HeadView& x, int n, int h0, int w0, uint4* tile) {
  constexpr int CG = CIN / 8;
  const int total = (kHTH + 2) * (kHTW + 2) * CG;
  for (int i = threadIdx.x; i < total; i += blockDim.x) {
    const int g = i % CG;
    const int px = i / CG;
    const int ty = px / (kHTW + 2), tx = px - ty * (kHTW + 2);
    const int h = h0 + ty - 1, w = w0 + tx - 1;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (h >= 0 && h < x.H && w >= 0 && w < x.W) v = ld16(x.ptr + n * x.sN + h * x.sH + w * x.sW + g * 8);
    tile[i] = v;
  }
}

// tile of the fp32 NCHW gradient (one class) with a one-pixel halo, zero outside the image
__device__ __forceinline__ void load_g_tile(const float* __restrict__ g, int H, int W, int h0, int w0, float* tile) {
  const int total = (kHTH + 2) * (kHTW + 2);
  for (int i = threadIdx.x; i < total; i += blockDim.x) {
    const int ty = i / (kHTW + 2), tx = i - ty * (kHTW + 2);
    const int h = h0 + ty - 1, w = w0 + tx - 1;
    tile[i] = (h >= 0 && h < H && w >= 0 && w < W) ? __ldg(g + (long long)h * W + w) : 0.f;
  }
}

// Register prefetch of the NEXT tile while the current one is computed from shared memory (the blocks are persistent:
// with one small tile per block the kernels were bound by the load -> barrier -> compute latency chain at ~1 TB/s).
// Tile index -> (image, tile row, tile column), advanced by gridDim.x tiles per step without divisions (the ncu source
// page showed the per-tile index arithmetic — divisions, 64-bit address products, bounds tests — costing more warp
// instructions than the MMAs: r02 profiles).
struct TileWalk {
  int tw, th, n, dw, dh, dn, tiles_w, tiles_h;
  __device__ __forceinline__ TileWalk(int first, int step, int tiles_w_, int tiles_h_) : tiles_w(tiles_w_), tiles_h(tiles_h_) {
    tw = first % tiles_w;
    first /= tiles_w;
    th = first % tiles_h;
    n = first / tiles_h;
    dw = step % tiles_w;
    step /= tiles_w;
    dh = step % tiles_h;
    dn = step / tiles_h;
  }
  __device__ __forceinline__ void advance() {
    tw += dw;
    if (tw >= tiles_w) {
      tw -= tiles_w;
      ++th;
    }
    th += dh;
    if (th >= tiles_h) {
      th -= tiles_h;
      ++n;
    }
    n += dn;
  }
  __device__ __forceinline__ int h0() const { return th * kHTH; }
  __device__ __forceinline__ int w0() const { return tw * kHTW; }
};

// HALO = 1: (kHTH+2) x (kHTW+2) pixels starting one pixel up-left of the tile; HALO = 0: the tile's own pixels.
// The position of each of this thread's chunks inside the tile is loop-invariant: computed once (init).
template <int CIN, int HALO>
struct XPrefetch {
  static constexpr int CG = CIN / 8;
  static constexpr int TWP = kHTW + 2 * HALO, THP = kHTH + 2 * HALO;
  static constexpr int TOTAL = TWP * THP * CG;
  static constexpr int NL = (TOTAL + 255) / 256;
  uint4 buf[NL];
  long long off[NL];  // element offset of the chunk from the tile origin pixel (h0, w0)
  int dyx[NL];        // (dy << 16) | (dx & 0xffff), dy / dx relative to (h0, w0); dy = 0x4000 marks an unused slot
  __device__ __forceinline__ void init(const HeadView& x) {
#pragma unroll
    for (int j = 0; j < NL; ++j) {
      const int i = threadIdx.x + j * 256;
      const int g = i % CG, px = i / CG;
      const int ty = px / TWP, tx = px - ty * TWP;
      const int dy = i < TOTAL ? ty - HALO : 0x4000, dx = tx - HALO;
      off[j] = (long long)dy * x.sH + (long long)dx * x.sW + g * 8;
      dyx[j] = (dy << 16) | (dx & 0xffff);
    }
  }
  __device__ __forceinline__ void issue(const HeadView& x, int n, int h0, int w0) {
    const bf16* base = x.ptr + n * x.sN + h0 * x.sH + w0 * x.sW;
    const bool interior = h0 >= HALO && w0 >= HALO && h0 + kHTH + HALO <= x.H && w0 + kHTW + HALO <= x.W;
#pragma unroll
    for (int j = 0; j < NL; ++j) {
      const int dy = dyx[j] >> 16, dx = (int)(short)(dyx[j] & 0xffff);
      bool ok = dy != 0x4000;
      if (!interior) {
        const int h = h0 + dy, w = w0 + dx;
        ok = ok && h >= 0 && h < x.H && w >= 0 && w < x.W;
      }
      buf[j] = ok ? ld16(base + off[j]) : make_uint4(0u, 0u, 0u, 0u);
    }
  }
  // 16-byte chunk g of tile pixel p lives at chunk slot g ^ swz(p): the 8 row addresses of an ldmatrix 8x8 block
  // (8 consecutive pixels, one chunk column) then fall into 8 different 16-byte bank groups. With the plain layout
  // (pitch 32 or 64 bytes) ncu counted an 8-way bank conflict per shared load and the L1 pipe at 96 %.
  static __device__ __forceinline__ int swz(int p) { return CG == 2 ? (p >> 2) & 1 : (CG == 4 ? (p >> 1) & 3 : 0); }
  static __device__ __forceinline__ uint32_t chunk_addr(uint32_t base, int p, int g) {
    return base + (uint32_t)((p * CG + (g ^ swz(p))) * 16);
  }
  __device__ __forceinline__ void store(uint4* tile) const {
#pragma unroll
    for (int j = 0; j < NL; ++j) {
      const int i = threadIdx.x + j * 256;
      const int g = i % CG, px = i / CG;
      if (i < TOTAL) tile[px * CG + (g ^ swz(px))] = buf[j];
    }
  }
};
struct GPrefetch {  // haloed fp32 gradient tile of one class
  static constexpr int TOTAL = (kHTH + 2) * (kHTW + 2);
  static constexpr int NL = (TOTAL + 255) / 256;
  float buf[NL];
  int off[NL], dyx[NL];
  __device__ __forceinline__ void init(int W) {
#pragma unroll
    for (int j = 0; j < NL; ++j) {
      const int i = threadIdx.x + j * 256;
      const int ty = i / (kHTW + 2), tx = i - ty * (kHTW + 2);
      const int dy = i < TOTAL ? ty - 1 : 0x4000, dx = tx - 1;
      off[j] = dy * W + dx;
      dyx[j] = (dy << 16) | (dx & 0xffff);
    }
  }
  // g: the (image, class) plane
  __device__ __forceinline__ void issue(const float* __restrict__ g, int H, int W, int h0, int w0) {
    const float* base = g + (long long)h0 * W + w0;
    const bool interior = h0 >= 1 && w0 >= 1 && h0 + kHTH + 1 <= H && w0 + kHTW + 1 <= W;
#pragma unroll
    for (int j = 0; j < NL; ++j) {
      const int dy = dyx[j] >> 16, dx = (int)(short)(dyx[j] & 0xffff);
      bool ok = dy != 0x4000;
      if (!interior) {
        const int h = h0 + dy, w = w0 + dx;
        ok = ok && h >= 0 && h < H && w >= 0 && w < W;
      }
      buf[j] = ok ? __ldg(base + off[j]) : 0.f;
    }
  }
  __device__ __forceinline__ void store(float* tile) const {
#pragma unroll
    for (int j = 0; j < NL; ++j) {
      const int i = threadIdx.x + j * 256;
      if (i < TOTAL) tile[i] = buf[j];
    }
  }
};

// ---- warp-level tensor-core helpers (mma.sync m16n8k16, bf16 x bf16 -> fp32): the head is far too small for a
// tcgen05 pipeline (N = 1 or 2) but at 144 MACs per pixel the scalar version above is instruction-bound at ~1.1 TB/s;
// with the pixels as the M (forward) or K (weight gradient) dimension of warp MMAs it is HBM-bound.
__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_trans(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// Forward on warp MMAs: a warp owns one row of the 32 x 8 tile = two segments of 16 pixels (M); K = the CIN channels of
// one tap (ldmatrix straight from the haloed shared-memory tile, shifted by the tap); N = 8 with the classes in the
// first columns (weights rounded to bf16, as the tensor-core convolutions have them). fp32 accumulation over the 9 taps.
template <int CIN, int COUT>
__global__ void __launch_bounds__(256) head3x3_fwd_mma_kernel(HeadView x, const float* __restrict__ wgt,
                                                              const float* __restrict__ bias, float* __restrict__ out,
                                                              int tiles_w, int tiles_h, int n_tiles) {
  constexpr int CG = CIN / 8, KS = CIN / 16;
  __shared__ __align__(16) uint4 tile[(kHTH + 2) * (kHTW + 2) * CG];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int gid = lane >> 2, tid = lane & 3;
  // B fragments: B[k = ci][n = class]; this lane holds k = tid*2, tid*2+1 (b0) and +8 (b1) of column n = gid
  uint32_t bw[9][KS][2];
#pragma unroll
  for (int t = 0; t < 9; ++t)
#pragma unroll
    for (int ks = 0; ks < KS; ++ks)
#pragma unroll
      for (int hlf = 0; hlf < 2; ++hlf) {
        float v0 = 0.f, v1 = 0.f;
        if (gid < COUT) {
          const int ci = ks * 16 + hlf * 8 + tid * 2;
          v0 = wgt[((long long)gid * CIN + ci) * 9 + t];
          v1 = wgt[((long long)gid * CIN + ci + 1) * 9 + t];
        }
        bw[t][ks][hlf] = pack2(v0, v1);
      }
  float bv[COUT];
#pragma unroll
  for (int co = 0; co < COUT; ++co) bv[co] = bias ? bias[co] : 0.f;
  const uint32_t tbase = smem_u32_h(tile);
  const int ty = warp;            // tile row of this warp
  XPrefetch<CIN, 1> pf;
  pf.init(x);
  TileWalk wk(blockIdx.x, gridDim.x, tiles_w, tiles_h);
  if ((int)blockIdx.x < n_tiles) pf.issue(x, wk.n, wk.h0(), wk.w0());
  for (int tl = blockIdx.x; tl < n_tiles; tl += gridDim.x) {
    const int n = wk.n, h0 = wk.h0(), w0 = wk.w0();
    __syncthreads();  // the previous tile has been consumed
    pf.store(tile);
    __syncthreads();
    wk.advance();
    if (tl + (int)gridDim.x < n_tiles) pf.issue(x, wk.n, wk.h0(), wk.w0());  // flies while this tile is computed
#pragma unroll
    for (int seg = 0; seg < 2; ++seg) {
      float d[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int s = 0; s < 3; ++s) {
          // ldmatrix lane address: matrix j = lane / 8: pixels (j & 1) * 8 + lane % 8, channels (j >> 1) * 8 ..
          const int px = seg * 16 + s + (lane & 7) + ((lane >> 3) & 1) * 8;
          const int plin = (ty + r) * (kHTW + 2) + px;
#pragma unroll
          for (int ks = 0; ks < KS; ++ks) {
            uint32_t a[4];
            ldsm_x4(XPrefetch<CIN, 1>::chunk_addr(tbase, plin, ks * 2 + (lane >> 4)), a);
            mma_bf16_16816(d, a, bw[r * 3 + s][ks][0], bw[r * 3 + s][ks][1]);
          }
        }
      // d[0], d[1] = classes tid*2, tid*2+1 of pixel gid; d[2], d[3] of pixel gid + 8: the classes sit in columns 0..1
      if (tid == 0) {
        const int h = h0 + ty;
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          const int w = w0 + seg * 16 + gid + half * 8;
          if (h < x.H && w < x.W) {
#pragma unroll
            for (int co = 0; co < COUT; ++co)
              out[(((long long)n * COUT + co) * x.H + h) * x.W + w] = d[half * 2 + co] + bv[co];
          }
        }
      }
    }
  }
}

// Weight gradient on warp MMAs, one class per pass: D[m = tap][n = ci] += sum_k A[m][k] B[k][n] with k = the 16 pixels u
// of a segment, A[tap][k] = bf16(g[u_k - off_tap]) built from the haloed fp32 gradient tile and B[k][ci] = x[u_k][ci]
// (ldmatrix.trans from the pixel-major tile). The 16 x CIN accumulators stay in registers over all tiles of the
// (persistent) block; warps and blocks are then added in index order (reproducible).
template <int CIN, int COUT>
__global__ void __launch_bounds__(256) head3x3_wgrad_mma_kernel(HeadView x, const float* __restrict__ g,
                                                                float* __restrict__ rows, int tiles_w, int tiles_h,
                                                                int n_tiles) {
  constexpr int CG = CIN / 8, NT = CIN / 8;  // NT n-tiles of 8 input channels
  (void)CG;
  __shared__ __align__(16) uint4 tile[kHTH * kHTW * (CIN / 8)];  // source pixels of the tile (no halo), pixel-major
  __shared__ float gt[(kHTH + 2) * (kHTW + 2)];
  __shared__ float red[8][16 * CIN + 1];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int gid = lane >> 2, tid = lane & 3;
  const uint32_t tbase = smem_u32_h(tile);
  for (int co = 0; co < COUT; ++co) {
    float d[NT][4];
#pragma unroll
    for (int i = 0; i < NT; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) d[i][j] = 0.f;
    float gsum = 0.f;
    XPrefetch<CIN, 0> pfx;
    GPrefetch pfg;
    pfx.init(x);
    pfg.init(x.W);
    TileWalk wk(blockIdx.x, gridDim.x, tiles_w, tiles_h);  // gridDim.x <= n_tiles
    const long long plane = (long long)x.H * x.W;
    pfx.issue(x, wk.n, wk.h0(), wk.w0());
    pfg.issue(g + ((long long)wk.n * COUT + co) * plane, x.H, x.W, wk.h0(), wk.w0());
    for (int tl = blockIdx.x; tl < n_tiles; tl += gridDim.x) {
      __syncthreads();  // previous tile fully consumed
      pfx.store(tile);
      pfg.store(gt);
      __syncthreads();
      wk.advance();
      if (tl + (int)gridDim.x < n_tiles) {  // next tile's loads fly while this one is computed
        pfx.issue(x, wk.n, wk.h0(), wk.w0());
        pfg.issue(g + ((long long)wk.n * COUT + co) * plane, x.H, x.W, wk.h0(), wk.w0());
      }
      const int ty = warp;
#pragma unroll
      for (int seg = 0; seg < 2; ++seg) {
        // A fragment: rows gid (tap gid) and gid + 8 (tap 8 for gid == 0, else zero), k = tid*2, tid*2+1, +8, +9
        uint32_t a[4];
        const int r0 = gid / 3, s0 = gid - r0 * 3;
        const float* grow = gt + (ty + 2 - r0) * (kHTW + 2) + seg * 16 + 2 - s0;   // g[u - (r-1, s-1)]
        const float g0 = grow[tid * 2], g1 = grow[tid * 2 + 1], g2 = grow[tid * 2 + 8], g3 = grow[tid * 2 + 9];
        a[0] = pack2(g0, g1);
        a[2] = pack2(g2, g3);
        a[1] = a[3] = 0u;
        if (gid == 0) {  // tap 8 = (r, s) = (2, 2)
          const float* g8 = gt + (ty + 0) * (kHTW + 2) + seg * 16 + 0;
          a[1] = pack2(g8[tid * 2], g8[tid * 2 + 1]);
          a[3] = pack2(g8[tid * 2 + 8], g8[tid * 2 + 9]);
        }
        if (gid == 4) gsum += (g0 + g1) + (g2 + g3);  // tap 4 is the centre: every pixel of the segment exactly once
        // B fragments of all n-tiles: matrix j of an x4 load = pixels (j & 1) * 8 .., channels (j >> 1) * 8 ..
        const int px = seg * 16 + (lane & 7) + ((lane >> 3) & 1) * 8;
#pragma unroll
        for (int np = 0; np < NT / 2; ++np) {
          uint32_t bb[4];
          ldsm_x4_trans(XPrefetch<CIN, 0>::chunk_addr(tbase, ty * kHTW + px, np * 2 + (lane >> 4)), bb);
          mma_bf16_16816(d[np * 2], a, bb[0], bb[1]);
          mma_bf16_16816(d[np * 2 + 1], a, bb[2], bb[3]);
        }
      }
    }
    // warp -> shared -> fixed-order sum over the 8 warps -> this block's row [CIN*9 + 1]
    __syncthreads();
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      // d[nt][0], [1]: tap gid, channels nt*8 + tid*2, +1; d[nt][2], [3]: tap gid + 8
      red[warp][gid * CIN + nt * 8 + tid * 2] = d[nt][0];
      red[warp][gid * CIN + nt * 8 + tid * 2 + 1] = d[nt][1];
      red[warp][(gid + 8) * CIN + nt * 8 + tid * 2] = d[nt][2];
      red[warp][(gid + 8) * CIN + nt * 8 + tid * 2 + 1] = d[nt][3];
    }
    // bias gradient: the four tid lanes of gid == 4 hold partial sums: add them in lane order
    float gs = gsum;
    gs += __shfl_down_sync(0xffffffffu, gs, 2, 4);
    gs += __shfl_down_sync(0xffffffffu, gs, 1, 4);
    if (lane == 16) red[warp][16 * CIN] = gs;
    __syncthreads();
    float* row = rows + ((size_t)blockIdx.x * COUT + co) * (CIN * 9 + 1);
    for (int i = threadIdx.x; i < CIN * 9 + 1; i += blockDim.x) {
      float v = 0.f;
      if (i < CIN * 9) {
        const int ci = i / 9, t = i - ci * 9;  // state_dict order within a class: [ci][tap]
        for (int w8 = 0; w8 < 8; ++w8) v += red[w8][t * CIN + ci];
      } else {
        for (int w8 = 0; w8 < 8; ++w8) v += red[w8][16 * CIN];
      }
      row[i] = v;
    }
  }
}

template <int CIN, int COUT>
__global__ void __launch_bounds__(256) head3x3_fwd_kernel(HeadView x, const float* __restrict__ wgt,
                                                          const float* __restrict__ bias, float* __restrict__ out,
                                                          int tiles_w, int tiles_h) {
  constexpr int CG = CIN / 8;
  __shared__ uint4 tile[(kHTH + 2) * (kHTW + 2) * CG];
  __shared__ float w_s[COUT][9][CIN];
  int b = blockIdx.x;
  const int tw = b % tiles_w;
  b /= tiles_w;
  const int th = b % tiles_h;
  const int n = b / tiles_h;
  const int h0 = th * kHTH, w0 = tw * kHTW;
  for (int i = threadIdx.x; i < COUT * 9 * CIN; i += blockDim.x) {
    const int ci = i % CIN, t = (i / CIN) % 9, co = i / (9 * CIN);
    w_s[co][t][ci] = wgt[((long long)co * CIN + ci) * 9 + t];  // state_dict [co][ci][r][s]
  }
  load_x_tile<CIN>(x, n, h0, w0, tile);
  __syncthreads();
  const int tx = threadIdx.x % kHTW, ty = threadIdx.x / kHTW;
  const int h = h0 + ty, w = w0 + tx;
  float acc[COUT];
#pragma unroll
  for (int co = 0; co < COUT; ++co) acc[co] = bias ? bias[co] : 0.f;
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int s = 0; s < 3; ++s) {
      const uint4* px = tile + ((ty + r) * (kHTW + 2) + tx + s) * CG;
#pragma unroll
      for (int g = 0; g < CG; ++g) {
        const F8 v = unpack8(px[g]);
#pragma unroll
        for (int co = 0; co < COUT; ++co)
#pragma unroll
          for (int i = 0; i < 8; ++i) acc[co] = fmaf(v.v[i], w_s[co][r * 3 + s][g * 8 + i], acc[co]);
      }
    }
  if (h < x.H && w < x.W) {
#pragma unroll
    for (int co = 0; co < COUT; ++co) out[(((long long)n * COUT + co) * x.H + h) * x.W + w] = acc[co];
  }
}

// Data gradient on warp MMAs: D[m = pixel q][n = ci] = sum_co sum_k A_co[m][k] B_co[k][n] with k = the 9 taps (padded to
// 16), A_co[q][t] = bf16(g[co][q - off_t]) gathered from the haloed gradient tile and B_co[t][ci] = bf16(w[co][ci][t]).
// One MMA per 16 pixels, 8 input channels and class instead of 144 scalar FMAs per pixel (the scalar version below is
// instruction-bound at 2.7 TB/s of its 8 TB/s-class traffic).
template <int CIN, int COUT>
__global__ void __launch_bounds__(256) head3x3_dgrad_mma_kernel(const float* __restrict__ g, const float* __restrict__ wgt,
                                                                bf16* __restrict__ dx, long long dsN, long long dsH,
                                                                long long dsW, int N, int H, int W, int tiles_w,
                                                                int tiles_h, int n_tiles) {
  constexpr int NT = CIN / 8;
  __shared__ float gt[COUT][(kHTH + 2) * (kHTW + 2)];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int gid = lane >> 2, tid = lane & 3;
  // B fragments: b0 = taps tid*2, tid*2+1 of channel nt*8 + gid; b1 = taps tid*2+8, +9 (only tap 8 exists)
  uint32_t bw[COUT][NT][2];
#pragma unroll
  for (int co = 0; co < COUT; ++co)
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      const float* wp = wgt + ((long long)co * CIN + nt * 8 + gid) * 9;
      bw[co][nt][0] = pack2(wp[tid * 2], wp[tid * 2 + 1]);
      bw[co][nt][1] = tid == 0 ? pack2(wp[8], 0.f) : 0u;
    }
  // tile offsets of this lane's taps: tap t = (r, s) reads g at tile offset (ty + 2 - r, tx + 2 - s)
  const int t0 = tid * 2, t1 = tid * 2 + 1;
  const int o0 = (2 - t0 / 3) * (kHTW + 2) + 2 - t0 % 3, o1 = (2 - t1 / 3) * (kHTW + 2) + 2 - t1 % 3;
  GPrefetch pf[COUT];
#pragma unroll
  for (int co = 0; co < COUT; ++co) pf[co].init(W);
  TileWalk wk(blockIdx.x, gridDim.x, tiles_w, tiles_h);
  const long long plane = (long long)H * W;
  if ((int)blockIdx.x < n_tiles) {
#pragma unroll
    for (int co = 0; co < COUT; ++co) pf[co].issue(g + ((long long)wk.n * COUT + co) * plane, H, W, wk.h0(), wk.w0());
  }
  const int ty = warp;
  for (int tl = blockIdx.x; tl < n_tiles; tl += gridDim.x) {
    const int n = wk.n, h0 = wk.h0(), w0 = wk.w0();
    __syncthreads();
#pragma unroll
    for (int co = 0; co < COUT; ++co) pf[co].store(gt[co]);
    __syncthreads();
    wk.advance();
    if (tl + (int)gridDim.x < n_tiles) {
#pragma unroll
      for (int co = 0; co < COUT; ++co) pf[co].issue(g + ((long long)wk.n * COUT + co) * plane, H, W, wk.h0(), wk.w0());
    }
    const int h = h0 + ty;
#pragma unroll
    for (int seg = 0; seg < 2; ++seg) {
      float d[NT][4];
#pragma unroll
      for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int j = 0; j < 4; ++j) d[nt][j] = 0.f;
#pragma unroll
      for (int co = 0; co < COUT; ++co) {
        const float* gp = gt[co] + ty * (kHTW + 2) + seg * 16 + gid;  // pixel gid of the segment; + 8: pixel gid + 8
        uint32_t a[4];
        a[0] = pack2(gp[o0], gp[o1]);
        a[1] = pack2(gp[o0 + 8], gp[o1 + 8]);
        a[2] = a[3] = 0u;
        if (tid == 0) {  // tap 8 = (2, 2): tile offset (ty, tx)
          a[2] = pack2(gp[0], 0.f);
          a[3] = pack2(gp[8], 0.f);
        }
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) mma_bf16_16816(d[nt], a, bw[co][nt][0], bw[co][nt][1]);
      }
      if (h < H) {
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          const int w = w0 + seg * 16 + gid + half * 8;
          if (w < W) {
            bf16* o = dx + n * dsN + h * dsH + w * dsW + tid * 2;
#pragma unroll
            for (int nt = 0; nt < NT; ++nt)
              *reinterpret_cast<uint32_t*>(o + nt * 8) = pack2(d[nt][half * 2], d[nt][half * 2 + 1]);
          }
        }
      }
    }
  }
}

// dx[q][ci] = sum_co sum_{r,s} g[co][q - (r-1, s-1)] * w[co][ci][r][s]   (persistent blocks, next tile prefetched)
template <int CIN, int COUT>
__global__ void __launch_bounds__(256) head3x3_dgrad_kernel(const float* __restrict__ g, const float* __restrict__ wgt,
                                                            bf16* __restrict__ dx, long long dsN, long long dsH,
                                                            long long dsW, int N, int H, int W, int tiles_w,
                                                            int tiles_h, int n_tiles) {
  __shared__ float gt[COUT][(kHTH + 2) * (kHTW + 2)];
  __shared__ __align__(16) float w_s[COUT][9][CIN];
  for (int i = threadIdx.x; i < COUT * 9 * CIN; i += blockDim.x) {
    const int ci = i % CIN, t = (i / CIN) % 9, co = i / (9 * CIN);
    w_s[co][t][ci] = wgt[((long long)co * CIN + ci) * 9 + t];
  }
  GPrefetch pf[COUT];
#pragma unroll
  for (int co = 0; co < COUT; ++co) pf[co].init(W);
  TileWalk wk(blockIdx.x, gridDim.x, tiles_w, tiles_h);
  const long long plane = (long long)H * W;
  if ((int)blockIdx.x < n_tiles) {
#pragma unroll
    for (int co = 0; co < COUT; ++co) pf[co].issue(g + ((long long)wk.n * COUT + co) * plane, H, W, wk.h0(), wk.w0());
  }
  const int tx = threadIdx.x % kHTW, ty = threadIdx.x / kHTW;
  for (int tl = blockIdx.x; tl < n_tiles; tl += gridDim.x) {
    const int n = wk.n, h0 = wk.h0(), w0 = wk.w0();
    __syncthreads();
#pragma unroll
    for (int co = 0; co < COUT; ++co) pf[co].store(gt[co]);
    __syncthreads();
    wk.advance();
    if (tl + (int)gridDim.x < n_tiles) {
#pragma unroll
      for (int co = 0; co < COUT; ++co) pf[co].issue(g + ((long long)wk.n * COUT + co) * plane, H, W, wk.h0(), wk.w0());
    }
    const int h = h0 + ty, w = w0 + tx;
    if (h >= H || w >= W) continue;
    float acc[CIN];
#pragma unroll
    for (int i = 0; i < CIN; ++i) acc[i] = 0.f;
#pragma unroll
    for (int co = 0; co < COUT; ++co)
#pragma unroll
      for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int s = 0; s < 3; ++s) {
          // the output pixel q - (r-1, s-1) sits at tile offset (ty + 1 - (r-1), tx + 1 - (s-1))
          const float gv = gt[co][(ty + 2 - r) * (kHTW + 2) + tx + 2 - s];
          const float4* wv = reinterpret_cast<const float4*>(w_s[co][r * 3 + s]);
#pragma unroll
          for (int i = 0; i < CIN / 4; ++i) {
            const float4 q = wv[i];
            acc[4 * i] = fmaf(gv, q.x, acc[4 * i]);
            acc[4 * i + 1] = fmaf(gv, q.y, acc[4 * i + 1]);
            acc[4 * i + 2] = fmaf(gv, q.z, acc[4 * i + 2]);
            acc[4 * i + 3] = fmaf(gv, q.w, acc[4 * i + 3]);
          }
        }
    bf16* o = dx + n * dsN + h * dsH + w * dsW;
#pragma unroll
    for (int g8 = 0; g8 < CIN / 8; ++g8) {
      F8 f;
#pragma unroll
      for (int i = 0; i < 8; ++i) f.v[i] = acc[g8 * 8 + i];
      st16(o + g8 * 8, pack8(f));
    }
  }
}

// dw[co][ci][r][s] += sum_q x[q + (r-1, s-1)][ci] * g[co][q] = sum_u x[u][ci] * g[co][u - (r-1, s-1)];  dbias[co] += sum g.
// A thread owns 8 channels of one source pixel u per tile and keeps 9 x 8 partial products per class across ALL the
// tiles of its (persistent) block; the block then adds its threads in a fixed order and stores one row
// [COUT][CIN*9 + 1] of the partial table.
template <int CIN, int COUT>
__global__ void __launch_bounds__(256) head3x3_wgrad_kernel(HeadView x, const float* __restrict__ g,
                                                            float* __restrict__ rows, int tiles_w, int tiles_h,
                                                            int n_tiles) {
  constexpr int CG = CIN / 8;
  constexpr int PPT = 256 / CG;          // source pixels per pass of the block
  constexpr int TPX = kHTW * kHTH;       // pixels per tile
  __shared__ float gt[(kHTH + 2) * (kHTW + 2)];
  __shared__ float s_dw[CIN * 9 + 1];
  const int cgi = threadIdx.x % CG, pl = threadIdx.x / CG;
  for (int co = 0; co < COUT; ++co) {
    float acc[9][8];
    float gsum = 0.f;
#pragma unroll
    for (int t = 0; t < 9; ++t)
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[t][i] = 0.f;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
      int b = tile;
      const int tw = b % tiles_w;
      b /= tiles_w;
      const int th = b % tiles_h;
      const int n = b / tiles_h;
      const int h0 = th * kHTH, w0 = tw * kHTW;
      __syncthreads();  // previous tile's gradient values fully consumed
      load_g_tile(g + ((long long)n * COUT + co) * x.H * x.W, x.H, x.W, h0, w0, gt);
      __syncthreads();
      for (int p = pl; p < TPX; p += PPT) {
        const int ty = p / kHTW, tx = p - ty * kHTW;
        const int h = h0 + ty, w = w0 + tx;
        if (h >= x.H || w >= x.W) continue;
        const F8 v = unpack8(ld16(x.ptr + n * x.sN + h * x.sH + w * x.sW + cgi * 8));
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
          for (int s = 0; s < 3; ++s) {
            const float gv = gt[(ty + 2 - r) * (kHTW + 2) + tx + 2 - s];  // g[u - (r-1, s-1)], zero outside the image
#pragma unroll
            for (int i = 0; i < 8; ++i) acc[r * 3 + s][i] = fmaf(v.v[i], gv, acc[r * 3 + s][i]);
          }
        if (cgi == 0) gsum += gt[(ty + 1) * (kHTW + 2) + tx + 1];
      }
    }
    // fixed-order block reduction: one pixel lane adds per round (distinct channel groups -> distinct slots)
    __syncthreads();
    for (int i = threadIdx.x; i < CIN * 9 + 1; i += blockDim.x) s_dw[i] = 0.f;
    __syncthreads();
    for (int q = 0; q < PPT; ++q) {
      if (pl == q) {
#pragma unroll
        for (int t = 0; t < 9; ++t)
#pragma unroll
          for (int i = 0; i < 8; ++i) s_dw[(cgi * 8 + i) * 9 + t] += acc[t][i];
        if (cgi == 0) s_dw[CIN * 9] += gsum;
      }
      __syncthreads();
    }
    float* row = rows + ((size_t)blockIdx.x * COUT + co) * (CIN * 9 + 1);
    for (int i = threadIdx.x; i < CIN * 9 + 1; i += blockDim.x) row[i] = s_dw[i];
  }
}

// rows [nrows][COUT][CIN*9 + 1] -> dw[(co*CIN + ci)*9 + t] += ..., dbias[co] += ... (rows added in index order)
// 32 elements per block; warp k adds the rows k, k + 8, ... and warp 0 the eight sub-sums in index order (a fixed order;
// one thread per element walking all 592 rows was 76 us of dependent L2 round trips)
__global__ void __launch_bounds__(256) head3x3_wgrad_finalize_kernel(const float* __restrict__ rows, int nrows, int cout,
                                                                     int cin9, float* __restrict__ dw,
                                                                     float* __restrict__ dbias) {
  __shared__ float red[8][32];
  const int lane = threadIdx.x & 31, k = threadIdx.x >> 5;
  const int per = cin9 + 1, total = cout * per;
  const int i = blockIdx.x * 32 + lane;
  float a = 0.f;
  if (i < total) {
#pragma unroll 4
    for (int r = k; r < nrows; r += 8) a += rows[(size_t)r * total + i];
  }
  red[k][lane] = a;
  __syncthreads();
  if (k != 0 || i >= total) return;
  float v = red[0][lane];
#pragma unroll
  for (int kk = 1; kk < 8; ++kk) v += red[kk][lane];
  const int co = i / per, kq = i - co * per;
  if (kq < cin9)
    dw[(size_t)co * cin9 + kq] += v;
  else if (dbias)
    dbias[co] += v;
}

static int head_view(const PdacView* v, HeadView* o, const char* what) {
  PDAC_REQUIRE(v && v->ptr, "%s: null view", what);
  PDAC_REQUIRE(((uintptr_t)v->ptr & 15) == 0 && v->sW % 8 == 0 && v->sH % 8 == 0 && v->sN % 8 == 0,
               "%s: view must be 16-byte aligned", what);
  o->ptr = (const bf16*)v->ptr;
  o->sN = v->sN;
  o->sH = v->sH;
  o->sW = v->sW;
  o->N = v->N;
  o->H = v->H;
  o->W = v->W;
  return PDAC_OK;
}

}  // namespace pdac

using namespace pdac;

// PDAC_HEAD_SCALAR=1: the CUDA-core versions of forward / weight gradient (A-B timing, debugging)
static int head_scalar() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("PDAC_HEAD_SCALAR");
    v = (e && e[0] == '1') ? 1 : 0;
  }
  return v;
}

#define HEAD_DISPATCH(CALL)                                             \
  if (x->C == 16 && classes == 1) { CALL(16, 1); }                      \
  else if (x->C == 16 && classes == 2) { CALL(16, 2); }                 \
  else if (x->C == 32 && classes == 1) { CALL(32, 1); }                 \
  else if (x->C == 32 && classes == 2) { CALL(32, 2); }                 \
  else return fail(PDAC_E_INVALID, "head3x3: %d channels / %d classes unsupported (16 or 32 channels, 1 or 2 classes)", x->C, classes)

extern "C" int pdac_head3x3_supported(int channels, int classes) {
  return (channels == 16 || channels == 32) && (classes == 1 || classes == 2);
}

extern "C" int pdac_head3x3_fwd(const PdacView* x, const float* weight, const float* bias, int classes, float* logits,
                                void* stream) {
  PDAC_REQUIRE(x && weight && logits, "head3x3_fwd: null argument");
  HeadView xv;
  PDAC_TRY(head_view(x, &xv, "head x"));
  if ((long long)x->N * x->H * x->W == 0) return PDAC_OK;
  const int tiles_w = ceil_div(x->W, kHTW), tiles_h = ceil_div(x->H, kHTH);
  const int n_tiles = tiles_w * tiles_h * x->N;
  const int blocks = n_tiles;                                   // scalar version: one tile per block
  int pblocks = 6 * num_sms();                                  // warp-MMA version: persistent blocks
  if (pblocks > n_tiles) pblocks = n_tiles;
#define CALL(CI, CO)                                                                                              \
  if (head_scalar())                                                                                              \
    head3x3_fwd_kernel<CI, CO><<<blocks, 256, 0, (cudaStream_t)stream>>>(xv, weight, bias, logits, tiles_w, tiles_h); \
  else                                                                                                            \
    head3x3_fwd_mma_kernel<CI, CO><<<pblocks, 256, 0, (cudaStream_t)stream>>>(xv, weight, bias, logits, tiles_w, tiles_h, n_tiles)
  HEAD_DISPATCH(CALL);
#undef CALL
  PDAC_LAUNCH_CHECK("head3x3_fwd");
  return PDAC_OK;
}

extern "C" int pdac_head3x3_dgrad(const float* grad_logits, const float* weight, int classes, const PdacView* dx,
                                  void* stream) {
  PDAC_REQUIRE(grad_logits && weight && dx, "head3x3_dgrad: null argument");
  const PdacView* x = dx;
  HeadView dv;
  PDAC_TRY(head_view(dx, &dv, "head dx"));
  if ((long long)x->N * x->H * x->W == 0) return PDAC_OK;
  const int tiles_w = ceil_div(x->W, kHTW), tiles_h = ceil_div(x->H, kHTH);
  const int n_tiles = tiles_w * tiles_h * x->N;
  int blocks = 8 * num_sms();
  if (blocks > n_tiles) blocks = n_tiles;
#define CALL(CI, CO)                                                                                                  \
  if (head_scalar())                                                                                                  \
    head3x3_dgrad_kernel<CI, CO><<<blocks, 256, 0, (cudaStream_t)stream>>>(grad_logits, weight, (bf16*)dx->ptr, dx->sN, \
                                                                           dx->sH, dx->sW, x->N, x->H, x->W, tiles_w, tiles_h, n_tiles); \
  else                                                                                                                \
    head3x3_dgrad_mma_kernel<CI, CO><<<blocks, 256, 0, (cudaStream_t)stream>>>(grad_logits, weight, (bf16*)dx->ptr, dx->sN, \
                                                                               dx->sH, dx->sW, x->N, x->H, x->W, tiles_w, tiles_h, n_tiles)
  HEAD_DISPATCH(CALL);
#undef CALL
  PDAC_LAUNCH_CHECK("head3x3_dgrad");
  return PDAC_OK;
}

extern "C" int pdac_head3x3_wgrad(const PdacView* x, const float* grad_logits, int classes, float* dw, float* dbias,
                                  float* scratch, long long scratch_bytes, void* stream) {
  PDAC_REQUIRE(x && grad_logits && dw && scratch, "head3x3_wgrad: null argument (the scratch buffer is required)");
  HeadView xv;
  PDAC_TRY(head_view(x, &xv, "head x"));
  if ((long long)x->N * x->H * x->W == 0) return PDAC_OK;
  const int tiles_w = ceil_div(x->W, kHTW), tiles_h = ceil_div(x->H, kHTH);
  const int n_tiles = tiles_w * tiles_h * x->N;
  const int per = x->C * 9 + 1;
  int blocks = 4 * num_sms();
  if (blocks > n_tiles) blocks = n_tiles;
  PDAC_REQUIRE((long long)blocks * classes * per * (long long)sizeof(float) <= scratch_bytes,
               "head3x3_wgrad: scratch too small (%lld bytes needed)", (long long)blocks * classes * per * 4);
#define CALL(CI, CO)                                                                                                   \
  if (head_scalar())                                                                                                   \
    head3x3_wgrad_kernel<CI, CO><<<blocks, 256, 0, (cudaStream_t)stream>>>(xv, grad_logits, scratch, tiles_w, tiles_h, n_tiles); \
  else                                                                                                                 \
    head3x3_wgrad_mma_kernel<CI, CO><<<blocks, 256, 0, (cudaStream_t)stream>>>(xv, grad_logits, scratch, tiles_w, tiles_h, n_tiles)
  HEAD_DISPATCH(CALL);
#undef CALL
  PDAC_LAUNCH_CHECK("head3x3_wgrad");
  head3x3_wgrad_finalize_kernel<<<ceil_div(classes * per, 32), 256, 0, (cudaStream_t)stream>>>(scratch, blocks, classes,
                                                                                                x->C * 9, dw, dbias);
  PDAC_LAUNCH_CHECK("head3x3_wgrad_finalize");
  return PDAC_OK;
}
