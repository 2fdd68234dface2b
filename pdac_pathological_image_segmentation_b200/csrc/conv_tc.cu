// Tensor-core implicit-GEMM "tap convolution" for sm_100a (forward / data-gradient) and its weight gradient.
//
// Forward:  out[pixel, :] = act( sum_t  W_t[Cout x K] * X[pixel + off_t, K] + bias + addend )
//   GEMM view: M = 128 output pixels (a TWxTHxTN box of the NHWC tensor), N = BLOCK_N output channels,
//   K = taps x concatenated input channels. The A tile of every (tap, 64-channel chunk) is ONE tiled TMA box of the
//   input shifted by the tap offset (out-of-bounds pixels are zero-filled by TMA = the conv's zero padding), landing in
//   shared memory as the canonical K-major swizzled layout tcgen05.mma consumes. The B tile is a 2-D TMA box of the
//   packed weight matrix. fp32 accumulators live in TMEM; the epilogue (4 warps) applies bias/addend/ReLU, rounds to
//   bf16, stages the tile in shared memory, writes it out with coalesced 16-byte stores and accumulates the per-channel
//   sum / sum-of-squares BatchNorm needs.
//   Roles: warp 0 = TMA producer, warp 1 = TMEM allocator + MMA issuer, warps 2..5 = epilogue.
//
// Wgrad:   dw[t][a][b] += sum_pixels X[pixel + off_t, a] * dY[pixel, b]
//   GEMM view: M = 128 rows of (tap, input-channel) pairs ("slab"), N = output channels, K = 128-pixel blocks. Both
//   operands are pixel-major in HBM, i.e. MN-major for the MMA: the same TMA boxes are used, described to tcgen05 with
//   MN-major shared-memory descriptors. A CTA owns up to 512/BLOCK_N slabs (all TMEM columns), loops over its share of
//   the pixel blocks (split-K) and adds its partial sums to dw with fp32 atomics.
//
// Reference ops replaced: nn.Conv2d / nn.ConvTranspose2d forward+backward as instantiated by models/resunet.py:12-21
// and torchvision resnet.py:59-106 (see include/pdac_b200.h).
#include "common.h"
#include "ptx.cuh"
#include <string.h>

namespace pdac {

struct TapTbl {
  int8_t map[PDAC_MAX_TAPS];
  int8_t dh[PDAC_MAX_TAPS];
  int8_t dw[PDAC_MAX_TAPS];
};

struct TcMaps {
  CUtensorMap a[8];  // source views (index = tap.map + cat index)
  CUtensorMap w;     // forward: packed weights; wgrad: dY
};

struct ConvKParams {
  TapTbl taps;
  int n_taps, n_cat;
  int chunks[2];  // CK-channel chunks per concatenated source
  int c0;         // channels of source 0 (K offset of source 1)
  int ck, ksub;   // channels per sub-tile, sub-tiles per pipeline stage (ck*ksub == 64)
  int wrows, block_n, stages;
  uint32_t tmem_cols, idesc, layout_type, pitch;
  int tw_log2, th_log2, tiles_w, tiles_h, tiles_c, total_tiles;
  int N, H, W, Cout;
  bf16* out;
  long long osN, osH, osW;
  const bf16* addend;
  long long asN, asH, asW;
  const float* bias;
  const float* scale;
  float* stats;
  int relu;
  int* err;
};

static constexpr int kEpiWG = 2;                  // epilogue warpgroups of conv_tc_kernel
static constexpr int kThreads = 64 + kEpiWG * 128;  // conv_tc_kernel: TMA warp, MMA warp, 2 x 4 epilogue warps
static constexpr int kWgradThreads = 192;
static constexpr int kMaxStages = 8;

__device__ __forceinline__ float bf16lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf16hi(uint32_t u) { return __uint_as_float(u & 0xffff0000u); }
__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}

// Persistent: one CTA per SM loops over output tiles (tile = n-tile fastest, then w, h, image group). The TMEM accumulator
// is double buffered and the two epilogue warpgroups alternate tiles, so the TMA / MMA main loop of tile i+1 overlaps the
// epilogue of tile i; barrier setup, TMEM allocation and descriptor prefetch are paid once per CTA; BatchNorm partial
// sums live in registers across tiles and are flushed with one atomic per channel per CTA.
__global__ void __launch_bounds__(kThreads, 1) conv_tc_kernel(const __grid_constant__ TcMaps maps,
                                                              const __grid_constant__ ConvKParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[kMaxStages];
  __shared__ __align__(8) uint64_t empty_bar[kMaxStages];
  __shared__ __align__(8) uint64_t tmem_full_bar[2];
  __shared__ __align__(8) uint64_t tmem_empty_bar[2];
  __shared__ uint32_t tmem_base_slot;
  __shared__ long long s_rowoff[kEpiWG][128];

  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  const int a_sub = 128 * p.pitch;         // bytes of one A sub-tile
  const int b_sub = p.block_n * p.pitch;   // bytes of one B sub-tile
  const int a_stage = a_sub * p.ksub;      // 16 KB
  const int stage_bytes = a_stage + b_sub * p.ksub;
  const int total_sub = p.n_taps * (p.chunks[0] + (p.n_cat > 1 ? p.chunks[1] : 0));
  const int n_iters = (total_sub + p.ksub - 1) / p.ksub;
  const int TW = 1 << p.tw_log2, TH = 1 << p.th_log2;

  if (threadIdx.x == 0) {
    for (int i = 0; i < p.stages; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full_bar[i], 1);
      mbar_init(&tmem_empty_bar[i], 128);
    }
    fence_barrier_init();
  }
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&maps.w);
    tma_prefetch_desc(&maps.a[0]);
  }
  if (warp == 1) {
    tmem_alloc(&tmem_base_slot, p.tmem_cols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_slot;

#define DECODE_TILE(tile)                                                         \
  int bx_ = (tile);                                                               \
  const int n_tile = bx_ % p.tiles_c;                                             \
  bx_ /= p.tiles_c;                                                               \
  const int tile_w = bx_ % p.tiles_w;                                             \
  bx_ /= p.tiles_w;                                                               \
  const int tile_h = bx_ % p.tiles_h;                                             \
  const int tile_n = bx_ / p.tiles_h;                                             \
  const int w0 = tile_w * TW, h0 = tile_h * TH, img0 = tile_n << (7 - p.tw_log2 - p.th_log2); \
  const int n0 = n_tile * p.block_n;

  if (warp == 0) {
    if (lane == 0) {
      int it = 0;
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
        DECODE_TILE(tile)
        int t = 0, s = 0, c = 0;
        for (int k = 0; k < n_iters; ++k, ++it) {
          const int stage = it % p.stages;
          mbar_wait(&empty_bar[stage], ((it / p.stages) & 1) ^ 1, p.err);
          const int nsub = min(p.ksub, total_sub - k * p.ksub);
          mbar_expect_tx(&full_bar[stage], (uint32_t)(nsub * (a_sub + b_sub)));
          uint8_t* a_dst = smem + (size_t)stage * stage_bytes;
          uint8_t* b_dst = a_dst + a_stage;
          for (int j = 0; j < nsub; ++j) {
            const int mi = p.taps.map[t] + s;
            tma_load_4d(a_dst + j * a_sub, &maps.a[mi], &full_bar[stage], c * p.ck, w0 + p.taps.dw[t],
                        h0 + p.taps.dh[t], img0);
            tma_load_2d(b_dst + j * b_sub, &maps.w, &full_bar[stage], (s ? p.c0 : 0) + c * p.ck, t * p.wrows + n0);
            if (++c == p.chunks[s]) {
              c = 0;
              if (++s == p.n_cat) {
                s = 0;
                ++t;
              }
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t sbo = 8 * p.pitch;
      int it = 0, lt = 0;
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++lt) {
        const int as = lt & 1;
        mbar_wait(&tmem_empty_bar[as], ((lt >> 1) & 1) ^ 1, p.err);
        tc_fence_after();
        const uint32_t tacc = tmem_base + (uint32_t)(as * p.block_n);
        uint32_t accum = 0;
        for (int k = 0; k < n_iters; ++k, ++it) {
          const int stage = it % p.stages;
          mbar_wait(&full_bar[stage], (it / p.stages) & 1, p.err);
          tc_fence_after();
          const int nsub = min(p.ksub, total_sub - k * p.ksub);
          const uint32_t a_base = smem_u32(smem + (size_t)stage * stage_bytes);
          const uint32_t b_base = a_base + a_stage;
          for (int j = 0; j < nsub; ++j) {
            for (int kk = 0; kk < p.ck / 16; ++kk) {
              const uint64_t da = make_smem_desc(a_base + j * a_sub + kk * 32, 16, sbo, p.layout_type);
              const uint64_t db = make_smem_desc(b_base + j * b_sub + kk * 32, 16, sbo, p.layout_type);
              umma_bf16(tacc, da, db, p.idesc, accum);
              accum = 1;
            }
          }
          umma_commit(&empty_bar[stage]);
        }
        umma_commit(&tmem_full_bar[as]);
      }
    }
    __syncwarp();
  } else {
    // ---------------- epilogue: two warpgroups (warps 2..5 and 6..9) alternate tiles; TMEM lane quarter = warp % 4
    const int wg = (warp - 2) >> 2;
    const int q = warp & 3;
    const int m = q * 32 + lane;
    const int e = (threadIdx.x - 64) & 127;  // cooperative index inside the warpgroup
    const int spitch = p.block_n * 2 + 16;   // staging row pitch in bytes (padded against bank conflicts)
    uint8_t* stg = smem + (size_t)p.stages * stage_bytes + (size_t)wg * 128 * spitch;
    const int tw = m & (TW - 1), th = (m >> p.tw_log2) & (TH - 1), tn = m >> (p.tw_log2 + p.th_log2);
    const uint32_t bar_id = 1 + wg;
    // BatchNorm partial sums of this thread's channel pair, kept across tiles
    const int pairs = p.block_n >> 1;
    const int groups = 128 / pairs > 0 ? 128 / pairs : 1;
    const int pr = e % pairs, sg = e / pairs;
    const bool stat_thread = p.stats && sg < groups && pairs <= 128;
    float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
    int stat_n0 = -1;
    // Deterministic flush of the channel-pair partials of N tile `fn0` (called by all 128 threads of the warpgroup):
    // partials -> staging -> thread `pr` adds the `groups` row-groups in index order into this warpgroup's own row of
    // the [rows][2][Cout] table (row = CTA * 2 + warpgroup: a single writer per element, so plain adds).
    auto flush_stats = [&](int fn0) {
      asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");  // staging no longer read by the write-out
      float4* slot = reinterpret_cast<float4*>(stg);
      if (stat_thread) slot[sg * pairs + pr] = make_float4(s0, s1, q0, q1);
      asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
      if (e < pairs && pairs <= 128 && 2 * e < min(p.block_n, p.Cout - fn0)) {
        float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int g = 0; g < groups; ++g) {
          const float4 v = slot[g * pairs + e];
          t.x += v.x;
          t.y += v.y;
          t.z += v.z;
          t.w += v.w;
        }
        float* st = p.stats + (size_t)(blockIdx.x * kEpiWG + wg) * 2 * p.Cout + fn0 + 2 * e;
        st[0] += t.x;
        st[1] += t.y;
        st[p.Cout] += t.z;
        st[p.Cout + 1] += t.w;
      }
    };

    int lt = 0;
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++lt) {
      if ((lt & 1) != wg) continue;
      DECODE_TILE(tile)
      const int w = w0 + tw, h = h0 + th, img = img0 + tn;
      const bool valid = (w < p.W) && (h < p.H) && (img < p.N);
      const long long ooff = (long long)img * p.osN + (long long)h * p.osH + (long long)w * p.osW;
      const long long aoff = (long long)img * p.asN + (long long)h * p.asH + (long long)w * p.asW;
      const int nch = min(p.block_n, p.Cout - n0);  // valid channels of this N tile
      if (p.stats && stat_n0 != n0) {  // uniform over the warpgroup
        if (stat_n0 >= 0) flush_stats(stat_n0);
        s0 = s1 = q0 = q1 = 0.f;
        stat_n0 = n0;
      }
      asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");  // previous tile's staging fully consumed
      s_rowoff[wg][m] = valid ? ooff : -1;

      mbar_wait(&tmem_full_bar[wg], (lt >> 1) & 1, p.err);
      tc_fence_after();
      const uint32_t tacc = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(wg * p.block_n);
      for (int c = 0; c < p.block_n; c += 16) {
        uint32_t v[16];
        tmem_ld16(tacc + (uint32_t)c, v);
        tmem_ld_wait();
        float f[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) f[i] = __uint_as_float(v[i]);
        if (p.scale) {
#pragma unroll
          for (int i = 0; i < 16; ++i)
            if (c + i < nch) f[i] *= __ldg(p.scale + n0 + c + i);
        }
        if (p.bias) {
#pragma unroll
          for (int i = 0; i < 16; ++i)
            if (c + i < nch) f[i] += __ldg(p.bias + n0 + c + i);
        }
        if (p.addend && valid) {
          const bf16* ap = p.addend + aoff + n0 + c;
          if ((p.Cout & 7) == 0) {
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {
              if (c + hh * 8 < nch) {
                const uint4 u = __ldg(reinterpret_cast<const uint4*>(ap + hh * 8));
                const uint32_t uu[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                  f[hh * 8 + 2 * i] += bf16lo(uu[i]);
                  f[hh * 8 + 2 * i + 1] += bf16hi(uu[i]);
                }
              }
            }
          } else {
#pragma unroll
            for (int i = 0; i < 16; ++i)
              if (c + i < nch) f[i] += __bfloat162float(ap[i]);
          }
        }
        if (p.relu) {
#pragma unroll
          for (int i = 0; i < 16; ++i) f[i] = fmaxf(f[i], 0.f);
        }
        uint4 o0, o1;
        o0.x = pack_bf16(f[0], f[1]);
        o0.y = pack_bf16(f[2], f[3]);
        o0.z = pack_bf16(f[4], f[5]);
        o0.w = pack_bf16(f[6], f[7]);
        o1.x = pack_bf16(f[8], f[9]);
        o1.y = pack_bf16(f[10], f[11]);
        o1.z = pack_bf16(f[12], f[13]);
        o1.w = pack_bf16(f[14], f[15]);
        uint4* sp = reinterpret_cast<uint4*>(stg + (size_t)m * spitch + c * 2);
        sp[0] = o0;
        sp[1] = o1;
      }
      // the accumulator stage is free for the MMA warp as soon as every thread has drained its TMEM lanes
      tc_fence_before();
      mbar_arrive(&tmem_empty_bar[wg]);
      asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");

      // cooperative coalesced write-out
      if ((p.Cout & 7) == 0) {
        const int cpr = p.block_n >> 3;
        const int cvalid = nch >> 3;
        for (int id = e; id < 128 * cpr; id += 128) {
          const int row = id / cpr, ch = id - row * cpr;
          const long long off = s_rowoff[wg][row];
          if (off < 0 || ch >= cvalid) continue;
          const uint4 val = *reinterpret_cast<const uint4*>(stg + (size_t)row * spitch + ch * 16);
          *reinterpret_cast<uint4*>(p.out + off + n0 + ch * 8) = val;
        }
      } else {
        for (int id = e; id < 128 * nch; id += 128) {
          const int row = id / nch, ch = id - row * nch;
          const long long off = s_rowoff[wg][row];
          if (off < 0) continue;
          p.out[off + n0 + ch] = *reinterpret_cast<const bf16*>(stg + (size_t)row * spitch + ch * 2);
        }
      }
      // per-channel sum / sum of squares of the stored values (BatchNorm batch statistics)
      if (stat_thread && (nch & 1) == 0 && 2 * pr < nch) {
        for (int row = sg; row < 128; row += groups) {
          if (s_rowoff[wg][row] < 0) continue;
          const uint32_t u = *reinterpret_cast<const uint32_t*>(stg + (size_t)row * spitch + pr * 4);
          const float x0 = bf16lo(u), x1 = bf16hi(u);
          s0 += x0;
          q0 += x0 * x0;
          s1 += x1;
          q1 += x1 * x1;
        }
      }
    }
    if (p.stats && stat_n0 >= 0) flush_stats(stat_n0);
  }
#undef DECODE_TILE

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, p.tmem_cols);
  }
}

// ------------------------------------------------------------------------------------------------ wgrad kernel

struct WgradKParams {
  TapTbl taps;
  int n_taps, n_cat;
  int chunks[2];
  int c0, chunks_total;
  int ck, subs;  // channels per A box, boxes per 128-row slab
  int n_sub_total, n_slabs, slabs_per_cta;
  int ckb, nboxb;  // dY box channels, boxes per B tile
  int block_n, stages_a;
  uint32_t tmem_cols, idesc, layout_a, pitch_a, layout_b, pitch_b;
  int tw_log2, th_log2, tiles_w, tiles_h, n_kblocks;
  float* dw;
  long long tap_stride, a_stride, b_stride;
  int Cout;
  float* part;  // deterministic mode, see WhParams in wgrad_halo.cu
  long long part_stride;
  int part_A;
  int* err;
};

__global__ void __launch_bounds__(kWgradThreads) wgrad_tc_kernel(const __grid_constant__ TcMaps maps,
                                                            const __grid_constant__ WgradKParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t fullA[kMaxStages];
  __shared__ __align__(8) uint64_t emptyA[kMaxStages];
  __shared__ __align__(8) uint64_t fullB[2];
  __shared__ __align__(8) uint64_t emptyB[2];
  __shared__ __align__(8) uint64_t tmem_full_bar;
  __shared__ uint32_t tmem_base_slot;

  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  const int a_box = 128 * p.pitch_a;     // one [128 px][ck] box
  const int a_slab = a_box * p.subs;     // 32 KB
  const int b_box = 128 * p.pitch_b;
  const int b_tile = b_box * p.nboxb;
  uint8_t* smem_b = smem + (size_t)p.stages_a * a_slab;

  const int slab0 = blockIdx.y * p.slabs_per_cta;
  const int nslab = min(p.slabs_per_cta, p.n_slabs - slab0);
  const int n0 = blockIdx.z * p.block_n;
  // this CTA's pixel blocks: kb = blockIdx.x, blockIdx.x + gridDim.x, ...
  const int n_my_kb = (p.n_kblocks - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

  if (threadIdx.x == 0) {
    for (int i = 0; i < p.stages_a; ++i) {
      mbar_init(&fullA[i], 1);
      mbar_init(&emptyA[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&fullB[i], 1);
      mbar_init(&emptyB[i], 1);
    }
    mbar_init(&tmem_full_bar, 1);
    fence_barrier_init();
  }
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&maps.w);
    tma_prefetch_desc(&maps.a[0]);
  }
  if (warp == 1) {
    tmem_alloc(&tmem_base_slot, p.tmem_cols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_slot;
  const int TW = 1 << p.tw_log2, TH = 1 << p.th_log2;

  if (warp == 0) {
    if (lane == 0) {
      int ita = 0;
      for (int i = 0; i < n_my_kb; ++i) {
        int kb = blockIdx.x + i * gridDim.x;
        const int tile_w = kb % p.tiles_w;
        kb /= p.tiles_w;
        const int tile_h = kb % p.tiles_h;
        const int tile_n = kb / p.tiles_h;
        const int w0 = tile_w * TW, h0 = tile_h * TH, img0 = tile_n << (7 - p.tw_log2 - p.th_log2);
        // dY tile for this pixel block
        const int bs = i & 1;
        mbar_wait(&emptyB[bs], ((i >> 1) & 1) ^ 1, p.err);
        mbar_expect_tx(&fullB[bs], (uint32_t)b_tile);
        for (int j = 0; j < p.nboxb; ++j)
          tma_load_4d(smem_b + (size_t)bs * b_tile + (size_t)j * b_box, &maps.w, &fullB[bs], n0 + j * p.ckb, w0, h0,
                      img0);
        for (int g = 0; g < nslab; ++g, ++ita) {
          const int stage = ita % p.stages_a;
          mbar_wait(&emptyA[stage], ((ita / p.stages_a) & 1) ^ 1, p.err);
          const int u0 = (slab0 + g) * p.subs;
          const int nsub = min(p.subs, p.n_sub_total - u0);
          mbar_expect_tx(&fullA[stage], (uint32_t)(nsub * a_box));
          for (int j = 0; j < nsub; ++j) {
            const int u = u0 + j;
            const int t = u / p.chunks_total;
            int c = u - t * p.chunks_total;
            int s = 0;
            if (c >= p.chunks[0]) {
              c -= p.chunks[0];
              s = 1;
            }
            tma_load_4d(smem + (size_t)stage * a_slab + (size_t)j * a_box, &maps.a[p.taps.map[t] + s], &fullA[stage],
                        c * p.ck, w0 + p.taps.dw[t], h0 + p.taps.dh[t], img0);
          }
        }
      }
    }
  } else if (warp == 1) {
    // warp-converged walk, one elected lane issues (see conv_halo.cu: the issue path must stay below ~40 cycles / MMA)
    {
      const uint32_t leader = elect_one();
      const uint32_t idesc = p.idesc;
      const uint32_t hi_a = smem_desc_hi(8 * p.pitch_a, p.layout_a), hi_b = smem_desc_hi(8 * p.pitch_b, p.layout_b);
      const uint32_t ka16 = (16 * p.pitch_a) >> 4, kb16 = (16 * p.pitch_b) >> 4;  // 16 pixels along K, in 16-byte units
      const uint32_t a_lo0 = smem_desc_lo(smem_u32(smem), (uint32_t)a_box);
      const uint32_t b_lo0 = smem_desc_lo(smem_u32(smem_b), (uint32_t)b_box);
      const uint32_t a_slab16 = (uint32_t)a_slab >> 4, b_tile16 = (uint32_t)b_tile >> 4;
      const uint32_t block_n = (uint32_t)p.block_n;
      const int stages_a = p.stages_a;
      int ita = 0;
      for (int i = 0; i < n_my_kb; ++i) {
        const int bs = i & 1;
        mbar_wait(&fullB[bs], (i >> 1) & 1, p.err);
        tc_fence_after();
        const uint32_t b_lo = b_lo0 + (uint32_t)bs * b_tile16;
        for (int g = 0; g < nslab; ++g, ++ita) {
          const int stage = ita % stages_a;
          mbar_wait(&fullA[stage], (ita / stages_a) & 1, p.err);
          tc_fence_after();
          if (leader) {
            const uint32_t a_lo = a_lo0 + (uint32_t)stage * a_slab16;
            const uint32_t td = tmem_base + (uint32_t)g * block_n;
            umma_bf16_lohi(td, a_lo, hi_a, b_lo, hi_b, idesc, i > 0 ? 1u : 0u);
#pragma unroll
            for (int k = 1; k < 8; ++k)  // 8 x 16 pixels
              umma_bf16_lohi(td, a_lo + k * ka16, hi_a, b_lo + k * kb16, hi_b, idesc, 1u);
            umma_commit(&emptyA[stage]);
          }
        }
        if (leader) umma_commit(&emptyB[bs]);
        __syncwarp();
      }
      if (leader) umma_commit(&tmem_full_bar);
    }
    __syncwarp();
  } else {
    const int q = warp & 3;
    const int m = q * 32 + lane;
    if (n_my_kb > 0) {
      mbar_wait(&tmem_full_bar, 0, p.err);
      tc_fence_after();
      const int sub = m / p.ck, within = m - sub * p.ck;
      const int nch = min(p.block_n, p.Cout - n0);
      for (int g = 0; g < nslab; ++g) {
        const int u = (slab0 + g) * p.subs + sub;
        const bool rvalid = u < p.n_sub_total;
        int t = 0, a_ch = 0;
        if (rvalid) {
          t = u / p.chunks_total;
          int c = u - t * p.chunks_total;
          int koff = 0;
          if (c >= p.chunks[0]) {
            c -= p.chunks[0];
            koff = p.c0;
          }
          a_ch = koff + c * p.ck + within;
        }
        float* base = p.dw + (long long)t * p.tap_stride + (long long)a_ch * p.a_stride + (long long)n0 * p.b_stride;
        float* pbase = p.part ? p.part + (long long)blockIdx.x * p.part_stride +
                                    ((long long)t * p.Cout + n0) * p.part_A + a_ch
                              : nullptr;
        for (int c = 0; c < p.block_n; c += 16) {
          uint32_t v[16];
          tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(g * p.block_n + c), v);
          tmem_ld_wait();
          if (rvalid) {
            if (pbase) {
#pragma unroll
              for (int i = 0; i < 16; ++i)
                if (c + i < nch) pbase[(long long)(c + i) * p.part_A] = __uint_as_float(v[i]);
            } else {
#pragma unroll
              for (int i = 0; i < 16; ++i)
                if (c + i < nch) atomicAdd(base + (long long)(c + i) * p.b_stride, __uint_as_float(v[i]));
            }
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, p.tmem_cols);
  }
}

// ------------------------------------------------------------------------------------------------ host side

static int pick_ck(int n_cat, const PdacView* v0) {
  int ck = 64;
  for (int s = 0; s < n_cat; ++s) {
    while (ck > 8 && (v0[s].C % ck)) ck >>= 1;
  }
  return ck;
}

struct Tiling {
  int tw, th, tn, tiles_w, tiles_h, tiles_n;
};
static Tiling pick_tiling(int N, int H, int W) {
  Tiling t;
  t.tw = pow2_ceil(W) < 16 ? pow2_ceil(W) : 16;
  int rest = 128 / t.tw;
  t.th = pow2_ceil(H) < rest ? pow2_ceil(H) : rest;
  t.tn = rest / t.th;
  t.tiles_w = ceil_div(W, t.tw);
  t.tiles_h = ceil_div(H, t.th);
  t.tiles_n = ceil_div(N, t.tn);
  return t;
}

static int fill_taps(const PdacTapConv* tc, int n_views, const PdacView* views, const PdacView* ref, TapTbl* tbl) {
  PDAC_REQUIRE(tc->n_taps >= 1 && tc->n_taps <= PDAC_MAX_TAPS, "n_taps=%d out of range", tc->n_taps);
  PDAC_REQUIRE(tc->n_cat == 1 || tc->n_cat == 2, "n_cat must be 1 or 2");
  PDAC_REQUIRE(n_views >= 1 && n_views <= 8, "n_views=%d out of range (1..8)", n_views);
  for (int t = 0; t < tc->n_taps; ++t) {
    const PdacTap& tp = tc->taps[t];
    PDAC_REQUIRE(tp.src >= 0 && tp.src + tc->n_cat <= n_views, "tap %d: source index %d out of range", t, tp.src);
    PDAC_REQUIRE(tp.dh >= -127 && tp.dh <= 127 && tp.dw >= -127 && tp.dw <= 127, "tap %d: offset too large", t);
    for (int s = 0; s < tc->n_cat; ++s) {
      PDAC_REQUIRE(views[tp.src + s].C == views[tc->taps[0].src + s].C, "tap %d: channel count differs between taps",
                   t);
      PDAC_REQUIRE(views[tp.src + s].N == ref->N, "tap %d: batch size differs from the output's", t);
    }
    tbl->map[t] = (int8_t)tp.src;
    tbl->dh[t] = (int8_t)tp.dh;
    tbl->dw[t] = (int8_t)tp.dw;
  }
  return PDAC_OK;
}

int conv_halo_try(const PdacTapConv* tc, int n_views, const PdacView* views, const void* w_packed, const float* bias,
                  const PdacView* addend, const PdacView* out, float* stats, void* stream, int* handled);

int wgrad_halo_try(const PdacTapConv* tc, int n_views, const PdacView* views, const PdacView* dy, float* dw,
                   int64_t tap_stride, int64_t a_stride, int64_t b_stride, int b_valid, float* scratch,
                   long long scratch_bytes, void* stream, int* handled);

// Deterministic split-K: with a scratch buffer the split-K shares are stored as partial tensors [share][tap][B][A]
// (A = concatenated input channels, fastest) and added in share order by wgrad_reduce_kernel. The number of shares is
// capped by the scratch size; a single share still goes through the scratch (dw is accumulated into, not overwritten).
int wgrad_plan_partials(float* scratch, long long scratch_bytes, int n_taps, int A, int B, int* splits, float** part,
                        long long* part_stride) {
  *part = nullptr;
  *part_stride = 0;
  if (!scratch) return PDAC_OK;
  PDAC_REQUIRE(((uintptr_t)scratch & 15) == 0, "wgrad scratch must be 16-byte aligned");
  const long long per = (((long long)n_taps * A * B + 3) / 4) * 4;
  const long long fit = scratch_bytes / (per * (long long)sizeof(float));
  PDAC_REQUIRE(fit >= 1, "wgrad scratch too small: %lld bytes given, %lld needed per split-K share", scratch_bytes,
               per * (long long)sizeof(float));
  if (*splits > fit) *splits = (int)fit;
  *part = scratch;
  *part_stride = per;
  return PDAC_OK;
}

__global__ void __launch_bounds__(256) wgrad_reduce_kernel(const float* __restrict__ part, int splits,
                                                           long long part_stride, int B, int A, long long total,
                                                           float* __restrict__ dw, long long tap_stride,
                                                           long long a_stride, long long b_stride) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    float acc = 0.f;
    for (int s = 0; s < splits; ++s) acc += part[(long long)s * part_stride + i];
    const int a = (int)(i % A);
    const long long r = i / A;
    const int b = (int)(r % B);
    const long long t = r / B;
    dw[t * tap_stride + (long long)a * a_stride + (long long)b * b_stride] += acc;
  }
}

// Many shares of a small tensor (the 16 / 32 / 64-channel layers: up to 296 shares of a few thousand elements): one
// thread per element walked the shares serially, ~75 dependent L2 round trips = 20 us per call, 0.6 ms per step. Here
// warp k of a block adds the shares k, k + K, ... of 32 consecutive elements (coalesced 128-byte rows) and warp 0 adds
// the K sub-sums in index order: still a fixed order, K times fewer round trips.
template <int K>
__global__ void __launch_bounds__(32 * K) wgrad_reduce_par_kernel(const float* __restrict__ part, int splits,
                                                                  long long part_stride, int B, int A, long long total,
                                                                  float* __restrict__ dw, long long tap_stride,
                                                                  long long a_stride, long long b_stride) {
  __shared__ float red[K][32];
  const int lane = threadIdx.x & 31, k = threadIdx.x >> 5;
  for (long long base = (long long)blockIdx.x * 32; base < total; base += (long long)gridDim.x * 32) {
    const long long i = base + lane;
    float acc = 0.f;
    if (i < total) {
#pragma unroll 4
      for (int s = k; s < splits; s += K) acc += part[(long long)s * part_stride + i];
    }
    red[k][lane] = acc;
    __syncthreads();
    if (k == 0 && i < total) {
      float v = red[0][lane];
#pragma unroll
      for (int kk = 1; kk < K; ++kk) v += red[kk][lane];
      const int a = (int)(i % A);
      const long long r = i / A;
      const int b = (int)(r % B);
      const long long t = r / B;
      dw[t * tap_stride + (long long)a * a_stride + (long long)b * b_stride] += v;
    }
    __syncthreads();
  }
}

// Large 3x3 tensors into the state_dict layout [B][A][9] (tap fastest): with one thread per partial element every store
// was a 4-byte read-modify-write 36 bytes from its neighbour's. Here a warp owns 32 consecutive input channels of one
// output channel: coalesced 128-byte reads per (share, tap), the 288 results transposed through shared memory and
// added to 288 CONSECUTIVE floats of dw. Shares are still added in index order.
__global__ void __launch_bounds__(256) wgrad_reduce_oihw9_kernel(const float* __restrict__ part, int splits,
                                                                 long long part_stride, int B, int A, int units,
                                                                 float* __restrict__ dw) {
  __shared__ float tr[8][288];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int per_b = A >> 5;
  for (int u = blockIdx.x * 8 + w; u < units; u += gridDim.x * 8) {
    const int b = u / per_b, a0 = (u - b * per_b) << 5;
    float acc[9];
#pragma unroll
    for (int t = 0; t < 9; ++t) acc[t] = 0.f;
    const float* src = part + ((long long)b * A + a0 + lane);
    for (int sI = 0; sI < splits; ++sI, src += part_stride) {
#pragma unroll
      for (int t = 0; t < 9; ++t) acc[t] += src[(long long)t * B * A];
    }
#pragma unroll
    for (int t = 0; t < 9; ++t) tr[w][lane * 9 + t] = acc[t];
    __syncwarp();
    float* dst = dw + ((long long)b * A + a0) * 9;
#pragma unroll
    for (int k = 0; k < 9; ++k) dst[k * 32 + lane] += tr[w][k * 32 + lane];
    __syncwarp();
  }
}

int wgrad_reduce(const float* part, int splits, long long part_stride, int n_taps, int B, int A, float* dw,
                 long long tap_stride, long long a_stride, long long b_stride, void* stream) {
  const long long total = (long long)n_taps * B * A;
  // measured (r3i): pays from 256 -> 256 upwards (0.340 -> 0.322 ms for four, 512 -> 512 0.266 -> 0.230 for three), loses below
  if (total >= 400000 && n_taps == 9 && (A & 31) == 0 && tap_stride == 1 && a_stride == 9 && b_stride == (long long)A * 9) {
    const int units = B * (A >> 5);
    int blocks = (units + 7) / 8;
    if (blocks > num_sms() * 8) blocks = num_sms() * 8;
    wgrad_reduce_oihw9_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(part, splits, part_stride, B, A, units, dw);
    PDAC_LAUNCH_CHECK("wgrad_reduce_oihw9_kernel");
    return PDAC_OK;
  }
  if (splits > 8 && total <= 65536) {  // measured: larger tensors are faster with one thread per element (r2z)
    long long blocks = (total + 31) / 32;
    if (blocks > (long long)num_sms() * 8) blocks = (long long)num_sms() * 8;
    wgrad_reduce_par_kernel<8><<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(part, splits, part_stride, B, A, total, dw,
                                                                            tap_stride, a_stride, b_stride);
    PDAC_LAUNCH_CHECK("wgrad_reduce_par_kernel");
    return PDAC_OK;
  }
  long long blocks = (total + 255) / 256;
  if (blocks > (long long)num_sms() * 8) blocks = (long long)num_sms() * 8;
  wgrad_reduce_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(part, splits, part_stride, B, A, total, dw,
                                                                    tap_stride, a_stride, b_stride);
  PDAC_LAUNCH_CHECK("wgrad_reduce_kernel");
  return PDAC_OK;
}

static int check_dev_ptr(const void* p, const char* what) {
  cudaPointerAttributes at;
  cudaError_t e = cudaPointerGetAttributes(&at, p);
  if (e != cudaSuccess) return cuda_fail(e, "cudaPointerGetAttributes");
  if (at.type != cudaMemoryTypeDevice && at.type != cudaMemoryTypeManaged)
    return fail(PDAC_E_INVALID, "%s is not a device pointer (no CPU fallback)", what);
  return PDAC_OK;
}

}  // namespace pdac

using namespace pdac;

extern "C" int pdac_tapconv_fwd(const PdacTapConv* tc, int n_views, const PdacView* views, const void* w_packed,
                                const float* bias, const PdacView* addend, const PdacView* out, float* stats,
                                void* stream) {
  PDAC_REQUIRE(tc && views && w_packed && out, "null argument");
  PDAC_TRY(check_dev_ptr(out->ptr, "out"));
  TcMaps maps;
  ConvKParams p;
  memset(&p, 0, sizeof(p));
  PDAC_TRY(fill_taps(tc, n_views, views, out, &p.taps));
  {
    // stride-1 multi-tap convolutions whose taps share one source lattice take the halo-tile kernel (conv_halo.cu)
    int handled = 0;
    PDAC_TRY(conv_halo_try(tc, n_views, views, w_packed, bias, addend, out, stats, stream, &handled));
    if (handled) return PDAC_OK;
  }
  const PdacView* v0 = views + tc->taps[0].src;
  const int ck = pick_ck(tc->n_cat, v0);
  PDAC_REQUIRE(ck >= 16, "input channels must be multiples of 16 (got %d%s)", v0[0].C, tc->n_cat > 1 ? " + cat" : "");
  p.n_taps = tc->n_taps;
  p.n_cat = tc->n_cat;
  p.chunks[0] = v0[0].C / ck;
  p.chunks[1] = tc->n_cat > 1 ? v0[1].C / ck : 0;
  p.c0 = v0[0].C;
  const int ktot = v0[0].C + (tc->n_cat > 1 ? v0[1].C : 0);
  p.ck = ck;
  p.ksub = 64 / ck;
  p.wrows = tc->wrows;
  const int cout = out->C;
  int block_n = cout <= 128 ? ((cout + 15) / 16) * 16 : 128;
  PDAC_REQUIRE(tc->wrows >= ceil_div(cout, block_n) * block_n, "wrows=%d too small for Cout=%d (tile %d)", tc->wrows,
               cout, block_n);
  p.block_n = block_n;
  p.pitch = ck * 2;
  p.layout_type = swizzle_layout_type(p.pitch);
  p.idesc = make_idesc_bf16(128, block_n, 0, 0);
  p.tmem_cols = pow2_ceil(2 * block_n < 32 ? 32 : 2 * block_n);  // double-buffered accumulator
  const int stage_bytes = 128 * 128 + block_n * 128;
  const int staging = kEpiWG * 128 * (block_n * 2 + 16);
  int stages = (200 * 1024 - staging) / stage_bytes;
  if (stages > kMaxStages) stages = kMaxStages;
  if (stages < 2) stages = 2;
  p.stages = stages;
  const Tiling tl = pick_tiling(out->N, out->H, out->W);
  p.tw_log2 = ilog2(tl.tw);
  p.th_log2 = ilog2(tl.th);
  p.tiles_w = tl.tiles_w;
  p.tiles_h = tl.tiles_h;
  p.N = out->N;
  p.H = out->H;
  p.W = out->W;
  p.Cout = cout;
  p.out = (bf16*)out->ptr;
  p.osN = out->sN;
  p.osH = out->sH;
  p.osW = out->sW;
  if (addend) {
    PDAC_REQUIRE(addend->N == out->N && addend->H == out->H && addend->W == out->W && addend->C == out->C,
                 "addend shape differs from out");
    p.addend = (const bf16*)addend->ptr;
    p.asN = addend->sN;
    p.asH = addend->sH;
    p.asW = addend->sW;
  }
  if ((cout & 7) == 0)
    PDAC_REQUIRE(((uintptr_t)out->ptr & 15) == 0 && out->sW % 8 == 0 && out->sH % 8 == 0 && out->sN % 8 == 0 &&
                     (!addend || (((uintptr_t)addend->ptr & 15) == 0 && addend->sW % 8 == 0 && addend->sH % 8 == 0 &&
                                  addend->sN % 8 == 0)),
                 "out/addend view must be 16-byte aligned");
  p.bias = bias;
  p.scale = epi_scale();
  p.stats = stats;
  p.relu = tc->relu;
  p.err = device_err_flag();
  for (int i = 0; i < n_views; ++i) PDAC_TRY(encode_view_map(&maps.a[i], &views[i], ck, tl.tw, tl.th, tl.tn));
  {
    uint64_t dims[2] = {(uint64_t)ktot, (uint64_t)tc->n_taps * tc->wrows};
    uint64_t strides[1] = {(uint64_t)ktot * 2};
    uint32_t box[2] = {(uint32_t)ck, (uint32_t)block_n};
    PDAC_TRY(encode_tiled_bf16(&maps.w, w_packed, 2, dims, strides, box, p.pitch));
  }
  p.tiles_c = ceil_div(cout, block_n);
  p.total_tiles = tl.tiles_w * tl.tiles_h * tl.tiles_n * p.tiles_c;
  size_t smem = (size_t)stages * stage_bytes + staging + 1024;
  static bool attr_set = false;
  if (!attr_set) {
    PDAC_CUDA(cudaFuncSetAttribute(conv_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(210 * 1024)));
    attr_set = true;
  }
  int ctas = num_sms();
  if (ctas > p.total_tiles) ctas = p.total_tiles;
  if (stats && ctas * kEpiWG > PDAC_STAT_ROWS) ctas = PDAC_STAT_ROWS / kEpiWG;
  conv_tc_kernel<<<ctas, kThreads, smem, (cudaStream_t)stream>>>(maps, p);
  PDAC_LAUNCH_CHECK("conv_tc_kernel");
  note_kernel(KID_CONV_TC);
  return PDAC_OK;
}

extern "C" int pdac_tapconv_wgrad(const PdacTapConv* tc, int n_views, const PdacView* views, const PdacView* dy,
                                  float* dw, int64_t tap_stride, int64_t a_stride, int64_t b_stride, int b_valid,
                                  float* scratch, long long scratch_bytes, void* stream) {
  PDAC_REQUIRE(tc && views && dy && dw, "null argument");
  PDAC_REQUIRE(b_valid >= 0 && b_valid <= dy->C, "b_valid=%d out of range (dY has %d channels)", b_valid, dy->C);
  PDAC_TRY(check_dev_ptr(dw, "dw"));
  TcMaps maps;
  WgradKParams p;
  memset(&p, 0, sizeof(p));
  PDAC_TRY(fill_taps(tc, n_views, views, dy, &p.taps));
  {
    // stride-1 multi-tap weight gradients with <= 128 output channels take the halo-tile kernel (wgrad_halo.cu)
    int handled = 0;
    PDAC_TRY(wgrad_halo_try(tc, n_views, views, dy, dw, tap_stride, a_stride, b_stride, b_valid, scratch,
                            scratch_bytes, stream, &handled));
    if (handled) return PDAC_OK;
  }
  const PdacView* v0 = views + tc->taps[0].src;
  const int ck = pick_ck(tc->n_cat, v0);
  PDAC_REQUIRE(ck >= 16, "input channels must be multiples of 16 (got %d)", v0[0].C);
  PDAC_REQUIRE(dy->C % 16 == 0, "dY channels must be a multiple of 16 (got %d)", dy->C);
  p.n_taps = tc->n_taps;
  p.n_cat = tc->n_cat;
  p.chunks[0] = v0[0].C / ck;
  p.chunks[1] = tc->n_cat > 1 ? v0[1].C / ck : 0;
  p.chunks_total = p.chunks[0] + p.chunks[1];
  p.c0 = v0[0].C;
  p.ck = ck;
  p.subs = 128 / ck;
  p.n_sub_total = p.n_taps * p.chunks_total;
  p.n_slabs = ceil_div(p.n_sub_total, p.subs);
  const int cout = dy->C;
  int ckb = 64;
  while (cout % ckb) ckb >>= 1;
  int block_n = cout <= 256 ? cout : 256;
  if (cout > 256) PDAC_REQUIRE(cout % 256 == 0, "dY channels > 256 must be a multiple of 256 (got %d)", cout);
  p.ckb = ckb;
  p.nboxb = block_n / ckb;
  p.block_n = block_n;
  int spc = 512 / block_n;
  if (spc > p.n_slabs) spc = p.n_slabs;
  // keep enough CTAs in flight: shrink the slab group while the grid would be too small
  const Tiling tl = pick_tiling(dy->N, dy->H, dy->W);
  p.n_kblocks = tl.tiles_w * tl.tiles_h * tl.tiles_n;
  p.slabs_per_cta = spc;
  p.tmem_cols = pow2_ceil(spc * block_n < 32 ? 32 : spc * block_n);
  p.pitch_a = ck * 2;
  p.layout_a = swizzle_layout_type(p.pitch_a);
  p.pitch_b = ckb * 2;
  p.layout_b = swizzle_layout_type(p.pitch_b);
  p.idesc = make_idesc_bf16(128, block_n, 1, 1);
  p.tw_log2 = ilog2(tl.tw);
  p.th_log2 = ilog2(tl.th);
  p.tiles_w = tl.tiles_w;
  p.tiles_h = tl.tiles_h;
  p.dw = dw;
  p.tap_stride = tap_stride;
  p.a_stride = a_stride;
  p.b_stride = b_stride;
  p.Cout = b_valid > 0 ? b_valid : cout;
  p.err = device_err_flag();
  const int b_tile = 128 * block_n * 2;
  int stages_a = (int)((220 * 1024 - 2 * b_tile) / (32 * 1024));
  if (stages_a > 4) stages_a = 4;
  PDAC_REQUIRE(stages_a >= 2, "internal: wgrad shared memory budget");
  p.stages_a = stages_a;
  for (int i = 0; i < n_views; ++i) PDAC_TRY(encode_view_map(&maps.a[i], &views[i], ck, tl.tw, tl.th, tl.tn));
  PDAC_TRY(encode_view_map(&maps.w, dy, ckb, tl.tw, tl.th, tl.tn));
  const int y = ceil_div(p.n_slabs, spc), z = cout / block_n;
  int splits = (2 * num_sms()) / (y * z);
  if (splits < 1) splits = 1;
  if (splits > p.n_kblocks) splits = p.n_kblocks;
  const int a_total = p.chunks_total * ck;
  PDAC_TRY(wgrad_plan_partials(scratch, scratch_bytes, p.n_taps, a_total, p.Cout, &splits, &p.part, &p.part_stride));
  p.part_A = a_total;
  size_t smem = (size_t)stages_a * 32 * 1024 + 2 * (size_t)b_tile + 1024;
  static bool attr_set = false;
  if (!attr_set) {
    PDAC_CUDA(cudaFuncSetAttribute(wgrad_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(226 * 1024)));
    attr_set = true;
  }
  dim3 grid(splits, y, z);
  wgrad_tc_kernel<<<grid, kWgradThreads, smem, (cudaStream_t)stream>>>(maps, p);
  PDAC_LAUNCH_CHECK("wgrad_tc_kernel");
  if (p.part)
    PDAC_TRY(wgrad_reduce(p.part, splits, p.part_stride, p.n_taps, p.Cout, a_total, dw, tap_stride, a_stride, b_stride,
                          stream));
  note_kernel(KID_WGRAD_TC);
  return PDAC_OK;
}

// ------------------------------------------------------------------------------------------------ nn.Conv2d wrappers

namespace pdac {

static PdacView parity_view(const PdacView& v, int ph, int pw) {
  PdacView r = v;
  r.ptr = (bf16*)v.ptr + ph * v.sH + pw * v.sW;
  r.H = (v.H - ph + 1) / 2;
  r.W = (v.W - pw + 1) / 2;
  r.sH = v.sH * 2;
  r.sW = v.sW * 2;
  return r;
}
static inline int floor_div2(int o) { return (o - (o & 1)) / 2; }

// Builds the forward tap list of a conv (stride 1 or 2) over `n_src` concatenated sources; for stride 2 the views are
// the four parity sub-lattices of the (single) source.
static int build_fwd_taps(const PdacConv* cv, int n_src, const PdacView* src, PdacTapConv* tc, PdacView* views,
                          int* n_views) {
  PDAC_REQUIRE(cv->R * cv->S <= PDAC_MAX_TAPS, "kernel %dx%d has too many taps", cv->R, cv->S);
  PDAC_REQUIRE(cv->stride == 1 || cv->stride == 2, "stride must be 1 or 2");
  tc->n_taps = cv->R * cv->S;
  tc->wrows = cv->wrows;
  tc->relu = cv->relu;
  if (cv->stride == 1) {
    tc->n_cat = n_src;
    for (int s = 0; s < n_src; ++s) views[s] = src[s];
    *n_views = n_src;
    for (int r = 0; r < cv->R; ++r)
      for (int s = 0; s < cv->S; ++s) {
        PdacTap& t = tc->taps[r * cv->S + s];
        t.src = 0;
        t.dh = r * cv->dil - cv->pad;
        t.dw = s * cv->dil - cv->pad;
      }
  } else {
    PDAC_REQUIRE(n_src == 1, "stride-2 convolutions take a single source");
    tc->n_cat = 1;
    for (int ph = 0; ph < 2; ++ph)
      for (int pw = 0; pw < 2; ++pw) views[ph * 2 + pw] = parity_view(src[0], ph, pw);
    *n_views = 4;
    for (int r = 0; r < cv->R; ++r)
      for (int s = 0; s < cv->S; ++s) {
        const int oh = r * cv->dil - cv->pad, ow = s * cv->dil - cv->pad;
        PdacTap& t = tc->taps[r * cv->S + s];
        t.src = (oh & 1) * 2 + (ow & 1);
        t.dh = floor_div2(oh);
        t.dw = floor_div2(ow);
      }
  }
  return PDAC_OK;
}

// out view = add view (or zeros), V channels per thread (V = 8: one 16-byte access; V = 1: any alignment)
template <int V>
__global__ void fill_view_kernel(bf16* out, long long sN, long long sH, long long sW, int N, int H, int W, int C,
                                 const bf16* add, long long aN, long long aH, long long aW) {
  const int cv = C / V;
  const long long total = (long long)N * H * W * cv;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % cv) * V;
    long long r = i / cv;
    const int w = (int)(r % W);
    r /= W;
    const int h = (int)(r % H);
    const int n = (int)(r / H);
    bf16* o = out + n * sN + h * sH + w * sW + c;
    if (V == 8) {
      *reinterpret_cast<uint4*>(o) =
          add ? __ldg(reinterpret_cast<const uint4*>(add + n * aN + h * aH + w * aW + c)) : make_uint4(0, 0, 0, 0);
    } else {
      *o = add ? add[n * aN + h * aH + w * aW + c] : __float2bfloat16(0.f);
    }
  }
}

}  // namespace pdac

extern "C" int pdac_conv2d_fwd(const PdacConv* cv, int n_src, const PdacView* src, const void* w_packed,
                               const float* bias, const PdacView* addend, const PdacView* out, float* stats,
                               void* stream) {
  PDAC_REQUIRE(cv && src && out, "null argument");
  PDAC_REQUIRE(n_src == 1 || n_src == 2, "n_src must be 1 or 2");
  const int eh = (src[0].H + 2 * cv->pad - cv->dil * (cv->R - 1) - 1) / cv->stride + 1;
  const int ew = (src[0].W + 2 * cv->pad - cv->dil * (cv->S - 1) - 1) / cv->stride + 1;
  PDAC_REQUIRE(out->H == eh && out->W == ew && out->N == src[0].N, "output shape %dx%dx%d does not match conv (%dx%dx%d)",
               out->N, out->H, out->W, src[0].N, eh, ew);
  PdacTapConv tc;
  PdacView views[4];
  int nv = 0;
  PDAC_TRY(build_fwd_taps(cv, n_src, src, &tc, views, &nv));
  return pdac_tapconv_fwd(&tc, nv, views, w_packed, bias, addend, out, stats, stream);
}

extern "C" int pdac_conv2d_fwd_affine(const PdacConv* cv, int n_src, const PdacView* src, const void* w_packed,
                                      const float* scale, const float* shift, const PdacView* addend,
                                      const PdacView* out, void* stream) {
  PDAC_REQUIRE(scale && shift, "conv2d_fwd_affine: null scale / shift");
  set_epi_scale(scale);
  const int rc = pdac_conv2d_fwd(cv, n_src, src, w_packed, shift, addend, out, nullptr, stream);
  set_epi_scale(nullptr);
  return rc;
}

// conv + folded BatchNorm (+ ReLU) stored nearest-neighbour x2 upsampled: inference form of smp's decoder block output
// in front of the next block's F.interpolate(scale_factor=2, mode="nearest"). out2x is (N, 2Ho, 2Wo, Cout). *handled = 0:
// not eligible for the halo kernel's direct epilogue, nothing launched (run the conv and pdac_upsample_nearest2x_fwd).
extern "C" int pdac_conv2d_fwd_affine_up2(const PdacConv* cv, int n_src, const PdacView* src, const void* w_packed,
                                          const float* scale, const float* shift, const PdacView* out2x, int* handled,
                                          void* stream) {
  PDAC_REQUIRE(cv && src && w_packed && scale && shift && out2x && handled, "conv2d_fwd_affine_up2: null argument");
  *handled = 0;
  if (cv->stride != 1 || n_src < 1 || n_src > 2 || (out2x->H & 1) || (out2x->W & 1)) return PDAC_OK;
  PDAC_TRY(check_dev_ptr(out2x->ptr, "out2x"));
  PdacTapConv tc;
  PdacView views[4];
  int nv = 0;
  memset(&tc, 0, sizeof(tc));
  PDAC_TRY(build_fwd_taps(cv, n_src, src, &tc, views, &nv));
  PdacView out = *out2x;  // the even pixels
  out.H = out2x->H / 2;
  out.W = out2x->W / 2;
  out.sH = 2 * out2x->sH;
  out.sW = 2 * out2x->sW;
  if (out.H != src[0].H || out.W != src[0].W || out.N != src[0].N) return PDAC_OK;  // spatial-size-preserving convs only
  set_epi_scale(scale);
  set_epi_up2(out2x);
  const int rc = conv_halo_try(&tc, nv, views, w_packed, shift, nullptr, &out, nullptr, stream, handled);
  set_epi_up2(nullptr);
  set_epi_scale(nullptr);
  return rc;
}

extern "C" int pdac_tapconv_fwd_affine(const PdacTapConv* tc, int n_views, const PdacView* views, const void* w_packed,
                                       const float* scale, const float* shift, const PdacView* addend,
                                       const PdacView* out, void* stream) {
  PDAC_REQUIRE(scale && shift, "tapconv_fwd_affine: null scale / shift");
  set_epi_scale(scale);
  const int rc = pdac_tapconv_fwd(tc, n_views, views, w_packed, shift, addend, out, nullptr, stream);
  set_epi_scale(nullptr);
  return rc;
}

extern "C" int pdac_conv2d_wgrad(const PdacConv* cv, int n_src, const PdacView* src, const PdacView* dy, float* dw,
                                 int dw_layout, float* scratch, long long scratch_bytes, void* stream) {
  PDAC_REQUIRE(cv && src && dy && dw, "null argument");
  PdacTapConv tc;
  PdacView views[4];
  int nv = 0;
  PDAC_TRY(build_fwd_taps(cv, n_src, src, &tc, views, &nv));
  const int ctot = src[0].C + (n_src > 1 ? src[1].C : 0);
  const int cv_valid = cv->cout_valid > 0 ? cv->cout_valid : dy->C;
  if (dw_layout == 1)  // nn.Conv2d state_dict layout [Cout][Ctot][R][S]
    return pdac_tapconv_wgrad(&tc, nv, views, dy, dw, 1, (int64_t)cv->R * cv->S, (int64_t)ctot * cv->R * cv->S,
                              cv_valid, scratch, scratch_bytes, stream);
  return pdac_tapconv_wgrad(&tc, nv, views, dy, dw, (int64_t)dy->C * ctot, 1, ctot, cv_valid, scratch, scratch_bytes,
                            stream);
}

extern "C" int pdac_conv2d_dgrad(const PdacConv* cv, const PdacView* dy, const void* w_packed_t,
                                 const PdacView* addend, const PdacView* dx, void* stream) {
  PDAC_REQUIRE(cv && dy && w_packed_t && dx, "null argument");
  PDAC_REQUIRE(cv->stride == 1 || cv->stride == 2, "stride must be 1 or 2");
  PdacTapConv tc;
  memset(&tc, 0, sizeof(tc));
  tc.n_cat = 1;
  tc.wrows = cv->wrows;
  tc.relu = 0;
  if (cv->stride == 1) {
    tc.n_taps = cv->R * cv->S;
    for (int r = 0; r < cv->R; ++r)
      for (int s = 0; s < cv->S; ++s) {
        PdacTap& t = tc.taps[r * cv->S + s];
        t.src = 0;
        t.dh = cv->pad - r * cv->dil;
        t.dw = cv->pad - s * cv->dil;
      }
    return pdac_tapconv_fwd(&tc, 1, dy, w_packed_t, nullptr, addend, dx, nullptr, stream);
  }
  // stride 2: dx[2i+ph, 2j+pw] = sum over taps with (r*dil - pad - ph) even of dy[i + (ph + pad - r*dil)/2, ...] W[r,s].
  // The weight matrix keeps all taps; each parity class uses a sub-list, so the tap rows are addressed through
  // per-class packed copies: rows of unused taps are skipped by giving every class its own tap->row mapping. The
  // kernel addresses weight rows as t*wrows with t the position in the list, so the packed matrix for stride-2 dgrad
  // is laid out class by class: [class (ph,pw)][taps of that class in (r,s) order][wrows][Cout].
  int row_base = 0;
  for (int ph = 0; ph < 2; ++ph)
    for (int pw = 0; pw < 2; ++pw) {
      tc.n_taps = 0;
      for (int r = 0; r < cv->R; ++r) {
        if ((r * cv->dil - cv->pad - ph) & 1) continue;
        for (int s = 0; s < cv->S; ++s) {
          if ((s * cv->dil - cv->pad - pw) & 1) continue;
          PdacTap& t = tc.taps[tc.n_taps++];
          t.src = 0;
          t.dh = (ph + cv->pad - r * cv->dil) / 2;
          t.dw = (pw + cv->pad - s * cv->dil) / 2;
        }
      }
      PdacView dxp = parity_view(*dx, ph, pw);
      PdacView adp;
      if (addend) adp = parity_view(*addend, ph, pw);
      if (dxp.H <= 0 || dxp.W <= 0) continue;
      if (tc.n_taps == 0) {
        // no tap lands on this parity class (1x1 stride 2): dx = addend there, i.e. nothing to do when the gradient
        // is accumulated in place, else a copy / zero fill
        if (addend && adp.ptr == dxp.ptr && adp.sN == dxp.sN && adp.sH == dxp.sH && adp.sW == dxp.sW) continue;
        const bool v8 = dxp.C % 8 == 0 && ((uintptr_t)dxp.ptr & 15) == 0 && dxp.sW % 8 == 0 && dxp.sH % 8 == 0 &&
                        dxp.sN % 8 == 0 &&
                        (!addend || (((uintptr_t)adp.ptr & 15) == 0 && adp.sW % 8 == 0 && adp.sH % 8 == 0 && adp.sN % 8 == 0));
        const long long total = (long long)dxp.N * dxp.H * dxp.W * (v8 ? dxp.C / 8 : dxp.C);
        int blocks = (int)((total + 255) / 256);
        if (blocks > 148 * 8) blocks = 148 * 8;
        if (v8)
          fill_view_kernel<8><<<blocks, 256, 0, (cudaStream_t)stream>>>(
              (bf16*)dxp.ptr, dxp.sN, dxp.sH, dxp.sW, dxp.N, dxp.H, dxp.W, dxp.C,
              addend ? (const bf16*)adp.ptr : nullptr, addend ? adp.sN : 0, addend ? adp.sH : 0, addend ? adp.sW : 0);
        else
          fill_view_kernel<1><<<blocks, 256, 0, (cudaStream_t)stream>>>(
              (bf16*)dxp.ptr, dxp.sN, dxp.sH, dxp.sW, dxp.N, dxp.H, dxp.W, dxp.C,
              addend ? (const bf16*)adp.ptr : nullptr, addend ? adp.sN : 0, addend ? adp.sH : 0, addend ? adp.sW : 0);
        PDAC_LAUNCH_CHECK("fill_view_kernel");
        continue;
      }
      const bf16* wcls = (const bf16*)w_packed_t + (size_t)row_base * cv->wrows * dy->C;
      PDAC_TRY(pdac_tapconv_fwd(&tc, 1, dy, wcls, nullptr, addend ? &adp : nullptr, &dxp, nullptr, stream));
      row_base += tc.n_taps;
    }
  return PDAC_OK;
}

// Data gradient of a stride-1 "same" convolution whose first dx_pooled->C input channels were a nearest-neighbour x2
// upsampling (smp UnetDecoder: F.interpolate(x, scale_factor=2) -> cat([x, skip]) -> conv): those channels' gradient is
// wanted at half resolution, summed over each 2x2 block, and is formed in the halo kernel's epilogue from the staged
// bf16 tile instead of being written at full resolution and read back by pdac_upsample_nearest2x_bwd (same values bit
// for bit: the same bf16 roundings and the same summation order). *handled = 0: geometry not eligible, nothing was
// launched and the caller runs the two-kernel form.
extern "C" int pdac_conv2d_dgrad_pool2(const PdacConv* cv, const PdacView* dy, const void* w_packed_t,
                                       const PdacView* dx_pooled, const PdacView* dx_rest, int* handled, void* stream) {
  PDAC_REQUIRE(cv && dy && w_packed_t && dx_pooled && handled, "null argument");
  *handled = 0;
  if (cv->stride != 1) return PDAC_OK;
  PDAC_TRY(check_dev_ptr(dx_pooled->ptr, "dx_pooled"));
  if (dx_rest) PDAC_TRY(check_dev_ptr(dx_rest->ptr, "dx_rest"));
  PDAC_REQUIRE(dx_pooled->N == dy->N && 2 * dx_pooled->H == dy->H && 2 * dx_pooled->W == dy->W,
               "dgrad_pool2: dx_pooled must be (N, H/2, W/2, C) of dy's (N, H, W)");
  PDAC_REQUIRE(!dx_rest || (dx_rest->N == dy->N && dx_rest->H == dy->H && dx_rest->W == dy->W),
               "dgrad_pool2: dx_rest must have dy's N, H, W");
  PDAC_REQUIRE(2 * cv->pad == (cv->R - 1) * cv->dil && 2 * cv->pad == (cv->S - 1) * cv->dil,
               "dgrad_pool2: the convolution must keep the spatial size");
  if ((dx_pooled->C & 7) || (dx_rest && (dx_rest->C & 7))) return PDAC_OK;
  PdacTapConv tc;
  memset(&tc, 0, sizeof(tc));
  tc.n_cat = 1;
  tc.wrows = cv->wrows;
  tc.relu = 0;
  tc.n_taps = cv->R * cv->S;
  for (int r = 0; r < cv->R; ++r)
    for (int s = 0; s < cv->S; ++s) {
      PdacTap& t = tc.taps[r * cv->S + s];
      t.src = 0;
      t.dh = cv->pad - r * cv->dil;
      t.dw = cv->pad - s * cv->dil;
    }
  // the full-resolution output the kernel indexes: channels [0, pooled C) are never stored through it
  PdacView out;
  out.N = dy->N;
  out.H = dy->H;
  out.W = dy->W;
  out.C = dx_pooled->C + (dx_rest ? dx_rest->C : 0);
  if (dx_rest) {
    out.ptr = (bf16*)dx_rest->ptr - dx_pooled->C;
    out.sN = dx_rest->sN;
    out.sH = dx_rest->sH;
    out.sW = dx_rest->sW;
  } else {
    out.ptr = dx_pooled->ptr;
    out.sN = out.sH = out.sW = 0;
  }
  set_epi_pool(dx_pooled);
  const int rc = conv_halo_try(&tc, 1, dy, w_packed_t, nullptr, nullptr, &out, nullptr, stream, handled);
  set_epi_pool(nullptr);
  return rc;
}

extern "C" int pdac_convt2x2_fwd(const PdacView* x, const void* w_packed, int wrows, const float* bias, int relu,
                                 const PdacView* out, void* stream) {
  PDAC_REQUIRE(x && w_packed && out, "null argument");
  PDAC_REQUIRE(out->H == 2 * x->H && out->W == 2 * x->W && out->N == x->N, "convT output must be 2x the input");
  PdacTapConv tc;
  memset(&tc, 0, sizeof(tc));
  tc.n_taps = 1;
  tc.n_cat = 1;
  tc.wrows = wrows;
  tc.relu = relu;
  for (int a = 0; a < 2; ++a)
    for (int b = 0; b < 2; ++b) {
      PdacView op = parity_view(*out, a, b);
      const bf16* wt = (const bf16*)w_packed + (size_t)(a * 2 + b) * wrows * x->C;
      PDAC_TRY(pdac_tapconv_fwd(&tc, 1, x, wt, bias, nullptr, &op, nullptr, stream));
    }
  return PDAC_OK;
}

extern "C" int pdac_convt2x2_dgrad(const PdacView* dy, const void* w_packed_t, int wrows, const PdacView* dx,
                                   void* stream) {
  PDAC_REQUIRE(dy && w_packed_t && dx, "null argument");
  PDAC_REQUIRE(dy->H == 2 * dx->H && dy->W == 2 * dx->W, "convT dgrad: dy must be 2x dx");
  PdacTapConv tc;
  memset(&tc, 0, sizeof(tc));
  tc.n_taps = 4;
  tc.n_cat = 1;
  tc.wrows = wrows;
  PdacView views[4];
  for (int a = 0; a < 2; ++a)
    for (int b = 0; b < 2; ++b) {
      views[a * 2 + b] = parity_view(*dy, a, b);
      tc.taps[a * 2 + b].src = a * 2 + b;
    }
  return pdac_tapconv_fwd(&tc, 4, views, w_packed_t, nullptr, nullptr, dx, nullptr, stream);
}

extern "C" int pdac_convt2x2_wgrad(const PdacView* x, const PdacView* dy, float* dw, int dw_layout, float* scratch,
                                   long long scratch_bytes, void* stream) {
  PDAC_REQUIRE(x && dy && dw, "null argument");
  // dW[a,b][cout][cin] = sum_{i,j} dy[2i+a, 2j+b, cout] * x[i, j, cin]: tapped operand = dy parity views, fixed = x.
  PdacTapConv tc;
  memset(&tc, 0, sizeof(tc));
  tc.n_taps = 4;
  tc.n_cat = 1;
  PdacView views[4];
  for (int a = 0; a < 2; ++a)
    for (int b = 0; b < 2; ++b) {
      views[a * 2 + b] = parity_view(*dy, a, b);
      tc.taps[a * 2 + b].src = a * 2 + b;
    }
  if (dw_layout == 1)  // nn.ConvTranspose2d state_dict layout [Cin][Cout][2][2]
    return pdac_tapconv_wgrad(&tc, 4, views, x, dw, 1, 4, (int64_t)dy->C * 4, 0, scratch, scratch_bytes, stream);
  return pdac_tapconv_wgrad(&tc, 4, views, x, dw, (int64_t)dy->C * x->C, x->C, 1, 0, scratch, scratch_bytes, stream);
}
