// Halo-tile tensor-core tap convolution for sm_100a: the stride-1 fast path of pdac_tapconv_fwd (forward and data
// gradient of every 3x3 / dilated / 4x4-stem convolution whose taps all read the same source lattice).
//
// Why: the generic kernel (conv_tc.cu) fetches one 128-pixel A box per (tap, channel chunk), i.e. the input is pulled
// through L2->SM nine times for a 3x3 conv, and the weights once per 128-pixel tile. ncu showed that kernel bound by
// L2->SM traffic (7.4 TB/s, tensor pipe 18-35 % busy), and by the TMA request rate on 32-byte pixels (C = 16).
// Here a CTA loads ONE (TH+eh) x (TW+ew) halo box per channel chunk and forms every tap operand as a shared-memory
// descriptor shifted by (dh*PW + dw) pixel rows (scripts/probe/shift_desc.cu: swizzled K-major descriptors accept any
// whole-row shift because the swizzle is a function of the absolute address). The MMA rows enumerate the halo-pitch
// raster o = th*PW + tw, so the columns tw >= TW of every row are computed and thrown away (1 - TW/PW of the MMA work).
// A tile is nacc accumulators of 128 rows; TMEM holds two tiles (2 * nacc * N <= 512 columns) so the epilogue of tile i
// overlaps the main loop of tile i+1. Weights stay resident in shared memory for the whole (persistent) CTA when they
// fit, otherwise they stream through their own ring, each stage reused by the nacc accumulators of the tile.
//
// Roles: warp 0 = A (halo) TMA producer, warp 1 = TMEM allocator + MMA issuer, warps 2..9 = two epilogue warpgroups
// (accumulators of a tile alternate between them), warp 10 = weight TMA producer.
#include "common.h"
#include "ptx.cuh"
#include <stdlib.h>
#include <string.h>

namespace pdac {

struct HaloMaps {
  CUtensorMap a[2];  // the (up to two concatenated) source views, box (ck, PW, R, 1)
  CUtensorMap w;     // packed weights [n_taps*wrows][Ktot], box (ck, block_n)
};

struct HaloParams {
  int n_taps, n_chunks, chunks0, ck;
  int tap_shift[PDAC_MAX_TAPS];  // (dh - dh_min) * PW + (dw - dw_min), in pixel rows of the halo buffer
  int dh_min, dw_min;
  uint32_t pitch, layout_type, idesc, tmem_cols;
  int block_n, tiles_c, wrows;
  int nacc, TW, TH, PW, R;
  int tiles_w, tiles_h, tiles_sp;  // spatial tiles per image row / column, total over the batch
  int a_buf, sa;                   // bytes per halo stage, stages
  int b_resident, sb, b_sub;       // weights resident? else ring stages; bytes per (tap, chunk) weight tile
  int N, H, W, Cout;
  bf16* out;
  long long osN, osH, osW;
  const bf16* addend;
  long long asN, asH, asW;
  const float* bias;
  float* stats;
  int relu;
  int* err;
};

static constexpr int kHaloThreads = 352;
static constexpr int kHaloMaxSA = 6;
static constexpr int kHaloMaxSB = 8;

__device__ __forceinline__ float h_bf16lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float h_bf16hi(uint32_t u) { return __uint_as_float(u & 0xffff0000u); }
__device__ __forceinline__ uint32_t h_pack(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}

__global__ void __launch_bounds__(kHaloThreads, 1) conv_halo_kernel(const __grid_constant__ HaloMaps maps,
                                                                    const __grid_constant__ HaloParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_a[kHaloMaxSA];
  __shared__ __align__(8) uint64_t empty_a[kHaloMaxSA];
  __shared__ __align__(8) uint64_t full_b[kHaloMaxSB];
  __shared__ __align__(8) uint64_t empty_b[kHaloMaxSB];
  __shared__ __align__(8) uint64_t tmem_full_bar[2];
  __shared__ __align__(8) uint64_t tmem_empty_bar[2];
  __shared__ uint32_t tmem_base_slot;
  __shared__ long long s_rowoff[2][128];

  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + (size_t)p.sa * p.a_buf;
  const int b_bytes = (p.b_resident ? p.n_taps * p.n_chunks : p.sb) * p.b_sub;
  uint8_t* smem_stg = smem_b + b_bytes;

  // this CTA's share: a fixed N tile and every ctas_sp-th spatial tile
  const int n_tile = blockIdx.x % p.tiles_c;
  const int sp0 = blockIdx.x / p.tiles_c;
  const int sp_step = gridDim.x / p.tiles_c;
  const int n0 = n_tile * p.block_n;

  if (threadIdx.x == 0) {
    for (int i = 0; i < p.sa; ++i) {
      mbar_init(&full_a[i], 1);
      mbar_init(&empty_a[i], 1);
    }
    for (int i = 0; i < kHaloMaxSB; ++i) {
      mbar_init(&full_b[i], 1);
      mbar_init(&empty_b[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full_bar[i], 1);
      mbar_init(&tmem_empty_bar[i], 256);
    }
    fence_barrier_init();
  }
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&maps.a[0]);
    tma_prefetch_desc(&maps.w);
  }
  if (warp == 1) {
    tmem_alloc(&tmem_base_slot, p.tmem_cols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_slot;

#define HALO_DECODE(sp)                         \
  int r_ = (sp);                                \
  const int tile_w = r_ % p.tiles_w;            \
  r_ /= p.tiles_w;                              \
  const int tile_h = r_ % p.tiles_h;            \
  const int img = r_ / p.tiles_h;               \
  const int w0 = tile_w * p.TW, h0 = tile_h * p.TH;

  if (warp == 0) {
    // ------------------------------------------------------------------ halo producer
    if (lane == 0 && sp0 < sp_step) {
      const uint32_t halo_bytes = (uint32_t)(p.R * p.PW) * p.pitch;
      int ita = 0;
      for (int sp = sp0; sp < p.tiles_sp; sp += sp_step) {
        HALO_DECODE(sp)
        for (int c = 0; c < p.n_chunks; ++c, ++ita) {
          const int stage = ita % p.sa;
          mbar_wait(&empty_a[stage], ((ita / p.sa) & 1) ^ 1, p.err);
          mbar_expect_tx(&full_a[stage], halo_bytes);
          const int s = c >= p.chunks0;
          tma_load_4d(smem_a + (size_t)stage * p.a_buf, &maps.a[s], &full_a[stage], (s ? c - p.chunks0 : c) * p.ck,
                      w0 + p.dw_min, h0 + p.dh_min, img);
        }
      }
    }
  } else if (warp == 10) {
    // ------------------------------------------------------------------ weight producer
    if (lane == 0 && sp0 < sp_step) {
      if (p.b_resident) {
        mbar_expect_tx(&full_b[0], (uint32_t)b_bytes);
        for (int t = 0; t < p.n_taps; ++t)
          for (int c = 0; c < p.n_chunks; ++c)
            tma_load_2d(smem_b + (size_t)(t * p.n_chunks + c) * p.b_sub, &maps.w, &full_b[0], c * p.ck,
                        t * p.wrows + n0);
      } else {
        int itb = 0;
        for (int sp = sp0; sp < p.tiles_sp; sp += sp_step) {
          for (int c = 0; c < p.n_chunks; ++c) {
            for (int t = 0; t < p.n_taps; ++t, ++itb) {
              const int stage = itb % p.sb;
              mbar_wait(&empty_b[stage], ((itb / p.sb) & 1) ^ 1, p.err);
              mbar_expect_tx(&full_b[stage], (uint32_t)p.b_sub);
              tma_load_2d(smem_b + (size_t)stage * p.b_sub, &maps.w, &full_b[stage], c * p.ck, t * p.wrows + n0);
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0 && sp0 < sp_step) {
      const uint32_t sbo = 8 * p.pitch;
      const int ksteps = p.ck >> 4;
      if (p.b_resident) {
        mbar_wait(&full_b[0], 0, p.err);
        tc_fence_after();
      }
      const uint32_t b_base0 = smem_u32(smem_b);
      int ita = 0, itb = 0, lt = 0;
      for (int sp = sp0; sp < p.tiles_sp; sp += sp_step, ++lt) {
        const int g = lt & 1;
        mbar_wait(&tmem_empty_bar[g], ((lt >> 1) & 1) ^ 1, p.err);
        tc_fence_after();
        const uint32_t tacc = tmem_base + (uint32_t)(g * p.nacc * p.block_n);
        for (int c = 0; c < p.n_chunks; ++c, ++ita) {
          const int stage = ita % p.sa;
          mbar_wait(&full_a[stage], (ita / p.sa) & 1, p.err);
          tc_fence_after();
          const uint32_t a_base = smem_u32(smem_a + (size_t)stage * p.a_buf);
          for (int t = 0; t < p.n_taps; ++t) {
            uint32_t b_base;
            int bstage = 0;
            if (p.b_resident) {
              b_base = b_base0 + (uint32_t)((t * p.n_chunks + c) * p.b_sub);
            } else {
              bstage = itb % p.sb;
              mbar_wait(&full_b[bstage], (itb / p.sb) & 1, p.err);
              tc_fence_after();
              b_base = b_base0 + (uint32_t)(bstage * p.b_sub);
            }
            const uint64_t db0 = make_smem_desc(b_base, 16, sbo, p.layout_type);
            const uint32_t a_tap = a_base + (uint32_t)p.tap_shift[t] * p.pitch;
            for (int j = 0; j < p.nacc; ++j) {
              const uint64_t da0 = make_smem_desc(a_tap + (uint32_t)(j * 128) * p.pitch, 16, sbo, p.layout_type);
              for (int kk = 0; kk < ksteps; ++kk) {
                // +2 in the 16-byte-unit address field = +32 bytes = 16 bf16 along K
                umma_bf16(tacc + (uint32_t)(j * p.block_n), da0 + (uint64_t)(2 * kk), db0 + (uint64_t)(2 * kk), p.idesc,
                          (c | t | kk) ? 1u : 0u);
              }
            }
            if (!p.b_resident) {
              umma_commit(&empty_b[bstage]);
              ++itb;
            }
          }
          umma_commit(&empty_a[stage]);
        }
        umma_commit(&tmem_full_bar[g]);
      }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------------ epilogue (warps 2..9)
    const int wg = (warp - 2) >> 2;
    const int q = warp & 3;  // TMEM lane quarter this warp may read
    const int m = q * 32 + lane;
    const int e = (threadIdx.x - 64) & 127;
    const int spitch = p.block_n * 2 + 16;
    uint8_t* stg = smem_stg + (size_t)wg * 128 * spitch;
    const uint32_t bar_id = 1 + wg;
    const int nch = min(p.block_n, p.Cout - n0);
    const int pairs = p.block_n >> 1;
    const int groups = 128 / pairs > 0 ? 128 / pairs : 1;
    const int pr = e % pairs, sg = e / pairs;
    const bool stat_thread = p.stats && sg < groups && pairs <= 128 && (nch & 1) == 0 && 2 * pr < nch;
    float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;

    int lt = 0;
    if (sp0 < sp_step)
      for (int sp = sp0; sp < p.tiles_sp; sp += sp_step, ++lt) {
        HALO_DECODE(sp)
        const int g = lt & 1;
        mbar_wait(&tmem_full_bar[g], (lt >> 1) & 1, p.err);
        tc_fence_after();
        for (int j = wg; j < p.nacc; j += 2) {
          const int o = j * 128 + m;
          const int th = o / p.PW, tw = o - th * p.PW;
          const int w = w0 + tw, h = h0 + th;
          const bool valid = (tw < p.TW) && (th < p.TH) && (w < p.W) && (h < p.H);
          const long long ooff = (long long)img * p.osN + (long long)h * p.osH + (long long)w * p.osW;
          const long long aoff = (long long)img * p.asN + (long long)h * p.asH + (long long)w * p.asW;
          asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");  // previous staging fully consumed
          s_rowoff[wg][m] = valid ? ooff : -1;
          const uint32_t tacc =
              tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)((g * p.nacc + j) * p.block_n);
          for (int c = 0; c < p.block_n; c += 16) {
            uint32_t v[16];
            tmem_ld16(tacc + (uint32_t)c, v);
            tmem_ld_wait();
            float f[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) f[i] = __uint_as_float(v[i]);
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
                      f[hh * 8 + 2 * i] += h_bf16lo(uu[i]);
                      f[hh * 8 + 2 * i + 1] += h_bf16hi(uu[i]);
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
            o0.x = h_pack(f[0], f[1]);
            o0.y = h_pack(f[2], f[3]);
            o0.z = h_pack(f[4], f[5]);
            o0.w = h_pack(f[6], f[7]);
            o1.x = h_pack(f[8], f[9]);
            o1.y = h_pack(f[10], f[11]);
            o1.z = h_pack(f[12], f[13]);
            o1.w = h_pack(f[14], f[15]);
            uint4* sp4 = reinterpret_cast<uint4*>(stg + (size_t)m * spitch + c * 2);
            sp4[0] = o0;
            sp4[1] = o1;
          }
          if (j + 2 >= p.nacc) {
            // last accumulator of this warpgroup for the tile: its TMEM lanes are drained
            tc_fence_before();
            mbar_arrive(&tmem_empty_bar[g]);
          }
          asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");

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
          if (stat_thread) {
            for (int row = sg; row < 128; row += groups) {
              if (s_rowoff[wg][row] < 0) continue;
              const uint32_t u = *reinterpret_cast<const uint32_t*>(stg + (size_t)row * spitch + pr * 4);
              const float x0 = h_bf16lo(u), x1 = h_bf16hi(u);
              s0 += x0;
              q0 += x0 * x0;
              s1 += x1;
              q1 += x1 * x1;
            }
          }
        }
        if (wg >= p.nacc) {  // a warpgroup with no accumulator in this tile still owes its arrivals
          tc_fence_before();
          mbar_arrive(&tmem_empty_bar[g]);
        }
      }
    if (stat_thread) {
      float* st = p.stats + n0 + 2 * pr;
      atomicAdd(st, s0);
      atomicAdd(st + 1, s1);
      atomicAdd(st + p.Cout, q0);
      atomicAdd(st + p.Cout + 1, q1);
    }
  }
#undef HALO_DECODE

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, p.tmem_cols);
  }
}

// ------------------------------------------------------------------------------------------------ host side

static const int kSmemBudget = 224 * 1024;  // dynamic shared memory we allow ourselves (227 KB max per CTA incl. static)

struct HaloPlan {
  int TW, TH, nacc, sa, sb, b_resident;
  double cost;
};

// Picks the tile. Cost model (cycles per image, per SM share): MMA issue = tiles * nacc * taps * K16-steps * max(N/2, 10)
// against bytes moved (halo in + tile out) at ~40 B/cycle/SM; the larger of the two plus a per-tile overhead.
static bool plan_halo(int H, int W, int N, int n_taps, int n_chunks, int ck, int eh, int ew, int has_addend,
                      HaloPlan* best) {
  const int pitch = ck * 2;
  const int max_nacc = 256 / N < 4 ? 256 / N : 4;
  if (max_nacc < 1) return false;
  const long long w_bytes = (long long)n_taps * n_chunks * N * pitch;
  const int staging = 2 * 128 * (N * 2 + 16);
  best->cost = 1e30;
  const int tws[5] = {16, 32, 64, 128, W};
  for (int ti = 0; ti < 5; ++ti) {
    int TW = tws[ti];
    if (TW > W) TW = W;
    if (TW + ew > 256) continue;
    const int PW = TW + ew;
    for (int nacc = 1; nacc <= max_nacc; ++nacc) {
      int TH = (128 * nacc - TW) / PW + 1;
      if (128 * nacc < TW) continue;
      if (TH > H) TH = H;
      if (TH + eh > 256) TH = 256 - eh;
      const int nacc_eff = ((TH - 1) * PW + TW + 127) / 128;
      if (nacc_eff != nacc) continue;  // the same tile is reachable with fewer accumulators
      const int R = TH + eh;
      const int a_buf = ((nacc * 128 + eh * PW + ew) * pitch + 1023) & ~1023;
      for (int res = 1; res >= 0; --res) {
        int sa, sb = 0;
        long long fixed = staging + 2048;
        if (res) {
          if (w_bytes > 112 * 1024) continue;
          fixed += w_bytes;
          sa = (int)((kSmemBudget - fixed) / a_buf);
        } else {
          sb = 4;
          fixed += (long long)sb * N * pitch;
          sa = (int)((kSmemBudget - fixed) / a_buf);
          if (sa > 3) {  // spend the rest on a deeper weight ring
            sa = 3;
            sb = (int)((kSmemBudget - staging - 2048 - 3LL * a_buf) / (N * pitch));
            if (sb > kHaloMaxSB) sb = kHaloMaxSB;
          }
        }
        if (sa > kHaloMaxSA) sa = kHaloMaxSA;
        if (sa < 2) continue;
        const int tiles = ceil_div(W, TW) * ceil_div(H, TH);
        const double mma_cyc = (double)tiles * nacc * n_taps * n_chunks * (ck / 16) * (N / 2 > 10 ? N / 2 : 10);
        double bytes = (double)tiles * R * PW * n_chunks * pitch + (double)H * W * N * 2 * (has_addend ? 2 : 1);
        if (!res) bytes += (double)tiles * w_bytes;
        const double mem_cyc = bytes / 40.0;
        const double cost = (mma_cyc > mem_cyc ? mma_cyc : mem_cyc) + 0.15 * (mma_cyc < mem_cyc ? mma_cyc : mem_cyc) +
                            300.0 * tiles;
        if (cost < best->cost) {
          best->cost = cost;
          best->TW = TW;
          best->TH = TH;
          best->nacc = nacc;
          best->sa = sa;
          best->sb = sb;
          best->b_resident = res;
        }
        if (res) break;  // resident weights always beat streaming them for the same tile
      }
    }
  }
  return best->cost < 1e30;
}

// Returns PDAC_OK with *handled = 1 when the halo kernel was launched, *handled = 0 when the geometry is not eligible
// (caller falls back to the generic tap kernel), or an error code.
int conv_halo_try(const PdacTapConv* tc, int n_views, const PdacView* views, const void* w_packed, const float* bias,
                  const PdacView* addend, const PdacView* out, float* stats, void* stream, int* handled) {
  *handled = 0;
  static int disabled = -1;
  if (disabled < 0) {
    const char* e = getenv("PDAC_NO_HALO");
    disabled = (e && e[0] == '1') ? 1 : 0;
  }
  if (disabled) return PDAC_OK;
  if (tc->n_taps < 2 || n_views != tc->n_cat) return PDAC_OK;
  int dh_min = 127, dh_max = -127, dw_min = 127, dw_max = -127;
  for (int t = 0; t < tc->n_taps; ++t) {
    if (tc->taps[t].src != 0) return PDAC_OK;
    dh_min = tc->taps[t].dh < dh_min ? tc->taps[t].dh : dh_min;
    dh_max = tc->taps[t].dh > dh_max ? tc->taps[t].dh : dh_max;
    dw_min = tc->taps[t].dw < dw_min ? tc->taps[t].dw : dw_min;
    dw_max = tc->taps[t].dw > dw_max ? tc->taps[t].dw : dw_max;
  }
  const int eh = dh_max - dh_min, ew = dw_max - dw_min;
  int ck = 64;
  for (int s = 0; s < tc->n_cat; ++s) {
    if (views[s].N != out->N || views[s].H != out->H || views[s].W != out->W) return PDAC_OK;
    while (ck > 8 && (views[s].C % ck)) ck >>= 1;
  }
  if (ck < 16) return PDAC_OK;
  const int cout = out->C;
  const int block_n = cout <= 128 ? ((cout + 15) / 16) * 16 : 128;
  if (tc->wrows < ceil_div(cout, block_n) * block_n) return PDAC_OK;
  const int chunks0 = views[0].C / ck;
  const int n_chunks = chunks0 + (tc->n_cat > 1 ? views[1].C / ck : 0);
  const int ktot = n_chunks * ck;
  HaloPlan pl;
  if (!plan_halo(out->H, out->W, block_n, tc->n_taps, n_chunks, ck, eh, ew, addend != nullptr, &pl)) return PDAC_OK;

  HaloMaps maps;
  HaloParams p;
  memset(&p, 0, sizeof(p));
  p.n_taps = tc->n_taps;
  p.n_chunks = n_chunks;
  p.chunks0 = chunks0;
  p.ck = ck;
  p.TW = pl.TW;
  p.TH = pl.TH;
  p.PW = pl.TW + ew;
  p.R = pl.TH + eh;
  p.nacc = pl.nacc;
  p.dh_min = dh_min;
  p.dw_min = dw_min;
  for (int t = 0; t < tc->n_taps; ++t) p.tap_shift[t] = (tc->taps[t].dh - dh_min) * p.PW + (tc->taps[t].dw - dw_min);
  p.pitch = ck * 2;
  p.layout_type = swizzle_layout_type(p.pitch);
  p.idesc = make_idesc_bf16(128, block_n, 0, 0);
  p.block_n = block_n;
  p.tiles_c = ceil_div(cout, block_n);
  p.wrows = tc->wrows;
  p.tmem_cols = pow2_ceil(2 * pl.nacc * block_n < 32 ? 32 : 2 * pl.nacc * block_n);
  p.tiles_w = ceil_div(out->W, pl.TW);
  p.tiles_h = ceil_div(out->H, pl.TH);
  p.tiles_sp = p.tiles_w * p.tiles_h * out->N;
  p.a_buf = ((pl.nacc * 128 + eh * p.PW + ew) * (int)p.pitch + 1023) & ~1023;
  p.sa = pl.sa;
  p.b_resident = pl.b_resident;
  p.sb = pl.sb;
  p.b_sub = block_n * p.pitch;
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
  p.stats = stats;
  p.relu = tc->relu;
  p.err = device_err_flag();
  for (int s = 0; s < tc->n_cat; ++s) PDAC_TRY(encode_view_map(&maps.a[s], &views[s], ck, p.PW, p.R, 1));
  if (tc->n_cat == 1) maps.a[1] = maps.a[0];
  {
    uint64_t dims[2] = {(uint64_t)ktot, (uint64_t)tc->n_taps * tc->wrows};
    uint64_t strides[1] = {(uint64_t)ktot * 2};
    uint32_t box[2] = {(uint32_t)ck, (uint32_t)block_n};
    PDAC_TRY(encode_tiled_bf16(&maps.w, w_packed, 2, dims, strides, box, p.pitch));
  }
  const size_t b_bytes = (size_t)(p.b_resident ? p.n_taps * p.n_chunks : p.sb) * p.b_sub;
  const size_t smem = (size_t)p.sa * p.a_buf + b_bytes + 2 * 128 * (size_t)(block_n * 2 + 16) + 1024;
  PDAC_REQUIRE(smem <= 224 * 1024, "internal: halo conv shared memory plan (%zu bytes)", smem);
  static bool attr_set = false;
  if (!attr_set) {
    PDAC_CUDA(cudaFuncSetAttribute(conv_halo_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(224 * 1024)));
    attr_set = true;
  }
  int per_n = num_sms() / p.tiles_c;
  if (per_n < 1) per_n = 1;
  if (per_n > p.tiles_sp) per_n = p.tiles_sp;
  conv_halo_kernel<<<per_n * p.tiles_c, kHaloThreads, smem, (cudaStream_t)stream>>>(maps, p);
  PDAC_LAUNCH_CHECK("conv_halo_kernel");
  *handled = 1;
  return PDAC_OK;
}

}  // namespace pdac
