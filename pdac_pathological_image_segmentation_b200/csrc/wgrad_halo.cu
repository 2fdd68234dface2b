// Halo-tile weight gradient for sm_100a: the stride-1 fast path of pdac_tapconv_wgrad.
//
//   dw[t][a][b] += sum_pixels X[pixel + off_t, a] * dY[pixel, b]
//
// GEMM view (as in conv_tc.cu): M = 128 rows of (tap, input channel) pairs = one "slab", N = output channels,
// K = pixels; both operands are pixel-major in shared memory, i.e. MN-major tcgen05 descriptors. The generic kernel
// loads one 128-pixel X box per (tap, channel chunk), so X crosses L2->SM nine times for a 3x3 conv (and, for 16- and
// 32-channel tensors, as 32/64-byte TMA requests). Here a K block is a 16 x 8 pixel tile: ONE (16+ew) x (8+eh) halo box
// per channel chunk is loaded and a tap is a descriptor start shifted by (dh*PW + dw) pixel rows; K step k (16 pixels)
// is tile row k, i.e. halo row stride PW. A slab's 128 M rows are 128/ck sub-blocks of ck channels, `lbo` bytes apart:
//   ck = 64: two sub-blocks = the two chunks of a stage (lbo = halo buffer size) or two taps of one chunk
//   ck < 64: the taps of one kernel row (dw = 0, 1, ..: lbo = one pixel), the rest of the slab is ignored
// (the slab table is built on the host). All slabs of a CTA accumulate in TMEM over its share of the K blocks
// (split-K over CTAs), then fp32 atomics into the gradient arena slot in the parameter's own layout.
//
// Roles: warp 0 = TMA producer, warp 1 = TMEM allocator + MMA issuer (warp-converged, one elected lane), warps 2..5 =
// epilogue.
#include "common.h"
#include "ptx.cuh"
#include <stdlib.h>
#include <string.h>

namespace pdac {

static constexpr int kWhMaxSlabs = 64;
static constexpr int kWhMaxStages = 6;
static constexpr int kWhThreads = 192;

struct WhSlab {
  uint32_t a_off16;  // start of sub-block 0 inside the stage, 16-byte units (chunk buffer + tap shift)
  uint32_t lbo16;    // distance between consecutive sub-blocks, 16-byte units
  int8_t cg;         // chunk group (stage index within a K block)
  int8_t sub_t[8];   // tap of each sub-block, -1 = ignored rows
  int8_t sub_c[8];   // global channel chunk of each sub-block
};

struct WhMaps {
  CUtensorMap a[2];  // source views, box (ck, PW, R, 1)
  CUtensorMap dy;    // box (ckb, 16, 8, 1)
};

struct WhParams {
  WhSlab slabs[kWhMaxSlabs];
  uint8_t cg_first[kWhMaxSlabs + 1];  // slabs of chunk group cg: [cg_first[cg], cg_first[cg + 1])
  int n_slabs, spc;        // total slabs, slabs per CTA (grid.y = ceil(n_slabs / spc))
  int n_cg, cps;           // chunk groups per K block, chunks per group
  int chunks0, n_chunks, ck;
  int PW, R, dh_min, dw_min;
  int a_buf, stages_a;     // bytes per chunk halo buffer, A stages (each cps buffers)
  int ckb, nboxb, block_n;
  uint32_t pitch_a, layout_a, pitch_b, layout_b, idesc, tmem_cols;
  int tiles_w, tiles_h, n_kblocks;
  // K block = 16 pixels x tile_rows rows (one MMA K step per row). Row-shift mode (16 / 32-channel tensors, see
  // wgrad_halo_try): the kernel-row offsets are taken on the dY side, so dY is the haloed operand in h: its box has
  // b_rows = tile_rows + eh rows and starts b_row0 rows above the tile; column group i of the accumulator belongs to
  // kernel row rs_tap_r[i] and A sub-block j to kernel column j: tap = rs_tap[i * 8 + j].
  int tile_rows, b_rows, b_row0, b_tx_bytes, b_tile_bytes;
  int rowshift, rs_cout;
  uint32_t lbo_b;
  int8_t rs_tap[64];
  float* dw;
  long long tap_stride, a_stride, b_stride;
  int Cout;
  // deterministic mode: split-K share blockIdx.x stores its partial [tap][Cout][A] at part + blockIdx.x * part_stride
  // (plain coalesced stores; wgrad_reduce adds the shares in index order) instead of fp32 atomics into dw
  float* part;
  long long part_stride;
  int part_A;
  int* err;
};

// DUAL: two CTAs per SM (each <= 104 KB of shared memory and <= 256 TMEM columns). The single issuing thread spends a
// few hundred cycles per halo stage on barrier waits, slab-table reads and commits, during which the tensor pipe idles
// (the MMA queue is only a few instructions deep, scripts/probe/mma_issue.cu); a second resident CTA fills those gaps.
template <int DUAL>
__global__ void __launch_bounds__(kWhThreads, DUAL ? 2 : 1) wgrad_halo_kernel(const __grid_constant__ WhMaps maps,
                                                                              const __grid_constant__ WhParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t fullA[kWhMaxStages];
  __shared__ __align__(8) uint64_t emptyA[kWhMaxStages];
  __shared__ __align__(8) uint64_t fullB[2];
  __shared__ __align__(8) uint64_t emptyB[2];
  __shared__ __align__(8) uint64_t tmem_full_bar;
  __shared__ uint32_t tmem_base_slot;

  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  const int a_stage = p.a_buf * p.cps;
  const int b_box = p.b_tile_bytes / p.nboxb;
  const int b_tile = p.b_tile_bytes;
  uint8_t* smem_b = smem + (size_t)p.stages_a * a_stage;

  const int slab0 = blockIdx.y * p.spc;
  const int slab1 = min(slab0 + p.spc, p.n_slabs);
  const int cg_lo = p.slabs[slab0].cg, cg_hi = p.slabs[slab1 - 1].cg;
  const int n0 = blockIdx.z * p.block_n;
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
    tma_prefetch_desc(&maps.dy);
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

  if (warp == 0) {
    if (lane == 0) {
      const uint32_t halo_bytes = (uint32_t)(p.R * p.PW) * p.pitch_a;
      int stage = 0;
      uint32_t phase = 0;  // A ring position and wrap parity, carried incrementally
      for (int i = 0; i < n_my_kb; ++i) {
        int kb = blockIdx.x + i * gridDim.x;
        const int tile_w = kb % p.tiles_w;
        kb /= p.tiles_w;
        const int tile_h = kb % p.tiles_h;
        const int img = kb / p.tiles_h;
        const int w0 = tile_w * 16, h0 = tile_h * p.tile_rows;
        const int bs = i & 1;
        mbar_wait(&emptyB[bs], ((i >> 1) & 1) ^ 1, p.err);
        mbar_expect_tx(&fullB[bs], (uint32_t)p.b_tx_bytes);
        for (int j = 0; j < p.nboxb; ++j)
          tma_load_4d(smem_b + (size_t)bs * b_tile + (size_t)j * b_box, &maps.dy, &fullB[bs], n0 + j * p.ckb, w0,
                      h0 + p.b_row0, img);
        for (int cg = cg_lo; cg <= cg_hi; ++cg) {
          mbar_wait(&emptyA[stage], phase ^ 1, p.err);
          const int c_first = cg * p.cps;
          const int nc = min(p.cps, p.n_chunks - c_first);
          mbar_expect_tx(&fullA[stage], halo_bytes * (uint32_t)nc);
          for (int j = 0; j < nc; ++j) {
            const int c = c_first + j;
            const int s = c >= p.chunks0;
            tma_load_4d(smem + (size_t)stage * a_stage + (size_t)j * p.a_buf, &maps.a[s], &fullA[stage],
                        (s ? c - p.chunks0 : c) * p.ck, w0 + p.dw_min, h0 + p.dh_min, img);
          }
          if (++stage == p.stages_a) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    const uint32_t leader = elect_one();
    const uint32_t idesc = p.idesc;
    const uint32_t hi_a = smem_desc_hi(8 * p.pitch_a, p.layout_a), hi_b = smem_desc_hi(8 * p.pitch_b, p.layout_b);
    const uint32_t ka16 = ((uint32_t)p.PW * p.pitch_a) >> 4;  // next tile row of the halo raster
    const uint32_t kb16 = (16 * p.pitch_b) >> 4;              // next 16 pixels of the dense dY tile
    const uint32_t a_lo0 = smem_u32(smem) >> 4;
    const uint32_t b_lo0 = smem_desc_lo(smem_u32(smem_b), p.lbo_b);
    const int ksteps = p.tile_rows;
    const uint32_t a_stage16 = (uint32_t)a_stage >> 4, b_tile16 = (uint32_t)b_tile >> 4;
    const uint32_t block_n = (uint32_t)p.block_n;
    const int stages_a = p.stages_a;
    int stage = 0;
    uint32_t phase = 0;
    for (int i = 0; i < n_my_kb; ++i) {
      const int bs = i & 1;
      mbar_wait(&fullB[bs], (i >> 1) & 1, p.err);
      tc_fence_after();
      const uint32_t b_lo = b_lo0 + (uint32_t)bs * b_tile16;
      for (int cg = cg_lo; cg <= cg_hi; ++cg) {
        mbar_wait(&fullA[stage], phase, p.err);
        tc_fence_after();
        if (leader) {  // one divergent region per halo stage: all of its slabs back to back
          const uint32_t st_lo = a_lo0 + (uint32_t)stage * a_stage16;
          const int s_lo = max((int)p.cg_first[cg], slab0), s_hi = min((int)p.cg_first[cg + 1], slab1);
          for (int s = s_lo; s < s_hi; ++s) {
            const uint32_t a_lo = (st_lo + p.slabs[s].a_off16) | (p.slabs[s].lbo16 << 16);
            const uint32_t td = tmem_base + (uint32_t)(s - slab0) * block_n;
            umma_bf16_lohi(td, a_lo, hi_a, b_lo, hi_b, idesc, i > 0 ? 1u : 0u);
#pragma unroll
            for (int k = 1; k < 8; ++k) umma_bf16_lohi(td, a_lo + k * ka16, hi_a, b_lo + k * kb16, hi_b, idesc, 1u);
            for (int k0 = 8; k0 < ksteps; k0 += 8) {  // taller K blocks (row-shift mode): 8 more rows per trip
#pragma unroll
              for (int k = 0; k < 8; ++k)
                umma_bf16_lohi(td, a_lo + (k0 + k) * ka16, hi_a, b_lo + (k0 + k) * kb16, hi_b, idesc, 1u);
            }
          }
          umma_commit(&emptyA[stage]);
        }
        if (++stage == stages_a) {
          stage = 0;
          phase ^= 1;
        }
      }
      if (leader) umma_commit(&emptyB[bs]);
      __syncwarp();
    }
    if (leader) umma_commit(&tmem_full_bar);
    __syncwarp();
  } else {
    const int q = warp & 3;
    const int m = q * 32 + lane;
    if (n_my_kb > 0) {
      mbar_wait(&tmem_full_bar, 0, p.err);
      tc_fence_after();
      const int sub = m / p.ck, within = m - sub * p.ck;
      const int nch = min(p.block_n, p.Cout - n0);
      if (p.rowshift) {
        // accumulator column c = (kernel-row group i, output channel b), row m = (kernel column j = sub, input channel)
        for (int s = slab0; s < slab1; ++s) {
          const int a_ch = p.slabs[s].sub_c[sub] * p.ck + within;
          for (int c = 0; c < p.block_n; c += 16) {
            uint32_t v[16];
            tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)((s - slab0) * p.block_n + c), v);
            tmem_ld_wait();
            const int gi = c / p.rs_cout, b0 = c - gi * p.rs_cout;
            const int kc = p.slabs[s].sub_t[sub];  // kernel column of this A sub-block
            const int t = kc >= 0 ? p.rs_tap[gi * 8 + kc] : -1;
            if (t < 0) continue;
            if (p.part) {
              float* pb = p.part + (long long)blockIdx.x * p.part_stride + ((long long)t * p.Cout + b0) * p.part_A + a_ch;
#pragma unroll
              for (int i = 0; i < 16; ++i)
                if (b0 + i < p.Cout) pb[(long long)i * p.part_A] = __uint_as_float(v[i]);
            } else {
              float* base = p.dw + (long long)t * p.tap_stride + (long long)a_ch * p.a_stride + (long long)b0 * p.b_stride;
#pragma unroll
              for (int i = 0; i < 16; ++i)
                if (b0 + i < p.Cout) atomicAdd(base + (long long)i * p.b_stride, __uint_as_float(v[i]));
            }
          }
        }
      } else
      for (int s = slab0; s < slab1; ++s) {
        const int t = p.slabs[s].sub_t[sub];
        const bool rvalid = t >= 0;
        const int a_ch = p.slabs[s].sub_c[sub] * p.ck + within;
        float* base = p.dw + (long long)(rvalid ? t : 0) * p.tap_stride + (long long)a_ch * p.a_stride +
                      (long long)n0 * p.b_stride;
        float* pbase = p.part ? p.part + (long long)blockIdx.x * p.part_stride +
                                    ((long long)(rvalid ? t : 0) * p.Cout + n0) * p.part_A + a_ch
                              : nullptr;
        for (int c = 0; c < p.block_n; c += 16) {
          uint32_t v[16];
          tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)((s - slab0) * p.block_n + c), v);
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

}  // namespace pdac

// ------------------------------------------------------------------------------------------------ host side
namespace pdac {

// Returns PDAC_OK with *handled = 1 when the halo kernel was launched, *handled = 0 when the geometry is not eligible
// (the caller falls back to the generic kernel), or an error code.
int wgrad_halo_try(const PdacTapConv* tc, int n_views, const PdacView* views, const PdacView* dy, float* dw,
                   int64_t tap_stride, int64_t a_stride, int64_t b_stride, int b_valid, float* scratch,
                   long long scratch_bytes, void* stream, int* handled) {
  *handled = 0;
  static int disabled = -1;
  if (disabled < 0) {
    const char* e = getenv("PDAC_NO_HALO");
    disabled = (e && e[0] == '1') ? 1 : 0;
  }
  if (disabled) return PDAC_OK;
  if (tc->n_taps < 2 || n_views != tc->n_cat) return PDAC_OK;
  if (dy->W < 16 || dy->H < 8) return PDAC_OK;
  int dh_min = 127, dh_max = -127, dw_min = 127, dw_max = -127;
  for (int t = 0; t < tc->n_taps; ++t) {
    if (tc->taps[t].src != 0) return PDAC_OK;
    dh_min = tc->taps[t].dh < dh_min ? tc->taps[t].dh : dh_min;
    dh_max = tc->taps[t].dh > dh_max ? tc->taps[t].dh : dh_max;
    dw_min = tc->taps[t].dw < dw_min ? tc->taps[t].dw : dw_min;
    dw_max = tc->taps[t].dw > dw_max ? tc->taps[t].dw : dw_max;
  }
  const int eh = dh_max - dh_min, ew = dw_max - dw_min;
  if (16 + ew > 256 || 8 + eh > 256) return PDAC_OK;
  int ck = 64;
  for (int s = 0; s < tc->n_cat; ++s) {
    if (views[s].N != dy->N || views[s].H != dy->H || views[s].W != dy->W) return PDAC_OK;
    while (ck > 8 && (views[s].C % ck)) ck >>= 1;
  }
  if (ck < 16 || dy->C % 16) return PDAC_OK;
  const int cout = dy->C;
  if (cout > 128 && cout % 128) return PDAC_OK;
  int block_n = cout > 128 ? 128 : cout;  // wider outputs: 128-column pieces over grid.z (dY is split, X re-read)
  const int zgroups = cout / block_n;
  int ckb = 64;
  while (cout % ckb) ckb >>= 1;

  // Row-shift mode for 16 / 32-channel tensors (the 512^2 and 256^2 decoder layers). With ck <= 32 an M = 128 slab holds
  // only the taps of ONE kernel row (3 x 16 or 3 x 32 of its 128 rows), i.e. three MMAs per 16 pixels, each costing
  // ~32 + N/4 cycles whatever N is (scripts/probe/mma_rate.cu) - the layer is MMA-issue bound at a third of its HBM
  // rate. Here the kernel-ROW offsets move to the dY operand: dY is loaded with the h halo, the accumulator's N
  // dimension becomes (kernel row, output channel) = 3 x Cout columns through B sub-blocks one image row apart (LBO),
  // and the A operand keeps the kernel columns as sub-blocks one pixel apart: ONE MMA per 16 pixels covers all taps.
  // With 64-channel chunks (PDAC_WGRAD_ROWSHIFT=2) the A sub-blocks of a slab are the two chunks of a stage for one kernel
  // column, or two kernel columns of one chunk: 3 (or 2) slabs of N = 3 x Cout instead of 9 (or 5) of N = Cout.
  static int rs_ok = -1;
  if (rs_ok < 0) {
    const char* e = getenv("PDAC_WGRAD_ROWSHIFT");
    rs_ok = e ? atoi(e) : 2;
  }
  int rs = 0, rs_R = 0, rs_S = 0, rs_d = 1;
  if (rs_ok && (ck <= 32 || rs_ok >= 2) && cout == ckb && zgroups == 1) {
    int S = 1;
    while (S < tc->n_taps && tc->taps[S].dh == tc->taps[0].dh) ++S;
    const int Rr = tc->n_taps / S;
    if (Rr * S == tc->n_taps && Rr <= 8 && S <= 8) {
      int d = S > 1 ? tc->taps[1].dw - tc->taps[0].dw : (Rr > 1 ? tc->taps[S].dh - tc->taps[0].dh : 1);
      bool grid_ok = d > 0;
      for (int t = 0; t < tc->n_taps && grid_ok; ++t)
        grid_ok = tc->taps[t].dh == dh_min + (t / S) * d && tc->taps[t].dw == dw_min + (t % S) * d;
      const int n_chunks_ = views[0].C / ck + (tc->n_cat > 1 ? views[1].C / ck : 0);
      if (grid_ok && (S * ck <= 128 || ck == 64) && Rr * cout <= 256 && n_chunks_ * S <= kWhMaxSlabs &&
          d * 16 * ckb * 2 < (1 << 18)) {
        rs = 1;
        rs_R = Rr;
        rs_S = S;
        rs_d = d;
        block_n = Rr * cout;
      }
    }
  }

  WhMaps maps;
  WhParams p;
  memset(&p, 0, sizeof(p));
  p.ck = ck;
  p.chunks0 = views[0].C / ck;
  p.n_chunks = p.chunks0 + (tc->n_cat > 1 ? views[1].C / ck : 0);
  p.PW = 16 + ew;
  p.pitch_a = ck * 2;
  p.layout_a = swizzle_layout_type(p.pitch_a);
  p.pitch_b = ckb * 2;
  p.layout_b = swizzle_layout_type(p.pitch_b);
  p.dw_min = dw_min;
  p.rowshift = rs;
  p.cps = (ck == 64 && p.n_chunks >= 2) ? 2 : 1;
  p.n_cg = ceil_div(p.n_chunks, p.cps);
  if (rs) {
    // K block = 16 pixels x tile_rows rows: as tall as three A stages + two dY tiles allow (two CTAs per SM when the
    // slabs fit half of the tensor memory)
    int ns_est = 0;
    for (int cg = 0; cg < p.n_cg; ++cg) {
      const int nc = p.n_chunks - cg * p.cps < p.cps ? p.n_chunks - cg * p.cps : p.cps;
      ns_est += (ck == 64 && nc == 2) ? rs_S : nc * ceil_div(rs_S, 128 / ck);
    }
    const int budget = (ns_est * block_n <= 256 ? 104 : 218) * 1024;
    int tr = 32;
    while (tr > 8 && (tr > dy->H || ((tr * p.PW * (int)p.pitch_a + 1023) & ~1023) * p.cps * 3 +
                                        2 * (((tr + eh) * 16 * (int)p.pitch_b + 1023) & ~1023) + 2048 > budget))
      tr >>= 1;
    p.tile_rows = tr;
    p.R = tr;           // the source box has no h halo in this mode ...
    p.dh_min = 0;
    p.b_rows = tr + eh; // ... dY has it
    p.b_row0 = -dh_max;
    p.b_tx_bytes = p.b_rows * 16 * (int)p.pitch_b;
    p.b_tile_bytes = (p.b_tx_bytes + 1023) & ~1023;
    p.lbo_b = (uint32_t)(rs_d * 16 * (int)p.pitch_b);
    p.rs_cout = cout;
    for (int i = 0; i < 8; ++i)
      for (int j = 0; j < 8; ++j) p.rs_tap[i * 8 + j] = (int8_t)((i < rs_R && j < rs_S) ? (rs_R - 1 - i) * rs_S + j : -1);
  } else {
    p.tile_rows = 8;
    p.R = 8 + eh;
    p.dh_min = dh_min;
    p.b_rows = 8;
    p.b_row0 = 0;
    p.b_tx_bytes = p.b_tile_bytes = 128 * block_n * 2;
    p.lbo_b = 128u * p.pitch_b;
  }
  p.a_buf = (p.R * p.PW * (int)p.pitch_a + 1023) & ~1023;

  // taps in ascending halo-raster shift
  int order[PDAC_MAX_TAPS], shift[PDAC_MAX_TAPS];
  for (int t = 0; t < tc->n_taps; ++t) {
    order[t] = t;
    shift[t] = (tc->taps[t].dh - dh_min) * p.PW + (tc->taps[t].dw - dw_min);
  }
  for (int i = 1; i < tc->n_taps; ++i)
    for (int j = i; j > 0 && shift[order[j]] < shift[order[j - 1]]; --j) {
      const int tmp = order[j];
      order[j] = order[j - 1];
      order[j - 1] = tmp;
    }
  for (int i = 1; i < tc->n_taps; ++i)
    if (shift[order[i]] == shift[order[i - 1]]) return PDAC_OK;
  const int subs = 128 / ck;
  int ns = 0;
  auto new_slab = [&](int cg) -> WhSlab* {
    if (ns >= kWhMaxSlabs) return nullptr;
    WhSlab* s = &p.slabs[ns++];
    s->cg = (int8_t)cg;
    for (int i = 0; i < 8; ++i) {
      s->sub_t[i] = -1;
      s->sub_c[i] = 0;
    }
    return s;
  };
  for (int cg = 0; cg < p.n_cg; ++cg) {
    const int c_first = cg * p.cps;
    const int nc = p.n_chunks - c_first < p.cps ? p.n_chunks - c_first : p.cps;
    if (rs) {
      // row-shift slabs: every slab covers ALL kernel rows (they sit in the accumulator's N dimension); its A
      // sub-blocks are kernel COLUMNS (sub_t = column index, rs_d pixels apart) of one chunk, or, for 64-channel chunk
      // pairs, the two chunks of the stage for one kernel column (a_buf apart)
      if (ck == 64 && nc == 2) {
        for (int j = 0; j < rs_S; ++j) {
          WhSlab* sl = new_slab(cg);
          if (!sl) return PDAC_OK;
          sl->a_off16 = (uint32_t)(j * rs_d * (int)p.pitch_a) >> 4;
          sl->lbo16 = (uint32_t)p.a_buf >> 4;
          sl->sub_t[0] = sl->sub_t[1] = (int8_t)j;
          sl->sub_c[0] = (int8_t)c_first;
          sl->sub_c[1] = (int8_t)(c_first + 1);
        }
      } else {
        for (int jc = 0; jc < nc; ++jc)
          for (int j0 = 0; j0 < rs_S; j0 += subs) {
            WhSlab* sl = new_slab(cg);
            if (!sl) return PDAC_OK;
            sl->a_off16 = (uint32_t)(jc * p.a_buf + j0 * rs_d * (int)p.pitch_a) >> 4;
            sl->lbo16 = (uint32_t)(rs_d * (int)p.pitch_a) >> 4;
            for (int k = 0; k < subs && j0 + k < rs_S; ++k) {
              sl->sub_t[k] = (int8_t)(j0 + k);
              sl->sub_c[k] = (int8_t)(c_first + jc);
            }
          }
      }
      continue;
    }
    if (ck == 64 && nc == 2) {
      for (int i = 0; i < tc->n_taps; ++i) {  // M = (tap; chunk c_first | chunk c_first + 1)
        WhSlab* s = new_slab(cg);
        if (!s) return PDAC_OK;
        const int t = order[i];
        s->a_off16 = (uint32_t)(shift[t] * (int)p.pitch_a) >> 4;
        s->lbo16 = (uint32_t)p.a_buf >> 4;
        s->sub_t[0] = s->sub_t[1] = (int8_t)t;
        s->sub_c[0] = (int8_t)c_first;
        s->sub_c[1] = (int8_t)(c_first + 1);
      }
    } else {
      for (int j = 0; j < nc; ++j) {
        const int c = c_first + j;
        int i = 0;
        while (i < tc->n_taps) {  // greedy arithmetic runs of up to `subs` taps
          WhSlab* s = new_slab(cg);
          if (!s) return PDAC_OK;
          int len = 1, d = 1;
          if (i + 1 < tc->n_taps) {
            d = shift[order[i + 1]] - shift[order[i]];
            len = 2;
            while (len < subs && i + len < tc->n_taps && shift[order[i + len]] - shift[order[i + len - 1]] == d) ++len;
          }
          if ((long long)d * p.pitch_a >= (1 << 18)) {  // LBO field is 14 bits of 16-byte units
            len = 1;
            d = 1;
          }
          s->a_off16 = (uint32_t)(j * p.a_buf + shift[order[i]] * (int)p.pitch_a) >> 4;
          s->lbo16 = (uint32_t)(d * (int)p.pitch_a) >> 4;
          for (int k = 0; k < len; ++k) {
            s->sub_t[k] = (int8_t)order[i + k];
            s->sub_c[k] = (int8_t)c;
          }
          i += len;
        }
      }
    }
  }
  p.n_slabs = ns;
  {
    int s = 0;
    for (int cg = 0; cg <= p.n_cg; ++cg) {
      while (s < ns && p.slabs[s].cg < cg) ++s;
      p.cg_first[cg] = (uint8_t)s;
    }
  }
  static int dual_ok = -1;
  if (dual_ok < 0) {
    const char* e = getenv("PDAC_WGRAD_DUAL");
    dual_ok = e ? atoi(e) : 1;
  }
  // two CTAs per SM when all slabs fit in half of the tensor memory (the 16- and 32-channel layers) or, for N <= 64,
  // in two slab groups over grid.y (64 -> 64: 0.136 -> 0.119 ms; N = 128 loses, 32 -> 128: 0.343 -> 0.377 ms).
  // PDAC_WGRAD_DUAL=2 (experiment): whenever the shared memory allows
  int dual = dual_ok && (ns * block_n <= 256 || (block_n <= 64 && ns * block_n <= 512) || dual_ok == 2);
  {
    const int a_stage_ = p.a_buf * p.cps;
    if (dual && (104 * 1024 - 2 * p.b_tile_bytes - 1024) / a_stage_ < 3) dual = 0;
  }
  int spc_max = (dual ? 256 : 512) / block_n;
  if (spc_max > ns) spc_max = ns;
  const int ygroups = ceil_div(ns, spc_max);
  p.spc = ceil_div(ns, ygroups);
  p.tmem_cols = pow2_ceil(p.spc * block_n < 32 ? 32 : p.spc * block_n);
  p.ckb = ckb;
  p.nboxb = rs ? 1 : block_n / ckb;
  p.block_n = block_n;
  p.idesc = make_idesc_bf16(128, block_n, 1, 1);
  p.tiles_w = ceil_div(dy->W, 16);
  p.tiles_h = ceil_div(dy->H, p.tile_rows);
  p.n_kblocks = p.tiles_w * p.tiles_h * dy->N;
  p.dw = dw;
  p.tap_stride = tap_stride;
  p.a_stride = a_stride;
  p.b_stride = b_stride;
  p.Cout = b_valid > 0 ? b_valid : cout;
  p.err = device_err_flag();
  const int b_tile = p.b_tile_bytes;
  const int a_stage = p.a_buf * p.cps;
  int stages_a = ((dual ? 104 : 220) * 1024 - 2 * b_tile - 1024) / a_stage;
  if (stages_a > kWhMaxStages) stages_a = kWhMaxStages;
  if (stages_a < 2) return PDAC_OK;
  p.stages_a = stages_a;
  for (int s = 0; s < tc->n_cat; ++s) PDAC_TRY(encode_view_map(&maps.a[s], &views[s], ck, p.PW, p.R, 1));
  if (tc->n_cat == 1) maps.a[1] = maps.a[0];
  PDAC_TRY(encode_view_map(&maps.dy, dy, ckb, 16, p.b_rows, 1));
  int splits = (dual ? 2 : 1) * num_sms() / (ygroups * zgroups);
  if (splits < 1) splits = 1;
  if (splits > p.n_kblocks) splits = p.n_kblocks;
  const int a_total = p.n_chunks * ck;
  PDAC_TRY(wgrad_plan_partials(scratch, scratch_bytes, tc->n_taps, a_total, p.Cout, &splits, &p.part, &p.part_stride));
  p.part_A = a_total;
  const size_t smem = (size_t)stages_a * a_stage + 2 * (size_t)b_tile + 1024;
  static bool attr_set = false;
  if (!attr_set) {
    PDAC_CUDA(cudaFuncSetAttribute(wgrad_halo_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(222 * 1024)));
    PDAC_CUDA(cudaFuncSetAttribute(wgrad_halo_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(105 * 1024)));
    PDAC_CUDA(cudaFuncSetAttribute(wgrad_halo_kernel<1>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                   cudaSharedmemCarveoutMaxShared));
    attr_set = true;
  }
  dim3 grid(splits, ygroups, zgroups);
  if (dual)
    wgrad_halo_kernel<1><<<grid, kWhThreads, smem, (cudaStream_t)stream>>>(maps, p);
  else
    wgrad_halo_kernel<0><<<grid, kWhThreads, smem, (cudaStream_t)stream>>>(maps, p);
  PDAC_LAUNCH_CHECK("wgrad_halo_kernel");
  if (p.part)
    PDAC_TRY(wgrad_reduce(p.part, splits, p.part_stride, tc->n_taps, p.Cout, a_total, dw, tap_stride, a_stride, b_stride,
                          stream));
  note_kernel(KID_WGRAD_HALO);
  *handled = 1;
  return PDAC_OK;
}

}  // namespace pdac
