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
// Roles: warp 0 = A (halo) TMA producer, warp 1 = TMEM allocator + MMA issuer, warps 2.. = two or four epilogue
// warpgroups (accumulators of a tile are dealt round-robin), last warp = weight TMA producer.
//
// MODE picks the epilogue (separate instantiations: each sits at the register limit of its launch bounds):
//   0 staged (TMEM -> registers -> bf16 staging tile -> 16-byte coalesced stores + BatchNorm partial sums)
//   1 staged, folded-BatchNorm inference (per-channel scale / shift)
//   2 staged, the leading channels summed over 2 x 2 pixel blocks (data gradient in front of a nearest x2 upsampling)
//   3 direct (TMEM -> registers -> 32-byte stores, no staging, no barriers): every launch without BatchNorm sums
//   4 direct, folded-BatchNorm inference
//   5 direct + BatchNorm sums of the plain N = 16 convolution (lane pairs split the channels: 16 sum registers)
//   7 staged + tap stacking (N <= 32 with BatchNorm sums); modes 3-5 take stacking through HaloParams::stack
//
// This header holds the device code; it is compiled in four translation units (conv_halo_k0..k3.cu, one group of
// epilogue modes each) so that `make -j` builds them in parallel, and conv_halo.cu (host planner + launch) looks the
// instantiations up through halo_kernel_k0..k3(). The diagnostic build (-DPDAC_HALO_PROF) is a single translation unit
// (conv_halo.cu includes the four files) because the per-CTA counters are plain __device__ globals.
#pragma once
#include "common.h"
#include "ptx.cuh"
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

namespace pdac {

struct HaloMaps {
  CUtensorMap a[2];  // the (up to two concatenated) source views, box (ck, PW, R, 1)
  CUtensorMap w;     // packed weights [n_taps*wrows][Ktot], box (ck, block_n)
};

struct HaloParams {
  int n_taps, n_chunks, chunks0, ck;
  // taps form an affine grid: tap t = r * tap_S + s reads the halo tap_base + r * tap_step_r + s * tap_step_s pixel rows
  // further on ((dh - dh_min) * PW + (dw - dw_min))
  int tap_R, tap_S, tap_base, tap_step_r, tap_step_s;
  int dh_min, dw_min;
  uint32_t pitch, layout_type, idesc, tmem_cols;
  int block_n, tiles_c, wrows;
  int nacc, TW, TH, PW, R;
  uint32_t inv_pw;  // ceil(2^20 / PW): o / PW == (o * inv_pw) >> 20 for o < 1024
  int n_wg;         // epilogue warpgroups in use (2 or 4)
  int w_warp;       // index of the weight-producer warp (= 2 + 4 * n_wg in the two-CTAs-per-SM variant, else 18)
  int tiles_w, tiles_h, tiles_sp;  // spatial tiles per image row / column, total over the batch
  int a_buf, sa;                   // bytes per halo stage, stages
  int b_resident, sb, b_sub;       // weights resident? else ring stages; bytes per (tap, chunk) weight tile
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
  // POOL mode (data gradient in front of a nearest-neighbour x2 upsampling): output channels [0, pool_c) are summed
  // over 2x2 pixel blocks in the epilogue and stored to the half-resolution tensor pout; the remaining channels go
  // to `out` (whose pointer the host has moved back by pool_c channels). PW is a power of two (pw_log2) that divides
  // 128 with an even quotient, so every accumulator holds whole pixel-row pairs.
  int pool_c, pw_log2;
  // Tap stacking (direct epilogue, N = 16, one K chunk): the three taps of a kernel row share ONE MMA whose B operand is
  // their three weight tiles (contiguous in shared memory: N' = 3 N) — the 4 KB pixel tile, whose shared-memory read is
  // what an MMA costs at small N (32 of its 32 + N/4 cycles), is read once per kernel row instead of once per tap.
  // Accumulator column block s then holds the products of tap (r, s) for the pixel stack_d[s] raster positions to the
  // left, and the epilogue adds the three blocks across lanes (PW is 16 or 32: an image row never leaves its warp).
  int stack, acc_cols, issue_S, b_tap_mul;
  int stack_d[3];
  // direct epilogue only: store every output pixel to the 2 x 2 block it becomes under a nearest x2 upsampling (`out`
  // then addresses the even pixels of the (2H, 2W) tensor: osH / osW are twice its strides, up_dh / up_dw its own)
  int up2;
  long long up_dh, up_dw;
  bf16* pout;
  long long psN, psH, psW;
};

// -DPDAC_HALO_PROF (make prof): per-CTA cycle counters of where each role waits, read back with pdac_halo_prof_read
#ifdef PDAC_HALO_PROF
__device__ long long g_halo_prof[160 * 16];
__device__ int g_halo_dbg;  // bit 0: no global stores, 1: epilogue only releases TMEM, 2: no MMAs, 3: no statistics
#define DBG(bit) (dbg_mask & (1 << (bit)))
#define PROF_DECL(x) long long x = 0
#define PROF_WAIT(accum, stmt) \
  {                            \
    long long _t = clock64();  \
    stmt;                      \
    accum += clock64() - _t;   \
  }
#define PROF_PUT(slot, v) g_halo_prof[(blockIdx.x % 160) * 16 + (slot)] = (v)
#else
#define PROF_DECL(x)
#define PROF_WAIT(accum, stmt) \
  { stmt; }
#define PROF_PUT(slot, v)
#define DBG(bit) 0
#endif

static constexpr int kHaloThreads = 64 + 4 * 128 + 32;  // TMA warp, MMA warp, up to 4 epilogue warpgroups, weight warp
// Two-CTAs-per-SM variant (low channel counts: small tiles, small weights): TMA warp, MMA warp, 2 epilogue warpgroups,
// weight warp. tcgen05.mma issue is in-order and nearly synchronous with execution (scripts/probe/mma_issue.cu: the
// issuing thread is held at the execution rate, the queue is only a few instructions deep), so every cycle the single
// issuing thread spends on barrier waits, commits and loop overhead idles the tensor pipe; a second resident CTA has
// its own issuing warp and fills those gaps.
static constexpr int kHaloThreadsDual = 64 + 2 * 128 + 32;
static constexpr int kHaloMaxSA = 12;
static constexpr int kHaloMaxSB = 8;

__device__ __forceinline__ float h_bf16lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float h_bf16hi(uint32_t u) { return __uint_as_float(u & 0xffff0000u); }
__device__ __forceinline__ void sts128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ void stg256(void* ptr, const uint32_t (&r)[8]) {  // one 32-byte sector per thread
  asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(ptr), "r"(r[0]), "r"(r[1]), "r"(r[2]),
               "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
               : "memory");
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ unsigned short lds16(uint32_t addr) {
  unsigned short v;
  asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t h_pack(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}

// The NACC x KS MMAs of one (tap, channel chunk): accumulator j reads the halo 128 pixel rows further on.
template <int NACC, int KS>
__device__ __forceinline__ void halo_issue(uint32_t tacc, uint32_t acc_step, uint32_t a_lo, uint32_t j_step, uint32_t b_lo,
                                           uint32_t desc_hi, uint32_t idesc, uint32_t acc) {
#pragma unroll
  for (int j = 0; j < NACC; ++j) {
#pragma unroll
    for (int kk = 0; kk < KS; ++kk)  // +2 in the 16-byte-unit address field = 16 bf16 along K
      umma_bf16_lohi(tacc + (uint32_t)j * acc_step, a_lo + (uint32_t)j * j_step + 2 * kk, desc_hi, b_lo + 2 * kk, desc_hi,
                     idesc, kk ? 1u : acc);
  }
}

// MODE 1 (AFF): folded-BatchNorm inference epilogue (per-channel scale and shift); MODE 2 (POOL): 2x2 sum-pooled stores
// (see HaloParams::pool_c). Separate instantiations because the epilogue sits at the register limit and the training
// kernels must not pay for either.
template <int NACC, int KS, int DUAL, int MODE>
__global__ void __launch_bounds__(DUAL ? kHaloThreadsDual : kHaloThreads, DUAL ? 2 : 1)
    conv_halo_kernel(const __grid_constant__ HaloMaps maps, const __grid_constant__ HaloParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_a[kHaloMaxSA];
  __shared__ __align__(8) uint64_t empty_a[kHaloMaxSA];
  __shared__ __align__(8) uint64_t full_b[kHaloMaxSB];
  __shared__ __align__(8) uint64_t empty_b[kHaloMaxSB];
  __shared__ __align__(8) uint64_t tmem_full_bar[2];
  __shared__ __align__(8) uint64_t tmem_empty_bar[2];
  __shared__ uint32_t tmem_base_slot;
  __shared__ long long s_rowoff[4][128];
  __shared__ long long s_pooloff[MODE == 2 ? (DUAL ? 2 : 4) : 1][32];
  constexpr int AFF = MODE == 1 || MODE == 4;
  constexpr int POOL = MODE == 2;
  constexpr int DIRECT = MODE >= 3 && MODE <= 5;  // register -> global epilogue (see the epilogue role)
  constexpr int DSTATS = MODE == 5;
  constexpr int SSTACK = MODE == 7;  // staged epilogue (BatchNorm sums) of a tap-stacked N <= 32 layer  // ... of a plain N = 16 convolution in front of a BatchNorm: store + partial sums

  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#ifdef PDAC_HALO_PROF
  const int dbg_mask = g_halo_dbg;
  long long t_k0 = 0, g_k0 = 0;
  if (threadIdx.x == 0) {
    t_k0 = clock64();
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g_k0));
  }
#endif

  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + (size_t)p.sa * p.a_buf;
  const int b_bytes = (p.b_resident ? p.n_taps * p.n_chunks : p.sb * p.tap_S) * p.b_sub;
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
      mbar_init(&tmem_empty_bar[i], 128 * p.n_wg);
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

  // Spatial tile walker: tile index sp = (img * tiles_h + tile_h) * tiles_w + tile_w advances by sp_step per iteration;
  // the decomposition is carried incrementally (a handful of integer divisions per tile and thread used to cost more
  // than the epilogue proper on the 16-channel layers).
  struct Walk {
    int tw, th, img, dw, dh, dimg;
  };
  auto walk_init = [&](Walk& k) {
    int r = sp0;
    k.tw = r % p.tiles_w;
    r /= p.tiles_w;
    k.th = r % p.tiles_h;
    k.img = r / p.tiles_h;
    int d = sp_step;
    k.dw = d % p.tiles_w;
    d /= p.tiles_w;
    k.dh = d % p.tiles_h;
    k.dimg = d / p.tiles_h;
  };
  auto walk_next = [&](Walk& k) {
    k.tw += k.dw;
    if (k.tw >= p.tiles_w) {
      k.tw -= p.tiles_w;
      ++k.th;
    }
    k.th += k.dh;
    if (k.th >= p.tiles_h) {
      k.th -= p.tiles_h;
      ++k.img;
    }
    k.img += k.dimg;
  };

  if (warp == 0) {
    // ------------------------------------------------------------------ halo producer
    if (lane == 0 && sp0 < sp_step) {
      const uint32_t halo_bytes = (uint32_t)(p.R * p.PW) * p.pitch;
      int stage = 0;
      uint32_t phase = 0;  // ring position and its wrap parity, carried incrementally (no divisions in the loops)
      Walk wk;
      walk_init(wk);
      PROF_DECL(w_ea);
      for (int sp = sp0; sp < p.tiles_sp; sp += sp_step, walk_next(wk)) {
        const int img = wk.img, w0 = wk.tw * p.TW, h0 = wk.th * p.TH;
        for (int c = 0; c < p.n_chunks; ++c) {
          PROF_WAIT(w_ea, mbar_wait(&empty_a[stage], phase ^ 1, p.err));
          mbar_expect_tx(&full_a[stage], halo_bytes);
          const int s = c >= p.chunks0;
          tma_load_4d(smem_a + (size_t)stage * p.a_buf, &maps.a[s], &full_a[stage], (s ? c - p.chunks0 : c) * p.ck,
                      w0 + p.dw_min, h0 + p.dh_min, img);
          if (++stage == p.sa) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
      PROF_PUT(7, w_ea);
    }
  } else if (warp == p.w_warp) {
    // ------------------------------------------------------------------ weight producer
    if (lane == 0 && sp0 < sp_step) {
      if (p.b_resident) {
        mbar_expect_tx(&full_b[0], (uint32_t)b_bytes);
        for (int t = 0; t < p.n_taps; ++t)
          for (int c = 0; c < p.n_chunks; ++c)
            // tap stacking: the three taps of a kernel row lie side by side per (row, chunk) = one N' = 3 N operand
            tma_load_2d(smem_b + (size_t)(p.stack ? ((t / 3) * p.n_chunks + c) * 3 + t % 3 : t * p.n_chunks + c) * p.b_sub,
                        &maps.w, &full_b[0], c * p.ck, t * p.wrows + n0);
      } else {
        int stage = 0;
        uint32_t phase = 0;
        PROF_DECL(w_eb);
        for (int sp = sp0; sp < p.tiles_sp; sp += sp_step) {
          for (int c = 0; c < p.n_chunks; ++c) {
            for (int r = 0; r < p.tap_R; ++r) {  // one stage = the tap_S taps of a kernel row
              PROF_WAIT(w_eb, mbar_wait(&empty_b[stage], phase ^ 1, p.err));
              mbar_expect_tx(&full_b[stage], (uint32_t)(p.tap_S * p.b_sub));
              for (int s = 0; s < p.tap_S; ++s)
                tma_load_2d(smem_b + (size_t)(stage * p.tap_S + s) * p.b_sub, &maps.w, &full_b[stage], c * p.ck,
                            (r * p.tap_S + s) * p.wrows + n0);
              if (++stage == p.sb) {
                stage = 0;
                phase ^= 1;
              }
            }
          }
        }
        PROF_PUT(8, w_eb);
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    // The whole warp walks the loops (warp-uniform control flow: descriptors and TMEM addresses live in uniform
    // registers); one elected lane issues. Measured (scripts/probe/mma_rate.cu): one M=128 K=16 MMA executes in
    // ~32 + N/4 cycles, so the issue path must stay well below ~40 cycles per instruction: the (accumulator, k-step)
    // block is compile-time unrolled and the taps are walked as an affine (r, s) grid with no table look-ups.
    if (sp0 < sp_step) {
      const uint32_t leader = elect_one();
#ifdef PDAC_HALO_PROF
      // bit 4: every MMA computes N = 16 columns only (garbage results; issue-bound vs execution-bound test)
      const uint32_t idesc = DBG(4) ? ((p.idesc & ~(0x3Fu << 17)) | (2u << 17)) : p.idesc;
#else
      const uint32_t idesc = p.idesc;
#endif
      const uint32_t desc_hi = (uint32_t)(make_smem_desc(0, 16, 8 * p.pitch, p.layout_type) >> 32);
      const uint32_t lbo_lo = (16u >> 4) << 16;
      if (p.b_resident) {
        mbar_wait(&full_b[0], 0, p.err);
        tc_fence_after();
      }
      const uint32_t pix16 = p.pitch >> 4;  // 16-byte units per halo pixel row
      const uint32_t b_lo0 = (smem_u32(smem_b) >> 4) | lbo_lo;
      const uint32_t a_lo0 = (smem_u32(smem_a) >> 4) | lbo_lo;
      const uint32_t a_buf16 = (uint32_t)p.a_buf >> 4, b_sub16 = (uint32_t)p.b_sub >> 4;
      const uint32_t acc_step = (uint32_t)p.acc_cols;
      const uint32_t j_step = 128u * pix16;
      const uint32_t tap0 = (uint32_t)p.tap_base * pix16;
      const uint32_t step_r = (uint32_t)(p.tap_step_r * (int)pix16), step_s = (uint32_t)(p.tap_step_s * (int)pix16);
      const int TR = p.tap_R, TS = p.issue_S, n_chunks = p.n_chunks, sa = p.sa, sb = p.sb;
      const uint32_t b_tap16 = (uint32_t)(p.b_tap_mul * n_chunks) * b_sub16;
      int stage = 0, bstage = 0, lt = 0;
      uint32_t phase = 0, bphase = 0;
      PROF_DECL(w_te);
      PROF_DECL(w_fa);
      PROF_DECL(w_fb);
      PROF_DECL(w_cm);
      PROF_DECL(w_sw);
      PROF_DECL(w_fn);
      PROF_DECL(w_ld);
      PROF_DECL(t_mma0);
#ifdef PDAC_HALO_PROF
      t_mma0 = clock64();
#endif
      for (int sp = sp0; sp < p.tiles_sp; sp += sp_step, ++lt) {
        const int g = lt & 1;
        PROF_WAIT(w_te, mbar_wait(&tmem_empty_bar[g], ((lt >> 1) & 1) ^ 1, p.err));
        PROF_WAIT(w_fn, tc_fence_after());
        const uint32_t tacc = tmem_base + (uint32_t)(g * NACC) * acc_step;
        for (int c = 0; c < n_chunks; ++c) {
          PROF_WAIT(w_fa, mbar_wait(&full_a[stage], phase, p.err));
          PROF_WAIT(w_fn, tc_fence_after());
          const uint32_t a_stage = a_lo0 + (uint32_t)stage * a_buf16 + tap0;
          if (p.b_resident) {
#ifdef PDAC_HALO_PROF
            const long long t_ld0 = clock64();
#endif
            if (leader) {
              uint32_t b_lo = b_lo0 + (uint32_t)(c * p.b_tap_mul) * b_sub16;
              uint32_t acc = c ? 1u : 0u;
              uint32_t a_row = a_stage;
              for (int r = 0; r < TR; ++r, a_row += step_r) {
                uint32_t a_lo = a_row;
                for (int s = 0; s < TS; ++s, a_lo += step_s, b_lo += b_tap16) {
                  if (!DBG(2)) halo_issue<NACC, KS>(tacc, acc_step, a_lo, j_step, b_lo, desc_hi, idesc, acc);
                  acc = 1u;
                }
              }
              PROF_WAIT(w_cm, umma_commit(&empty_a[stage]));
            }
#ifdef PDAC_HALO_PROF
            w_ld += clock64() - t_ld0;
#endif
          } else {
            // streamed weights: one ring stage holds the TS taps of a kernel row, so the warp-wide barrier wait and
            // the re-divergence around the issue (~150 cycles) are paid once per TS * NACC * KS MMAs
            uint32_t acc = c ? 1u : 0u;
            uint32_t a_row = a_stage;
            for (int r = 0; r < TR; ++r, a_row += step_r) {
              PROF_WAIT(w_fb, mbar_wait(&full_b[bstage], bphase, p.err));
              PROF_WAIT(w_fn, tc_fence_after());
              if (leader) {
                uint32_t a_lo = a_row;
                uint32_t b_lo = b_lo0 + (uint32_t)(bstage * TS) * b_sub16;
                for (int s = 0; s < TS; ++s, a_lo += step_s, b_lo += b_sub16) {
                  if (!DBG(2)) halo_issue<NACC, KS>(tacc, acc_step, a_lo, j_step, b_lo, desc_hi, idesc, acc);
                  acc = 1u;
                }
                PROF_WAIT(w_cm, umma_commit(&empty_b[bstage]));
              }
              acc = 1u;
              if (++bstage == sb) {
                bstage = 0;
                bphase ^= 1;
              }
            }
            PROF_WAIT(w_cm, if (leader) umma_commit(&empty_a[stage]));
          }
          if (++stage == sa) {
            stage = 0;
            phase ^= 1;
          }
        }
        PROF_WAIT(w_cm, if (leader) umma_commit(&tmem_full_bar[g]));
        PROF_WAIT(w_sw, __syncwarp());
      }
#ifdef PDAC_HALO_PROF
      if (lane == 0) {
        PROF_PUT(0, clock64() - t_mma0);
        PROF_PUT(1, w_te);
        PROF_PUT(2, w_fa);
        PROF_PUT(3, w_fb);
        PROF_PUT(4, (long long)lt);
        PROF_PUT(10, w_sw);
        PROF_PUT(11, w_fn);
        PROF_PUT(12, w_ld);
      }
      if (leader) PROF_PUT(9, w_cm);
#endif
    }
    __syncwarp();
  } else if (warp - 2 < 4 * p.n_wg) {
    // ------------------------------------------------------------------ epilogue (n_wg warpgroups, warps 2..)
    // Accumulators are dealt round-robin to the warpgroups (global accumulator index % n_wg). Per accumulator and per
    // 64-column pass: TMEM -> registers -> bias/addend/ReLU -> bf16 -> staging (invalid rows as zeros), then each
    // thread walks a fixed 8-channel column of the staging tile: 16-byte coalesced stores + BatchNorm partial sums
    // kept in registers for the whole kernel.
    const int wg = (warp - 2) >> 2;
    const int q = warp & 3;  // TMEM lane quarter this warp may read
    const int m = q * 32 + lane;
    const int e = (threadIdx.x - 64) & 127;
    const int sc = p.block_n < 64 ? p.block_n : 64;  // staging columns per pass
    const int spitch = sc * 2 + 16;
    const uint32_t stg = smem_u32(smem_stg) + (uint32_t)(wg * 128 * spitch);
    const uint32_t my_row = stg + (uint32_t)(m * spitch);
    const uint32_t bar_id = 1 + wg;
    const int nch = min(p.block_n, p.Cout - n0);
    const bool vec = (p.Cout & 7) == 0;
    constexpr int NPASS = DUAL ? 1 : 2;  // 64-column passes per accumulator (the dual variant has N <= 32)
    float acc[NPASS][16];
#pragma unroll
    for (int ps = 0; ps < NPASS; ++ps)
#pragma unroll
      for (int i = 0; i < 16; ++i) acc[ps][i] = 0.f;

    int lt = 0;
    Walk wk;
    walk_init(wk);
    PROF_DECL(w_tf);
    PROF_DECL(t_epi0);
#ifdef PDAC_HALO_PROF
    t_epi0 = clock64();
#endif
    const int nacc_mod = p.nacc % p.n_wg;
    int wg_off = 0;  // (lt * nacc) % n_wg, carried incrementally
    if (DIRECT) {
      // Thread m owns MMA row m, i.e. one pixel's contiguous output bytes (32 for N = 16 ... 256 for N = 128), and a
      // warp's 32 rows are 32 consecutive pixels of an image row -> every 32-byte store below writes a whole sector and
      // a warp writes one contiguous 1-2 KB run: no staging tile, no warpgroup barriers, a third of the instructions
      // of the staged epilogue (which ncu / the knock-outs showed to be what bounds the 16-channel layers: 2000 cycles
      // per accumulator and warpgroup). Three instantiations, so that none spills (the first version kept 32 statistics
      // registers next to the generic epilogue and spilled 16 values into the loop: ncu showed 75 % of the epilogue's
      // stall samples on the consumers of those local loads): MODE 3 generic without statistics (data gradients,
      // bias / addend / ReLU layers), MODE 4 folded BatchNorm (inference), MODE 5 the plain N = 16 convolution in front
      // of a training BatchNorm: lane pairs swap half of their packed row so that each lane adds 8 channels of 2 rows
      // (16 statistics registers instead of 32).
      float ssum[DSTATS ? 8 : 1], ssq[DSTATS ? 8 : 1];
#pragma unroll
      for (int i = 0; i < (DSTATS ? 8 : 1); ++i) ssum[i] = ssq[i] = 0.f;
      const int nchunk = DSTATS ? 1 : (nch + 15) >> 4;  // 16-column chunks of an accumulator that hold real channels
      if (sp0 < sp_step)
        for (int sp = sp0; sp < p.tiles_sp; sp += sp_step, ++lt, walk_next(wk)) {
          const int img = wk.img, w0 = wk.tw * p.TW, h0 = wk.th * p.TH;
          const int g = lt & 1;
          PROF_WAIT(w_tf, mbar_wait(&tmem_full_bar[g], (lt >> 1) & 1, p.err));
          tc_fence_after();
          const long long obase = (long long)img * p.osN + (long long)h0 * p.osH + (long long)w0 * p.osW;
          const long long abase = (long long)img * p.asN + (long long)h0 * p.asH + (long long)w0 * p.asW;
          int j = wg - wg_off;
          if (j < 0) j += p.n_wg;
          wg_off += nacc_mod;
          if (wg_off >= p.n_wg) wg_off -= p.n_wg;
          if (j >= p.nacc || DBG(1)) {
            tc_fence_before();
            mbar_arrive(&tmem_empty_bar[g]);
            if (DBG(1)) continue;
          }
          for (; j < p.nacc; j += p.n_wg) {
            const int o = j * 128 + m;
            const int th = (int)(((uint32_t)o * p.inv_pw) >> 20), tw = o - th * p.PW;
            const bool valid = (tw < p.TW) && (th < p.TH) && (w0 + tw < p.W) && (h0 + th < p.H);
            bf16* orow = p.out + obase + (long long)th * p.osH + (long long)tw * p.osW + n0;
            const bf16* arow = p.addend + abase + (long long)th * p.asH + (long long)tw * p.asW + n0;
            const uint32_t tacc = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)((g * p.nacc + j) * p.acc_cols);
            const bool last = j + p.n_wg >= p.nacc;
            if (DSTATS) {
              uint32_t v[16];
              tmem_ld16(tacc, v);
              tmem_ld_wait();
              if (p.stack) {  // add the other two taps of the kernel row, each from the lane stack_d[s] to the right
                if (p.stack_d[0]) {
#pragma unroll
                  for (int i = 0; i < 16; ++i) v[i] = __shfl_down_sync(0xffffffffu, v[i], p.stack_d[0]);
                }
#pragma unroll
                for (int sl = 1; sl < 3; ++sl) {
                  uint32_t u[16];
                  tmem_ld16(tacc + (uint32_t)(sl * 16), u);
                  tmem_ld_wait();
#pragma unroll
                  for (int i = 0; i < 16; ++i)
                    v[i] = __float_as_uint(__uint_as_float(v[i]) +
                                           __uint_as_float(__shfl_down_sync(0xffffffffu, u[i], p.stack_d[sl])));
                }
              }
              if (last) {
                tc_fence_before();
                mbar_arrive(&tmem_empty_bar[g]);
              }
              uint32_t r[8];
#pragma unroll
              for (int i = 0; i < 8; ++i)
                r[i] = valid ? h_pack(__uint_as_float(v[2 * i]), __uint_as_float(v[2 * i + 1])) : 0u;
              if (valid && !DBG(0)) stg256(orow, r);
              if (!DBG(3)) {
                // even lanes keep channels 0..7 of both rows of the pair, odd lanes channels 8..15
                const bool odd = lane & 1;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                  const uint32_t give = odd ? r[i] : r[4 + i], keep = odd ? r[4 + i] : r[i];
                  const uint32_t got = __shfl_xor_sync(0xffffffffu, give, 1);
                  const float a0 = h_bf16lo(keep), a1 = h_bf16hi(keep), b0 = h_bf16lo(got), b1 = h_bf16hi(got);
                  ssum[2 * i] += a0;
                  ssum[2 * i] += b0;
                  ssum[2 * i + 1] += a1;
                  ssum[2 * i + 1] += b1;
                  ssq[2 * i] = fmaf(a0, a0, ssq[2 * i]);
                  ssq[2 * i] = fmaf(b0, b0, ssq[2 * i]);
                  ssq[2 * i + 1] = fmaf(a1, a1, ssq[2 * i + 1]);
                  ssq[2 * i + 1] = fmaf(b1, b1, ssq[2 * i + 1]);
                }
              }
              continue;
            }
#pragma unroll 1
            for (int ck16 = 0; ck16 < nchunk; ++ck16) {
              const int cb = ck16 * 16;
              uint32_t v[16];
              tmem_ld16(tacc + (uint32_t)cb, v);
              tmem_ld_wait();
              if (p.stack) {  // see DSTATS above; slot s of a stacked accumulator starts at column s * N
                if (p.stack_d[0]) {
#pragma unroll
                  for (int i = 0; i < 16; ++i) v[i] = __shfl_down_sync(0xffffffffu, v[i], p.stack_d[0]);
                }
#pragma unroll
                for (int sl = 1; sl < 3; ++sl) {
                  uint32_t u[16];
                  tmem_ld16(tacc + (uint32_t)(sl * p.block_n + cb), u);
                  tmem_ld_wait();
#pragma unroll
                  for (int i = 0; i < 16; ++i)
                    v[i] = __float_as_uint(__uint_as_float(v[i]) +
                                           __uint_as_float(__shfl_down_sync(0xffffffffu, u[i], p.stack_d[sl])));
                }
              }
              if (last && ck16 == nchunk - 1) {  // last TMEM read of this warpgroup in this tile
                tc_fence_before();
                mbar_arrive(&tmem_empty_bar[g]);
              }
              float f[16];
#pragma unroll
              for (int i = 0; i < 16; ++i) f[i] = __uint_as_float(v[i]);
              if (AFF && p.bias && cb + 16 <= nch) {
                const float4* sp4 = reinterpret_cast<const float4*>(p.scale + n0 + cb);
                const float4* bp4 = reinterpret_cast<const float4*>(p.bias + n0 + cb);
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                  const float4 sv = __ldg(sp4 + i), bv = __ldg(bp4 + i);
                  f[4 * i] = fmaf(f[4 * i], sv.x, bv.x);
                  f[4 * i + 1] = fmaf(f[4 * i + 1], sv.y, bv.y);
                  f[4 * i + 2] = fmaf(f[4 * i + 2], sv.z, bv.z);
                  f[4 * i + 3] = fmaf(f[4 * i + 3], sv.w, bv.w);
                }
              } else {
                if (p.scale) {
#pragma unroll
                  for (int i = 0; i < 16; ++i)
                    if (cb + i < nch) f[i] *= __ldg(p.scale + n0 + cb + i);
                }
                if (p.bias) {
#pragma unroll
                  for (int i = 0; i < 16; ++i)
                    if (cb + i < nch) f[i] += __ldg(p.bias + n0 + cb + i);
                }
              }
              if (p.addend && valid) {
#pragma unroll
                for (int hh = 0; hh < 2; ++hh) {
                  if (cb + hh * 8 < nch) {
                    const uint4 u = __ldg(reinterpret_cast<const uint4*>(arow + cb + hh * 8));
                    const uint32_t uu[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                      f[hh * 8 + 2 * i] += h_bf16lo(uu[i]);
                      f[hh * 8 + 2 * i + 1] += h_bf16hi(uu[i]);
                    }
                  }
                }
              }
              if (p.relu) {
#pragma unroll
                for (int i = 0; i < 16; ++i) f[i] = fmaxf(f[i], 0.f);
              }
              uint32_t r[8];
#pragma unroll
              for (int i = 0; i < 8; ++i) r[i] = h_pack(f[2 * i], f[2 * i + 1]);
              if (valid && !DBG(0)) {
                if (cb + 16 <= nch) {
                  stg256(orow + cb, r);
                  if (p.up2) {
                    stg256(orow + cb + p.up_dw, r);
                    stg256(orow + cb + p.up_dh, r);
                    stg256(orow + cb + p.up_dh + p.up_dw, r);
                  }
                } else if (cb + 8 <= nch) {
                  const uint4 r4 = make_uint4(r[0], r[1], r[2], r[3]);
                  *reinterpret_cast<uint4*>(orow + cb) = r4;
                  if (p.up2) {
                    *reinterpret_cast<uint4*>(orow + cb + p.up_dw) = r4;
                    *reinterpret_cast<uint4*>(orow + cb + p.up_dh) = r4;
                    *reinterpret_cast<uint4*>(orow + cb + p.up_dh + p.up_dw) = r4;
                  }
                }
              }
            }
          }
        }
      if (DSTATS && p.stats) {
        // fixed order: lanes of equal parity by xor-shuffles, then the warp's slot (added over the warps after the CTA
        // barrier); lane 0 ends with channels 0..7, lane 1 with channels 8..15
        float* wslot = reinterpret_cast<float*>(smem_stg + (size_t)wg * 128 * spitch) + q * 256;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          float a = ssum[i], b = ssq[i];
#pragma unroll
          for (int o = 16; o >= 2; o >>= 1) {
            a += __shfl_xor_sync(0xffffffffu, a, o);
            b += __shfl_xor_sync(0xffffffffu, b, o);
          }
          const int chn = (lane & 1) * 8 + i;
          if (lane < 2 && chn < nch) {
            wslot[chn] = a;
            wslot[128 + chn] = b;
          }
        }
      }
    }
    if (!DIRECT && sp0 < sp_step)
      for (int sp = sp0; sp < p.tiles_sp; sp += sp_step, ++lt, walk_next(wk)) {
        const int img = wk.img, w0 = wk.tw * p.TW, h0 = wk.th * p.TH;
        const int g = lt & 1;
        PROF_WAIT(w_tf, mbar_wait(&tmem_full_bar[g], (lt >> 1) & 1, p.err));
        tc_fence_after();
        const long long obase = (long long)img * p.osN + (long long)h0 * p.osH + (long long)w0 * p.osW;
        const long long abase = (long long)img * p.asN + (long long)h0 * p.asH + (long long)w0 * p.asW;
        int j = wg - wg_off;
        if (j < 0) j += p.n_wg;
        wg_off += nacc_mod;
        if (wg_off >= p.n_wg) wg_off -= p.n_wg;
        if (j >= p.nacc || DBG(1)) {  // nothing of this tile is ours: the TMEM stage is free as far as we are concerned
          tc_fence_before();
          mbar_arrive(&tmem_empty_bar[g]);
          if (DBG(1)) continue;
        }
        for (; j < p.nacc; j += p.n_wg) {
          const int o = j * 128 + m;
          const int th = (int)(((uint32_t)o * p.inv_pw) >> 20), tw = o - th * p.PW;
          const bool valid = (tw < p.TW) && (th < p.TH) && (w0 + tw < p.W) && (h0 + th < p.H);
          const long long ooff = obase + (long long)th * p.osH + (long long)tw * p.osW;
          const bf16* arow = p.addend + abase + (long long)th * p.asH + (long long)tw * p.asW + n0;
          asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");  // previous staging fully consumed
          s_rowoff[wg][m] = valid ? ooff : -1;
          if (POOL && m < 32) {  // the 32 pooled pixels of this accumulator: (128 / PW) / 2 row pairs x PW / 2 column pairs
            const int hw = p.pw_log2 - 1;
            const int pr = m >> hw, pc = m & ((1 << hw) - 1);
            const int th2 = j * (128 >> p.pw_log2) + 2 * pr, tw2 = 2 * pc;
            const bool ok = (tw2 < p.TW) && (th2 < p.TH) && (w0 + tw2 < p.W) && (h0 + th2 < p.H);
            s_pooloff[wg][m] = ok ? (long long)img * p.psN + (long long)((h0 + th2) >> 1) * p.psH +
                                        (long long)((w0 >> 1) + pc) * p.psW
                                  : -1;
          }
          const uint32_t tacc =
              tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)((g * p.nacc + j) * p.acc_cols);
#pragma unroll
          for (int pass = 0; pass < NPASS; ++pass) {
            const int c0 = pass * 64;
            if (c0 < p.block_n) {
              const int cols = min(64, p.block_n - c0);
              if (pass) asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
              for (int c = 0; c < cols; c += 16) {
                uint32_t v[16];
                tmem_ld16(tacc + (uint32_t)(c0 + c), v);
                tmem_ld_wait();
                if (SSTACK) {  // N <= 32: add the other two taps of the kernel row from the lanes to the right
                  if (p.stack_d[0]) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) v[i] = __shfl_down_sync(0xffffffffu, v[i], p.stack_d[0]);
                  }
#pragma unroll
                  for (int sl = 1; sl < 3; ++sl) {
                    uint32_t u[16];
                    tmem_ld16(tacc + (uint32_t)(sl * p.block_n + c), u);
                    tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 16; ++i)
                      v[i] = __float_as_uint(__uint_as_float(v[i]) +
                                             __uint_as_float(__shfl_down_sync(0xffffffffu, u[i], p.stack_d[sl])));
                  }
                }
                float f[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) f[i] = __uint_as_float(v[i]);
                if (AFF && p.bias && c0 + c + 16 <= nch) {
                  // folded BatchNorm (inference): 8 vector loads instead of 32 scalar ones per 16 outputs
                  const float4* sp = reinterpret_cast<const float4*>(p.scale + n0 + c0 + c);
                  const float4* bp = reinterpret_cast<const float4*>(p.bias + n0 + c0 + c);
#pragma unroll
                  for (int i = 0; i < 4; ++i) {
                    const float4 sv = __ldg(sp + i), bv = __ldg(bp + i);
                    f[4 * i] = fmaf(f[4 * i], sv.x, bv.x);
                    f[4 * i + 1] = fmaf(f[4 * i + 1], sv.y, bv.y);
                    f[4 * i + 2] = fmaf(f[4 * i + 2], sv.z, bv.z);
                    f[4 * i + 3] = fmaf(f[4 * i + 3], sv.w, bv.w);
                  }
                } else {
                  if (p.scale) {
#pragma unroll
                    for (int i = 0; i < 16; ++i)
                      if (c0 + c + i < nch) f[i] *= __ldg(p.scale + n0 + c0 + c + i);
                  }
                  if (p.bias) {
#pragma unroll
                    for (int i = 0; i < 16; ++i)
                      if (c0 + c + i < nch) f[i] += __ldg(p.bias + n0 + c0 + c + i);
                  }
                }
                if (p.addend && valid) {
                  const bf16* ap = arow + c0 + c;
                  if (vec) {
#pragma unroll
                    for (int hh = 0; hh < 2; ++hh) {
                      if (c0 + c + hh * 8 < nch) {
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
                      if (c0 + c + i < nch) f[i] += __bfloat162float(ap[i]);
                  }
                }
                if (p.relu) {
#pragma unroll
                  for (int i = 0; i < 16; ++i) f[i] = fmaxf(f[i], 0.f);
                }
                if (!valid) {
#pragma unroll
                  for (int i = 0; i < 16; ++i) f[i] = 0.f;
                }
                sts128(my_row + (uint32_t)(c * 2), h_pack(f[0], f[1]), h_pack(f[2], f[3]), h_pack(f[4], f[5]),
                       h_pack(f[6], f[7]));
                sts128(my_row + (uint32_t)(c * 2 + 16), h_pack(f[8], f[9]), h_pack(f[10], f[11]),
                       h_pack(f[12], f[13]), h_pack(f[14], f[15]));
              }
              if (c0 + 64 >= p.block_n && j + p.n_wg >= p.nacc) {
                // last TMEM read of this warpgroup in this tile
                tc_fence_before();
                mbar_arrive(&tmem_empty_bar[g]);
              }
              asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
              if (vec) {
                // thread -> fixed 8-channel column `ch`, rows row0, row0 + step, ...
                const int cpr_log2 = cols == 64 ? 3 : (cols == 32 ? 2 : 1);
                const int ch = e & ((1 << cpr_log2) - 1);
                const int step = 128 >> cpr_log2;
                const bool ch_ok = c0 + ch * 8 < nch;
                bf16* gcol = p.out + n0 + c0 + ch * 8;
                uint32_t saddr = stg + (uint32_t)((e >> cpr_log2) * spitch + ch * 16);
                const uint32_t sstep = (uint32_t)(step * spitch);
                int row = e >> cpr_log2;
                if (POOL && n0 + c0 + ch * 8 < p.pool_c) {
                  // this 8-channel column is pooled: rows < 32 index pooled pixels; (a + b) + (c + d) in fp32 of the
                  // bf16-rounded values, the order of up2_bwd_kernel (pool.cu) -> bit-identical to the unfused pair
                  const int hw = p.pw_log2 - 1;
                  const uint32_t colb = stg + (uint32_t)(ch * 16);
                  for (; row < 32; row += step) {
                    const int pr = row >> hw, pc = row & ((1 << hw) - 1);
                    const uint32_t a00 = colb + (uint32_t)((((2 * pr) << p.pw_log2) + 2 * pc) * spitch);
                    const uint4 va = lds128(a00), vb = lds128(a00 + (uint32_t)spitch);
                    const uint4 vc = lds128(a00 + (uint32_t)(spitch << p.pw_log2));
                    const uint4 vd = lds128(a00 + (uint32_t)(spitch << p.pw_log2) + (uint32_t)spitch);
                    const uint32_t ua[4] = {va.x, va.y, va.z, va.w}, ub[4] = {vb.x, vb.y, vb.z, vb.w};
                    const uint32_t uc[4] = {vc.x, vc.y, vc.z, vc.w}, ud[4] = {vd.x, vd.y, vd.z, vd.w};
                    uint32_t r[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                      r[i] = h_pack((h_bf16lo(ua[i]) + h_bf16lo(ub[i])) + (h_bf16lo(uc[i]) + h_bf16lo(ud[i])),
                                    (h_bf16hi(ua[i]) + h_bf16hi(ub[i])) + (h_bf16hi(uc[i]) + h_bf16hi(ud[i])));
                    const long long off = s_pooloff[wg][row];
                    if (off >= 0 && ch_ok && !DBG(0))
                      *reinterpret_cast<uint4*>(p.pout + n0 + c0 + ch * 8 + off) = make_uint4(r[0], r[1], r[2], r[3]);
                  }
                } else {
#pragma unroll 2
                for (int it = 0; it < (1 << cpr_log2); ++it, row += step, saddr += sstep) {
                  const long long off = s_rowoff[wg][row];
                  const uint4 val = lds128(saddr);
                  if (off >= 0 && ch_ok && !DBG(0)) *reinterpret_cast<uint4*>(gcol + off) = val;
                  if (p.stats && !DBG(3)) {
                    const uint32_t uu[4] = {val.x, val.y, val.z, val.w};
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                      const float x0 = h_bf16lo(uu[i]), x1 = h_bf16hi(uu[i]);
                      acc[pass][2 * i] += x0;
                      acc[pass][2 * i + 1] += x1;
                      acc[pass][8 + 2 * i] = fmaf(x0, x0, acc[pass][8 + 2 * i]);
                      acc[pass][8 + 2 * i + 1] = fmaf(x1, x1, acc[pass][8 + 2 * i + 1]);
                    }
                  }
                }
                }
              } else {
                // channel count not a multiple of 8 (1-class logit heads): scalar stores, no statistics
                const int nv = min(cols, nch - c0);
                for (int id = e; id < 128 * nv; id += 128) {
                  const int row = id / nv, chn = id - row * nv;
                  const long long off = s_rowoff[wg][row];
                  if (off < 0) continue;
                  p.out[off + n0 + c0 + chn] = __ushort_as_bfloat16(lds16(stg + (uint32_t)(row * spitch + chn * 2)));
                }
              }
            }
          }
        }
      }
#ifdef PDAC_HALO_PROF
    if (threadIdx.x == 64) {
      PROF_PUT(5, clock64() - t_epi0);
      PROF_PUT(6, w_tf);
    }
#endif
    // BatchNorm partial sums, in a FIXED order (bit-reproducible statistics; the reference trains with
    // cudnn.deterministic, train_main.py:40): lanes sharing a channel column -> one lane per column -> this warp's
    // slot in the (now idle) staging area -> after the final CTA barrier one thread per channel adds the warps'
    // slots in index order and stores the CTA's row of the [rows][2][Cout] partial table (no atomics anywhere).
    if (!DIRECT && p.stats && vec) {
      asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");  // the last tile's staging has been consumed
      float* wslot = reinterpret_cast<float*>(smem_stg + (size_t)wg * 128 * spitch) + q * 256;
#pragma unroll
      for (int pass = 0; pass < NPASS; ++pass) {
        const int c0 = pass * 64;
        if (c0 < p.block_n) {
          const int cols = min(64, p.block_n - c0);
          const int cpr = cols >> 3;
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            float v = acc[pass][i];
            for (int o = 16; o >= cpr; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            acc[pass][i] = v;
          }
          if (lane < cpr) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const int chn = c0 + lane * 8 + i;
              if (chn < nch) {
                wslot[chn] = acc[pass][i];
                wslot[128 + chn] = acc[pass][8 + i];
              }
            }
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (p.stats && (int)threadIdx.x < 2 * p.block_n) {
    const int which = threadIdx.x >= p.block_n, chn = threadIdx.x - which * p.block_n;
    if (chn < min(p.block_n, p.Cout - n0)) {
      const int sc = p.block_n < 64 ? p.block_n : 64;
      const int spitch = sc * 2 + 16;
      float v = 0.f;
      for (int w = 0; w < p.n_wg; ++w)
        for (int qq = 0; qq < 4; ++qq)
          v += reinterpret_cast<const float*>(smem_stg + (size_t)w * 128 * spitch)[qq * 256 + which * 128 + chn];
      p.stats[((size_t)sp0 * 2 + which) * p.Cout + n0 + chn] = v;  // row sp0 belongs to this CTA column alone
    }
  }
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, p.tmem_cols);
  }
#ifdef PDAC_HALO_PROF
  if (threadIdx.x == 0) {
    long long g_k1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g_k1));
    PROF_PUT(13, g_k0);
    PROF_PUT(14, g_k1);
    PROF_PUT(15, clock64() - t_k0);
  }
#endif
}


typedef void (*HaloKernel)(const HaloMaps, const HaloParams);
// [nacc - 1][variant]: variant = {ck 16, 32, 64} for one CTA per SM, then {ck 16, 32} for two CTAs per SM
#define HALO_ROW(n, mode)                                                                                     \
  {                                                                                                            \
    conv_halo_kernel<n, 1, 0, mode>, conv_halo_kernel<n, 2, 0, mode>, conv_halo_kernel<n, 4, 0, mode>,         \
        conv_halo_kernel<n, 1, 1, mode>, conv_halo_kernel<n, 2, 1, mode>                                       \
  }
// the statistics variant of the direct epilogue (mode 5) exists for the two-CTAs-per-SM kernels only
#define HALO_ROW_D(n, mode) \
  { nullptr, nullptr, nullptr, conv_halo_kernel<n, 1, 1, mode>, conv_halo_kernel<n, 2, 1, mode> }
#define HALO_MODE(mode) {HALO_ROW(1, mode), HALO_ROW(2, mode), HALO_ROW(3, mode), HALO_ROW(4, mode)}
#define HALO_MODE_D(mode) {HALO_ROW_D(1, mode), HALO_ROW_D(2, mode), HALO_ROW_D(3, mode), HALO_ROW_D(4, mode)}
HaloKernel halo_kernel_k0(int mode, int ki, int kj);  // modes 0 (training, staged) and 1 (folded BatchNorm, staged)
HaloKernel halo_kernel_k1(int mode, int ki, int kj);  // modes 2 (pooled data gradient) and 7 (staged + tap stacking)
HaloKernel halo_kernel_k2(int mode, int ki, int kj);  // modes 3 and 4 (direct epilogue, generic / folded BatchNorm)
HaloKernel halo_kernel_k3(int mode, int ki, int kj);  // mode 5 (direct epilogue with BatchNorm sums)

}  // namespace pdac
