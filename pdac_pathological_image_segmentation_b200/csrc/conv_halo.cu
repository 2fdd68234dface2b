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
                if (cb + 16 <= nch)
                  stg256(orow + cb, r);
                else if (cb + 8 <= nch)
                  *reinterpret_cast<uint4*>(orow + cb) = make_uint4(r[0], r[1], r[2], r[3]);
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

// ------------------------------------------------------------------------------------------------ host side

static const int kSmemBudget = 221 * 1024;  // dynamic shared memory we allow ourselves (227 KB max per CTA incl. static)

struct HaloPlan {
  int TW, TH, nacc, sa, sb, b_resident, n_wg;
  double cost;
};

// Picks the tile. Cost model (cycles per image, per SM share): MMA issue = tiles * nacc * taps * K16-steps * max(N/2, 10)
// against bytes moved (halo in + tile out) at ~40 B/cycle/SM; the larger of the two plus a per-tile overhead.
static bool plan_halo(int H, int W, int N, int n_taps, int ts, int n_chunks, int ck, int eh, int ew, int has_addend,
                      int dual, int pool, int no_staging, int stack, HaloPlan* best) {
  const int pitch = ck * 2;
  // dual: two CTAs per SM, each with half of the shared memory (228 KB per SM, 1 KB reserved + ~6.5 KB static per CTA)
  // and half of the tensor memory
  const long long kSmemBudget = dual ? 104 * 1024 : pdac::kSmemBudget;
  const int tmem_half = dual ? 128 : 256;
  const int ncols = stack ? 3 * N : N;  // accumulator columns (tap stacking: the three taps of a kernel row side by side)
  const int max_nacc = tmem_half / ncols < 4 ? tmem_half / ncols : 4;
  if (max_nacc < 1) return false;
  const long long w_bytes = (long long)n_taps * n_chunks * N * pitch;
  // staging bytes per epilogue warpgroup (none with the direct register -> global epilogue)
  const int stg_wg = no_staging ? 0 : 128 * ((N < 64 ? N : 64) * 2 + 16);
  best->cost = 1e30;
  // tuning overrides (debug): PDAC_HALO_TW / PDAC_HALO_NACC / PDAC_HALO_SA / PDAC_HALO_WG
  static int f_tw = -1, f_nacc = -1, f_sa = -1, f_wg = -1;
  if (f_tw < 0) {
    const char* e;
    f_tw = (e = getenv("PDAC_HALO_TW")) ? atoi(e) : 0;
    f_nacc = (e = getenv("PDAC_HALO_NACC")) ? atoi(e) : 0;
    f_sa = (e = getenv("PDAC_HALO_SA")) ? atoi(e) : 0;
    f_wg = (e = getenv("PDAC_HALO_WG")) ? atoi(e) : 0;
  }
  // pool: the halo pitch is a power of two that divides 128 into an even number of pixel rows per accumulator (every
  // 2x2 block then lies inside one accumulator's staging tile); the tile is nacc * 128 / PW rows high
  // stack: an image row of the halo raster must not leave its warp (PW = 16 or 32)
  const int tws[5] = {pool || stack ? 16 - ew : 16, pool || stack ? 32 - ew : 32, stack ? 0 : (pool ? 64 - ew : 64),
                      pool || stack ? 0 : 128, pool || stack ? 0 : W};
  for (int ti = 0; ti < 5; ++ti) {
    int TW = tws[ti];
    if (TW <= 0) continue;
    if (TW > W && !pool && !stack) TW = W;
    if (TW + ew > 256) continue;
    if (f_tw && TW != f_tw && !pool && !stack) continue;
    const int PW = TW + ew;
    for (int nacc = 1; nacc <= max_nacc; ++nacc) {
      int TH;
      if (pool) {
        TH = nacc * (128 / PW);
        if (nacc > 1 && TH - 128 / PW >= H) continue;  // fewer accumulators already cover the image height
        if (TH + eh > 256) continue;
      } else {
        TH = (128 * nacc - TW) / PW + 1;
        if (128 * nacc < TW) continue;
        if (TH > H) TH = H;
        if (TH + eh > 256) TH = 256 - eh;
        const int nacc_eff = ((TH - 1) * PW + TW + 127) / 128;
        if (nacc_eff != nacc) continue;  // the same tile is reachable with fewer accumulators
      }
      if (f_nacc && nacc != f_nacc) continue;
      const int R = TH + eh;
      const int a_buf = ((nacc * 128 + eh * PW + ew) * pitch + 1023) & ~1023;
      for (int res = 1; res >= 0; --res) {
        if (res && w_bytes > 112 * 1024) continue;
        int sa = 0, sb = 0, n_wg = dual ? 2 : (f_wg ? f_wg : 4);
        for (; n_wg >= 2; n_wg -= 2) {  // four epilogue warpgroups when their staging fits, else two
          const long long fixed = (long long)n_wg * stg_wg + 2048;
          if (res) {
            sa = (int)((kSmemBudget - fixed - w_bytes) / a_buf);
            sb = 0;
          } else {
            const long long row_bytes = (long long)ts * N * pitch;  // one ring stage = a kernel row of taps
            sb = 2;
            sa = (int)((kSmemBudget - fixed - sb * row_bytes) / a_buf);
            if (sa > 3) {  // spend the rest on a deeper weight ring
              sa = 3;
              sb = (int)((kSmemBudget - fixed - 3LL * a_buf) / row_bytes);
              if (sb > kHaloMaxSB) sb = kHaloMaxSB;
            }
          }
          if (sa >= 2) break;
        }
        if (sa > kHaloMaxSA) sa = kHaloMaxSA;
        if (f_sa && sa > f_sa) sa = f_sa;
        if (sa < 2) continue;
        const int tiles = ceil_div(W, TW) * ceil_div(H, TH);
        // measured: ~32 + N/4 cycles per M=128 K=16 MMA (the shared-memory reads of its operands)
        const double mma_cyc = stack ? (double)tiles * nacc * (n_taps / 3) * n_chunks * (ck / 16) * (32 + 3 * N / 4)
                                     : (double)tiles * nacc * n_taps * n_chunks * (ck / 16) * (32 + N / 4 > 40 ? 32 + N / 4 : 40);
        double bytes = (double)tiles * R * PW * n_chunks * pitch + (double)H * W * N * 2 * (has_addend ? 2 : 1);
        if (!res) bytes += (double)tiles * w_bytes;
        const double mem_cyc = bytes / 40.0;
        double cost = (mma_cyc > mem_cyc ? mma_cyc : mem_cyc) + 0.5 * (mma_cyc < mem_cyc ? mma_cyc : mem_cyc) +
                      300.0 * tiles;
        if (!res) cost += 150.0 * tiles * (n_taps / ts) * n_chunks;  // per weight stage: barrier wait + commit in the issue path
        if (cost < best->cost) {
          best->cost = cost;
          best->TW = TW;
          best->TH = TH;
          best->nacc = nacc;
          best->sa = sa;
          best->sb = sb;
          best->b_resident = res;
          best->n_wg = n_wg;
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
  if ((cout & 7) == 0) {
    // the vectorised epilogue walks 8-channel columns of 16/32/64-wide passes
    if (block_n != 16 && block_n != 32 && block_n != 64 && block_n != 80 && block_n != 96 && block_n != 128)
      return PDAC_OK;
  } else if (stats) {
    return PDAC_OK;
  }
  const int chunks0 = views[0].C / ck;
  const int n_chunks = chunks0 + (tc->n_cat > 1 ? views[1].C / ck : 0);
  const int ktot = n_chunks * ck;
  int tap_s = 1;  // taps per kernel row (the tap list is an (r, s) grid, verified below)
  while (tap_s < tc->n_taps && tc->taps[tap_s].dh == tc->taps[0].dh) ++tap_s;
  if (tc->n_taps % tap_s) return PDAC_OK;
  HaloPlan pl = {0, 0, 0, 0, 0, 0, 0, 0.0};
  // Low channel counts (K chunk and N <= 32): two CTAs per SM, see kHaloThreadsDual. PDAC_HALO_DUAL=0 disables.
  static int dual_ok = -1;
  if (dual_ok < 0) {
    const char* e = getenv("PDAC_HALO_DUAL");
    dual_ok = (e && e[0] == '0') ? 0 : 1;
  }
  // POOL mode (pdac_conv2d_dgrad_pool2): the leading pool->C output channels are stored 2x2-summed at half resolution
  const PdacView* pool = epi_pool();
  if (pool) {
    if ((cout & 7) || (pool->C & 7) || pool->C <= 0 || pool->C > cout || (out->H & 1) || (out->W & 1) || (ew & 1) ||
        pool->N != out->N || 2 * pool->H != out->H || 2 * pool->W != out->W || addend || stats || bias || epi_scale() ||
        tc->relu)
      return PDAC_OK;
  }
  // Direct register -> global epilogue (no staging tile): 32-byte aligned output pixels; without BatchNorm sums for any
  // N, with them for the plain N = 16 convolution of the two-CTAs-per-SM kernels (the per-thread sums fit the register
  // budget there only). PDAC_HALO_DIRECT=0 keeps the staged epilogue.
  static int direct_ok = -1;
  if (direct_ok < 0) {
    const char* e = getenv("PDAC_HALO_DIRECT");
    direct_ok = (e && e[0] == '0') ? 0 : 1;
  }
  const bool aligned32 = ((uintptr_t)out->ptr & 31) == 0 && out->sW % 16 == 0 && out->sH % 16 == 0 && out->sN % 16 == 0;
  const bool plain = !bias && !addend && !epi_scale() && !tc->relu;
  const bool direct_ns = direct_ok && !pool && (cout & 7) == 0 && aligned32 && !stats;
  const bool direct_st = direct_ok && !pool && cout == 16 && plain && aligned32 && stats;
  int dual = dual_ok && ck <= 32 && block_n <= 32;
  if (dual && !plan_halo(out->H, out->W, block_n, tc->n_taps, tap_s, n_chunks, ck, eh, ew, addend != nullptr, 1,
                         pool != nullptr, direct_ns, 0, &pl))
    dual = 0;
  if (!dual && !plan_halo(out->H, out->W, block_n, tc->n_taps, tap_s, n_chunks, ck, eh, ew, addend != nullptr, 0,
                          pool != nullptr, direct_ns, 0, &pl))
    return PDAC_OK;
  const bool direct = direct_ns || (direct_st && dual);
  // Tap stacking (see HaloParams::stack): direct epilogue, N = 16, a single K chunk, 3 taps per kernel row, resident
  // weights; taken when the cost model prefers its plan. PDAC_HALO_STACK=0 disables.
  static int stack_ok = -1;
  if (stack_ok < 0) {
    const char* e = getenv("PDAC_HALO_STACK");
    stack_ok = (e && e[0] == '0') ? 0 : 1;
  }
  bool stack = false;
  // ck >= 32 only: with a 16-channel source the layer is HBM-bound already and the smaller stacked tile costs more than
  // the MMAs it saves (measured: 16 -> 16 @512^2 0.213 -> 0.283 ms, 32 -> 16 0.392 -> 0.330 ms). Direct epilogues take it
  // as they are; a layer with BatchNorm sums that is not the N = 16 direct case gets the staged variant (mode 7).
  // (two CTAs per SM with N = 32, i.e. the 32 -> 32 layers: TMEM holds one stacked accumulator per tile stage only, and
  // the 128-row tile made both the forward, 0.163 -> 0.194 ms, and the data gradient, 0.119 -> 0.160 ms, slower)
  static int stack_maxn = -1;  // experiment switch: PDAC_HALO_STACK_MAXN (default 32)
  if (stack_maxn < 0) {
    const char* e = getenv("PDAC_HALO_STACK_MAXN");
    stack_maxn = e ? atoi(e) : 32;
  }
  if (stack_ok && !pool && (cout & 7) == 0 && block_n <= stack_maxn && 3 * block_n <= 256 && ck >= 32 && tap_s == 3 &&
      ew <= 14 && pl.b_resident &&
      (direct || !epi_scale()) && !(dual && block_n == 32)) {
    HaloPlan ps = {0, 0, 0, 0, 0, 0, 0, 0.0};
    if (plan_halo(out->H, out->W, block_n, tc->n_taps, tap_s, n_chunks, ck, eh, ew, addend != nullptr, dual, 0, direct_ns,
                  1, &ps) &&
        ps.b_resident && ps.cost < pl.cost) {
      pl = ps;
      stack = true;
    }
  }

  {
    static int dbg = -1;
    if (dbg < 0) {
      const char* e = getenv("PDAC_HALO_DEBUG");
      dbg = (e && e[0] == '1') ? 1 : 0;
    }
    if (dbg)
      fprintf(stderr, "[conv_halo] %dx%dx%d K=%dx%d N=%d taps=%d: TW=%d TH=%d nacc=%d sa=%d sb=%d resident=%d n_wg=%d dual=%d direct=%d stack=%d cost=%.0f\n",
              out->N, out->H, out->W, n_chunks, ck, block_n, tc->n_taps, pl.TW, pl.TH, pl.nacc, pl.sa, pl.sb,
              pl.b_resident, pl.n_wg, dual, (int)direct, (int)stack, pl.cost);
  }
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
  p.n_wg = pl.n_wg;
  p.w_warp = dual ? 2 + 4 * pl.n_wg : 18;
  p.inv_pw = (uint32_t)(((1u << 20) + (uint32_t)p.PW - 1) / (uint32_t)p.PW);
  p.dh_min = dh_min;
  p.dw_min = dw_min;
  {
    // the issue loop walks the taps as an affine (r, s) grid; anything else goes to the generic kernel
    int S = 1;
    while (S < tc->n_taps && tc->taps[S].dh == tc->taps[0].dh) ++S;
    if (tc->n_taps % S) return PDAC_OK;
    const int Rr = tc->n_taps / S;
    auto shift = [&](int t) { return (tc->taps[t].dh - dh_min) * p.PW + (tc->taps[t].dw - dw_min); };
    p.tap_R = Rr;
    p.tap_S = S;
    p.tap_base = shift(0);
    p.tap_step_s = S > 1 ? shift(1) - shift(0) : 0;
    p.tap_step_r = Rr > 1 ? shift(S) - shift(0) : 0;
    for (int t = 0; t < tc->n_taps; ++t)
      if (shift(t) != p.tap_base + (t / S) * p.tap_step_r + (t % S) * p.tap_step_s) return PDAC_OK;
  }
  p.acc_cols = block_n;
  p.issue_S = p.tap_S;
  p.b_tap_mul = 1;
  if (stack) {
    const int lo = p.tap_step_s < 0 ? 2 * p.tap_step_s : 0;  // the leftmost tap of a kernel row
    for (int sidx = 0; sidx < 3; ++sidx) p.stack_d[sidx] = sidx * p.tap_step_s - lo;
    PDAC_REQUIRE(p.stack_d[0] >= 0 && p.stack_d[2] >= 0 && p.stack_d[0] + p.stack_d[2] <= ew && (p.PW == 16 || p.PW == 32),
                 "internal: stacked halo plan (PW=%d step=%d)", p.PW, p.tap_step_s);
    p.tap_base += lo;
    p.stack = 1;
    p.acc_cols = 3 * block_n;
    p.issue_S = 1;
    p.b_tap_mul = 3;
  }
  p.pitch = ck * 2;
  p.layout_type = swizzle_layout_type(p.pitch);
  p.idesc = make_idesc_bf16(128, p.acc_cols, 0, 0);
  p.block_n = block_n;
  p.tiles_c = ceil_div(cout, block_n);
  p.wrows = tc->wrows;
  p.tmem_cols = pow2_ceil(2 * pl.nacc * p.acc_cols < 32 ? 32 : 2 * pl.nacc * p.acc_cols);
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
  p.scale = epi_scale();
  if (p.scale && bias)
    PDAC_REQUIRE(((uintptr_t)p.scale & 15) == 0 && ((uintptr_t)bias & 15) == 0, "affine scale / shift must be 16-byte aligned");
  p.stats = stats;
  p.relu = tc->relu;
  p.err = device_err_flag();
  if (pool) {
    PDAC_REQUIRE((128 % p.PW) == 0 && ((128 / p.PW) & 1) == 0 && p.TH == p.nacc * (128 / p.PW),
                 "internal: pooled halo plan (PW=%d TH=%d nacc=%d)", p.PW, p.TH, p.nacc);
    PDAC_REQUIRE(((uintptr_t)pool->ptr & 15) == 0 && pool->sW % 8 == 0 && pool->sH % 8 == 0 && pool->sN % 8 == 0,
                 "pooled gradient view must be 16-byte aligned");
    p.pool_c = pool->C;
    p.pw_log2 = ilog2(p.PW);
    p.pout = (bf16*)pool->ptr;
    p.psN = pool->sN;
    p.psH = pool->sH;
    p.psW = pool->sW;
  }
  for (int s = 0; s < tc->n_cat; ++s) PDAC_TRY(encode_view_map(&maps.a[s], &views[s], ck, p.PW, p.R, 1));
  if (tc->n_cat == 1) maps.a[1] = maps.a[0];
  {
    uint64_t dims[2] = {(uint64_t)ktot, (uint64_t)tc->n_taps * tc->wrows};
    uint64_t strides[1] = {(uint64_t)ktot * 2};
    uint32_t box[2] = {(uint32_t)ck, (uint32_t)block_n};
    PDAC_TRY(encode_tiled_bf16(&maps.w, w_packed, 2, dims, strides, box, p.pitch));
  }
  const size_t b_bytes = (size_t)(p.b_resident ? p.n_taps * p.n_chunks : p.sb * p.tap_S) * p.b_sub;
  const size_t smem = (size_t)p.sa * p.a_buf + b_bytes +
                      (direct_ns ? 0 : (size_t)p.n_wg * 128 * ((block_n < 64 ? block_n : 64) * 2 + 16)) + 1024;
  PDAC_REQUIRE(smem <= (size_t)(dual ? 105 : 221) * 1024, "internal: halo conv shared memory plan (%zu bytes)", smem);
  int per_n = (dual ? 2 : 1) * num_sms() / p.tiles_c;
  if (per_n < 1) per_n = 1;
  if (per_n > p.tiles_sp) per_n = p.tiles_sp;
  PDAC_REQUIRE(!stats || per_n <= PDAC_STAT_ROWS, "internal: %d statistics rows exceed PDAC_STAT_ROWS", per_n);
  typedef void (*HaloKernel)(const HaloMaps, const HaloParams);
#define HALO_ROW(n, aff)                                                                                       \
  {                                                                                                             \
    conv_halo_kernel<n, 1, 0, aff>, conv_halo_kernel<n, 2, 0, aff>, conv_halo_kernel<n, 4, 0, aff>,             \
        conv_halo_kernel<n, 1, 1, aff>, conv_halo_kernel<n, 2, 1, aff>                                          \
  }
  // the statistics variant of the direct epilogue (mode 5) exists for the two-CTAs-per-SM kernels only
#define HALO_ROW_D(n, mode) \
  { nullptr, nullptr, nullptr, conv_halo_kernel<n, 1, 1, mode>, conv_halo_kernel<n, 2, 1, mode> }
  static const HaloKernel table[8][4][5] = {{HALO_ROW(1, 0), HALO_ROW(2, 0), HALO_ROW(3, 0), HALO_ROW(4, 0)},
                                            {HALO_ROW(1, 1), HALO_ROW(2, 1), HALO_ROW(3, 1), HALO_ROW(4, 1)},
                                            {HALO_ROW(1, 2), HALO_ROW(2, 2), HALO_ROW(3, 2), HALO_ROW(4, 2)},
                                            {HALO_ROW(1, 3), HALO_ROW(2, 3), HALO_ROW(3, 3), HALO_ROW(4, 3)},
                                            {HALO_ROW(1, 4), HALO_ROW(2, 4), HALO_ROW(3, 4), HALO_ROW(4, 4)},
                                            {HALO_ROW_D(1, 5), HALO_ROW_D(2, 5), HALO_ROW_D(3, 5), HALO_ROW_D(4, 5)},
                                            {{}, {}, {}, {}},
                                            {HALO_ROW(1, 7), HALO_ROW(2, 7), HALO_ROW(3, 7), HALO_ROW(4, 7)}};
#undef HALO_ROW
#undef HALO_ROW_D
  static bool attr_set[8][4][5] = {};
  const int ka = pool ? 2 : (direct ? (stats ? 5 : (p.scale ? 4 : 3)) : (stack ? 7 : (p.scale ? 1 : 0)));
  const int ki = pl.nacc - 1, kj = (ck == 16 ? 0 : (ck == 32 ? 1 : 2)) + (dual ? 3 : 0);
  HaloKernel kern = table[ka][ki][kj];
  PDAC_REQUIRE(kern != nullptr, "internal: no halo kernel for mode %d (nacc %d, variant %d)", ka, pl.nacc, kj);
  if (!attr_set[ka][ki][kj]) {
    PDAC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)((dual ? 105 : 221) * 1024)));
    if (dual)
      PDAC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    attr_set[ka][ki][kj] = true;
  }
  kern<<<per_n * p.tiles_c, dual ? kHaloThreadsDual : kHaloThreads, smem, (cudaStream_t)stream>>>(maps, p);
  PDAC_LAUNCH_CHECK("conv_halo_kernel");
  note_kernel(KID_CONV_HALO);
  *handled = 1;
  return PDAC_OK;
}

}  // namespace pdac

#ifdef PDAC_HALO_PROF
// Diagnostic build only (make prof): copies the 160 x 16 per-CTA cycle counters of the last conv_halo launch to the host.
extern "C" int pdac_halo_dbg_set(int v) {
  return cudaMemcpyToSymbol(pdac::g_halo_dbg, &v, sizeof(int)) == cudaSuccess ? 0 : -1;
}
extern "C" int pdac_halo_prof_read(long long* host) {
  return cudaMemcpyFromSymbol(host, pdac::g_halo_prof, sizeof(long long) * 160 * 16) == cudaSuccess ? 0 : -1;
}
#endif
