// Host side of the halo-tile convolution (planner, tensor maps, launch); the kernel lives in conv_halo_kernel.cuh.
#include "conv_halo_kernel.cuh"
#ifdef PDAC_HALO_PROF  // diagnostic build: one translation unit (the per-CTA counters are __device__ globals)
#include "conv_halo_k0.cu"
#include "conv_halo_k1.cu"
#include "conv_halo_k2.cu"
#include "conv_halo_k3.cu"
#endif

namespace pdac {
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
  const PdacView* up2 = epi_up2();  // `out` = the even pixels of this (2H, 2W) tensor
  const bool aligned32 = ((uintptr_t)out->ptr & 31) == 0 && out->sW % 16 == 0 && out->sH % 16 == 0 && out->sN % 16 == 0 &&
                         (!up2 || (up2->sW % 16 == 0 && up2->sH % 16 == 0));
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
  if (up2 && !direct_ns) return PDAC_OK;  // the replicated stores exist in the statistics-free direct epilogue only
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
  if (up2) {
    p.up2 = 1;
    p.up_dh = up2->sH;
    p.up_dw = up2->sW;
  }
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
  static bool attr_set[8][4][5] = {};
  const int ka = pool ? 2 : (direct ? (stats ? 5 : (p.scale ? 4 : 3)) : (stack ? 7 : (p.scale ? 1 : 0)));
  const int ki = pl.nacc - 1, kj = (ck == 16 ? 0 : (ck == 32 ? 1 : 2)) + (dual ? 3 : 0);
  HaloKernel kern = ka <= 1 ? halo_kernel_k0(ka, ki, kj)
                    : (ka == 2 || ka == 7 ? halo_kernel_k1(ka, ki, kj)
                                          : (ka <= 4 ? halo_kernel_k2(ka, ki, kj) : halo_kernel_k3(ka, ki, kj)));
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
