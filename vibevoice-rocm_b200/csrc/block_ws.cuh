// block_ws.cuh — whole Block1D in ONE kernel at C = 128 (the middle stage): block_tma_kernel's TMA-staged mixer front end
// joined to ffn_stream_kernel's streamed-weight FFN, so neither y (fp32) nor a (bf16) nor the hidden tensor reach HBM.
//
//   y   = x + gamma * ( dwconv7( RMSNorm(x) * norm_w ) + conv_bias )                      (reference :927-932 / :667-672)
//   out = y + ffn_gamma * ( W2 * gelu_erf( W1 * (RMSNorm(y) * ffn_norm_w) + b1 ) + b2 )   (:936-942 / :675-682)
//
// Why: round 2's stage 4 ran mixer_kernel (10 B per element: x in, y and a out) and then ffn_stream_kernel (10 B: a and y in,
// out out) -- 20 B per element where the block needs 8, 21 % of the K2 step.  At C = 128 the two weight matrices (128 KB
// each) do not fit beside the tiles, which is why block_tma_kernel stops at C = 64; here they are streamed in 64-unit hidden
// chunks through two TMA rings (they stay L2-resident).  What makes the rest fit, and keeps shared-memory BANDWIDTH (the
// tightest resource of this kernel: the first build moved ~1.1 MB per 128-row tile through a 128 B/clk port) in check:
//   * ONE x/y slot of [8 halo + 128 rows, 128 ch] fp32 (68 KB).  Its cycle (TMA load -> first statistic -> conv in place ->
//     second statistic) is shorter than a tile's GEMM phase, so a single slot keeps the tensor pipe fed; the loader refills it
//     as soon as the row warps have read y.
//   * The GEMM-1 operand a = bf16(y * rstd2) never touches shared memory: the row warps (thread = row = TMEM lane) write it
//     with tcgen05.st into 64 columns of TENSOR memory and GEMM 1 takes its A operand from there (like GEMM 2 takes H), which
//     removes 8 re-reads of a 32 KB tile per tile from the shared-memory port.  The same warps preload the D2 accumulator
//     with y + ffn_gamma * b2 (as block_tma_kernel), so E2 adds nothing and y is dead once they are done.
//   * A separate output staging tile (64 KB): E2 writes the finished accumulator there and a dedicated thread TMA-stores it,
//     off the slot's cycle.
//   * W1 ring of 3 and W2 ring of 2 chunks (16 KB each).  Tensor memory: D1 2 x 64, H 2 x 32, D2 2 x 128, A 64 = 512 columns.
// Warp roles (28 warps, 3 of them idle): E1 x 8 (GELU, H back into tensor memory), E2 x 4 (accumulator -> staging), R x 4 (thread = row: both
// statistics, stream context, a and the D2 preload), S2 x 4 (depthwise conv, one warp = one 32-row strip, lane = 4 channels),
// and five single-thread roles: weight loader, GEMM 1, GEMM 2, x loader, store.  All hand-offs are mbarriers.
// Streaming (STATE) as in block_tma_kernel: same state layout.
#pragma once
#include <cuda.h>
#include "ptx.cuh"
#include "epilogue.cuh"
#include "mixer.cuh"        // StateRef
#include "block_fused.cuh"  // reg_dec / reg_inc, rstd_of
#include "block_tma.cuh"    // reg_set, xy_cell

namespace vv {

struct BlockWsArgs {
  int nbatch, rows_pb;
  int tiles_pb;              // ceil(rows_pb / 128)
  float eps;
  const float* norm_w;       // [C]
  const float* dw_w;         // [7][C] tap-major
  const float* dw_b;         // [C]
  const float* gamma;        // [C]
  float b1c[512];            // [4C]  0.5 * b1 (W1 is packed pre-scaled by 0.5 and by ffn_norm_w): constant bank
  float gb2c[128];           // [C]   ffn_gamma * b2
  __nv_bfloat16* out_bf16;   // optional padded bf16 copy: [nbatch, ob_rpb, C], row t of an item lands at row ob_off + t
  long long ob_rpb;
  int ob_off;
  StateRef st;               // mixer context: fp32 [6][C] raw rows per stream (null base = offline)
};

#ifndef VV_BW_NW1
#define VV_BW_NW1 3
#endif
#ifndef VV_BW_NW2
#define VV_BW_NW2 2
#endif

struct BlockWsCfg {
  static constexpr int C = 128;
  static constexpr int HC = 64;                      // hidden units per chunk
  static constexpr int NCH = 4 * C / HC;             // 8 chunks per tile
  static constexpr int KB = C / 64;                  // 64-wide K blocks of a W1 chunk
  static constexpr int NBOX = C / 32;
  static constexpr int HALO = 8, XROWS = 128 + HALO;
  static constexpr int XBOX_BYTES = XROWS * 128;     // 17 KB
  static constexpr int X_BYTES = NBOX * XBOX_BYTES;  // 68 KB
  static constexpr int OBOX_BYTES = 128 * 128;       // 16 KB
  static constexpr int OUT_BYTES = NBOX * OBOX_BYTES;
  static constexpr int NW1 = VV_BW_NW1, NW2 = VV_BW_NW2, WLAG = 2;
  static constexpr int W1_BYTES = HC * C * 2;        // KB blocks of [64 rows x 128 B]
  static constexpr int W2_BYTES = C * HC * 2;        // [C rows x 128 B]
  static constexpr int E1W = 8, E2W = 4, RW = 4, S2W = 4, CTRL = 8;   // 5 control roles + 3 idle warps: see REG_LAUNCH
  static constexpr int SROWS = 128 / S2W;            // rows per conv strip: 32
  static constexpr int WARPS = E1W + E2W + RW + S2W + CTRL;
  static constexpr int THREADS = WARPS * 32;
  static constexpr int RSTD_FLOATS = 144;
  static constexpr int OFF_X = 0;
  static constexpr int OFF_OUT = OFF_X + X_BYTES;
  static constexpr int OFF_W1 = OFF_OUT + OUT_BYTES;
  static constexpr int OFF_W2 = OFF_W1 + NW1 * W1_BYTES;
  static constexpr int OFF_RSTD = OFF_W2 + NW2 * W2_BYTES;
  static constexpr int OFF_BAR = OFF_RSTD + RSTD_FLOATS * 4;
  static constexpr int ND2 = 2;
  static constexpr int NBARS = 8 + 2 * NW1 + 2 * NW2 + 8 + 3 * ND2;
  static constexpr int TOTAL = OFF_BAR + 512 + 1024;
  static_assert(NBARS * 8 + 8 <= 512, "barrier block");
  static_assert(OFF_OUT % 1024 == 0 && OFF_W1 % 1024 == 0 && OFF_W2 % 1024 == 0 && XBOX_BYTES % 1024 == 0,
                "swizzled tiles need 1 KB aligned bases");
  static_assert(TOTAL <= 232448, "exceeds the 227 KB dynamic shared memory limit");
  static constexpr uint32_t TMEM_COLS = 512;         // D1: 2 x 64 at 0; H: 2 x 32 at 128; D2: 2 x 128 at 192; A: 64 at 448
  static constexpr uint32_t H_COL0 = 2 * HC, D2_COL0 = 3 * HC, A_COL0 = 3 * HC + ND2 * C;
  static_assert(A_COL0 + C / 2 <= 512, "tensor memory columns");
  // Registers are allocated per SM sub-partition (warp id mod 4), so the warp count is kept a multiple of 4 and every
  // sub-partition holds the same mix (2 E1, 1 E2, 1 R, 1 S2, 2 control): 2 x 64 + 64 + 128 + 120 + 2 x 32 = 504 = 7 x 72.
  static constexpr int REG_LAUNCH = (65536 / THREADS) / 8 * 8;      // 72
  static_assert(WARPS % 4 == 0, "warps per sub-partition must be equal");
  static constexpr int REG_E1 = 64, REG_E2 = 64, REG_R = 128, REG_S2 = 120, REG_CTRL = 32;
  static_assert(32 * (E1W * REG_E1 + E2W * REG_E2 + RW * REG_R + S2W * REG_S2 + CTRL * REG_CTRL) <= THREADS * REG_LAUNCH,
                "setmaxnreg budget exceeds the CTA's launch allocation");
};

template <bool STATE>
__global__ void __launch_bounds__(BlockWsCfg::THREADS, 1)
block_ws_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmO,
                const __grid_constant__ CUtensorMap tmW1, const __grid_constant__ CUtensorMap tmW2,
                const BlockWsArgs p) {
  using L = BlockWsCfg;
  constexpr int C = L::C, NCH = L::NCH, NBOX = L::NBOX, HC = L::HC, KB = L::KB, ND2 = L::ND2;
  constexpr int NW1 = L::NW1, NW2 = L::NW2;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_align_1k(smem_raw);
  uint8_t* sX = smem + L::OFF_X;
  uint8_t* sOut = smem + L::OFF_OUT;
  uint8_t* sW1 = smem + L::OFF_W1;
  uint8_t* sW2 = smem + L::OFF_W2;
  float* sRstd = (float*)(smem + L::OFF_RSTD);
  uint64_t* bars = (uint64_t*)(smem + L::OFF_BAR);
  uint64_t* x_full = bars;                 // TMA landed the x tile
  uint64_t* x_free = bars + 1;             // R has read y: the slot may be refilled
  uint64_t* st_ready = bars + 2;           // R wrote rstd[] (and patched / saved the stream context)
  uint64_t* y_full = bars + 3;             // S2 wrote y in place
  uint64_t* a_full = bars + 4;             // R wrote the a tile into tensor memory
  uint64_t* a_empty = bars + 5;            // GEMM 1 has read it
  uint64_t* out_ready = bars + 6;          // E2 wrote the staging tile
  uint64_t* out_free = bars + 7;           // the TMA store has read it
  uint64_t* w1_full = bars + 8;            // [NW1]
  uint64_t* w1_empty = w1_full + NW1;      // [NW1]
  uint64_t* w2_full = w1_empty + NW1;      // [NW2]
  uint64_t* w2_empty = w2_full + NW2;      // [NW2]
  uint64_t* d1_full = w2_empty + NW2;      // [2]
  uint64_t* d1_empty = d1_full + 2;        // [2]
  uint64_t* h_full = d1_empty + 2;         // [2]
  uint64_t* h_empty = h_full + 2;          // [2]
  uint64_t* d2_full = h_empty + 2;         // [ND2] GEMM 2 finished the tile's accumulator
  uint64_t* d2_empty = d2_full + ND2;      // [ND2] E2 has read it
  uint64_t* d2_init = d2_empty + ND2;      // [ND2] R preloaded it with y + gb2
  uint32_t* tmem_slot = (uint32_t*)(d2_init + ND2);

  // hardware warps: [0, 8) E1, [8, 12) E2, [12, 16) R, [16, 20) S2, [20, 25): weights, G1, G2, x loader, store, [25, 28) idle
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  constexpr int WARP_E2 = L::E1W, WARP_R = WARP_E2 + L::E2W, WARP_S2 = WARP_R + L::RW, WARP_CTRL = WARP_S2 + L::S2W;
  const int ntiles = p.nbatch * p.tiles_pb;
  const int my_tiles = (ntiles > (int)blockIdx.x) ? (ntiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  const int total_g = my_tiles * NCH;
  auto tile_of = [&](int it, int& b, int& t0) {
    const int tile = (int)blockIdx.x + it * (int)gridDim.x;
    b = tile / p.tiles_pb;
    t0 = (tile - b * p.tiles_pb) * 128;
  };

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmX);
    tma_prefetch_desc(&tmO);
    tma_prefetch_desc(&tmW1);
    tma_prefetch_desc(&tmW2);
    mbar_init(x_full, 1);
    mbar_init(x_free, L::RW);
    mbar_init(st_ready, L::RW);
    mbar_init(y_full, L::S2W);
    mbar_init(a_full, L::RW);
    mbar_init(a_empty, 1);
    mbar_init(out_ready, L::E2W);
    mbar_init(out_free, 1);
    for (int i = 0; i < NW1; ++i) { mbar_init(&w1_full[i], 1); mbar_init(&w1_empty[i], 1); }
    for (int i = 0; i < NW2; ++i) { mbar_init(&w2_full[i], 1); mbar_init(&w2_empty[i], 1); }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&d1_full[i], 1);
      mbar_init(&d1_empty[i], L::E1W / 2);
      mbar_init(&h_full[i], L::E1W / 2);
      mbar_init(&h_empty[i], 1);
    }
    for (int i = 0; i < ND2; ++i) {
      mbar_init(&d2_full[i], 1);
      mbar_init(&d2_empty[i], L::E2W);
      mbar_init(&d2_init[i], L::RW);
    }
    fence_mbar_init();
  }
  if (warp == WARP_CTRL + 1) {
    tmem_alloc(tmem_slot, L::TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch();
  pdl_wait();

  if (warp >= WARP_CTRL) {
    reg_set<L::REG_LAUNCH, L::REG_CTRL>();
    const int role = warp - WARP_CTRL;
    if (role == 0) {
      // ===== weight chunks: W1 chunk [64 hidden, C] (KB blocks of [64 rows x 128 B]) and, WLAG chunks behind, W2 chunk
      // [C, 64 hidden].  One in-order producer cannot deadlock: every wait depends only on loads issued earlier.
      if (elect_one()) {
        auto load_w2 = [&](int gp) {
          const int s2 = gp % NW2;
          mbar_wait_sleep(&w2_empty[s2], (uint32_t)(((gp / NW2) & 1) ^ 1), 20);
          mbar_expect_tx(&w2_full[s2], L::W2_BYTES);
          tma_load_2d(sW2 + s2 * L::W2_BYTES, &tmW2, &w2_full[s2], (gp % NCH) * HC, 0);
        };
        for (int g = 0; g < total_g; ++g) {
          const int s1 = g % NW1;
          mbar_wait_sleep(&w1_empty[s1], (uint32_t)(((g / NW1) & 1) ^ 1), 20);
          mbar_expect_tx(&w1_full[s1], L::W1_BYTES);
#pragma unroll
          for (int kb = 0; kb < KB; ++kb)
            tma_load_2d(sW1 + s1 * L::W1_BYTES + kb * (HC * 128), &tmW1, &w1_full[s1], kb * 64, (g % NCH) * HC);
          if (g >= L::WLAG) load_w2(g - L::WLAG);
        }
        for (int gp = (total_g > L::WLAG ? total_g - L::WLAG : 0); gp < total_g; ++gp) load_w2(gp);
      }
    } else if (role == 1) {
      // ===== GEMM 1 issuer: D1[g & 1] = a * W1_chunk^T, A operand from tensor memory ===================================
      if (elect_one()) {
        constexpr uint32_t idesc1 = umma_idesc_bf16(128, HC);
        const uint64_t dw0 = umma_smem_desc<128>(smem_u32(sW1));
        const uint32_t aa = tmem_base + L::A_COL0;
        int g = 0;
        for (int it = 0; it < my_tiles; ++it) {
          mbar_wait_sleep(a_full, (uint32_t)(it & 1), 32);
          VV_BWT(8, it);
#pragma unroll 1
          for (int j = 0; j < NCH; ++j, ++g) {
            const int db = g & 1, s1 = g % NW1;
            mbar_wait2(&d1_empty[db], (uint32_t)(((g >> 1) & 1) ^ 1), &w1_full[s1], (uint32_t)((g / NW1) & 1));
            tc_fence_after();
            const uint64_t dw = dw0 + (uint64_t)((s1 * L::W1_BYTES) >> 4);
            const uint32_t d = tmem_base + (uint32_t)(db * HC);
#pragma unroll
            for (int kb = 0; kb < KB; ++kb) {
#pragma unroll
              for (int kk = 0; kk < 4; ++kk)       // channels [64 kb + 16 kk, +16) = 8 packed columns of the a tile
                umma_bf16_ts(d, aa + (uint32_t)(kb * 32 + kk * 8), dw + (uint64_t)(((kb * HC * 128) >> 4) + kk * 2), idesc1,
                             (kb | kk) != 0);
            }
            umma_commit(&d1_full[db]);
            umma_commit(&w1_empty[s1]);
          }
          umma_commit(a_empty);
          VV_BWT(9, it);
        }
      }
    } else if (role == 2) {
      // ===== GEMM 2 issuer: D2[it % 2] (preloaded with y + gb2) += H * W2_chunk^T, A operand from tensor memory ===========
      if (elect_one()) {
        constexpr uint32_t idesc2 = umma_idesc_bf16(128, C);
        const uint64_t dw0 = umma_smem_desc<128>(smem_u32(sW2));
        int g = 0;
        for (int it = 0; it < my_tiles; ++it) {
          const int d2b = it % ND2;
          mbar_wait_sleep(&d2_init[d2b], (uint32_t)((it / ND2) & 1), 32);
          tc_fence_after();
          const uint32_t d = tmem_base + L::D2_COL0 + (uint32_t)(d2b * C);
#pragma unroll 1
          for (int j = 0; j < NCH; ++j, ++g) {
            const int hb = g & 1, s2 = g % NW2;
            mbar_wait2(&w2_full[s2], (uint32_t)((g / NW2) & 1), &h_full[hb], (uint32_t)((g >> 1) & 1));
            tc_fence_after();
            const uint32_t ah = tmem_base + L::H_COL0 + (uint32_t)(hb * (HC / 2));
            const uint64_t dw = dw0 + (uint64_t)((s2 * L::W2_BYTES) >> 4);
#pragma unroll
            for (int kk = 0; kk < HC / 16; ++kk) umma_bf16_ts(d, ah + (uint32_t)(kk * 8), dw + (uint64_t)(kk * 2), idesc2, 1u);
            umma_commit(&h_empty[hb]);
            umma_commit(&w2_empty[s2]);
          }
          umma_commit(&d2_full[d2b]);
          VV_BWT(10, it);
        }
      }
    } else if (role == 3) {
      // ===== x loader: one [XROWS, 32] box per 32 channels, rows t0 - 8 .. t0 + 127 (rows outside the item read as zero) ====
      if (elect_one()) {
        for (int it = 0; it < my_tiles; ++it) {
          int b, t0;
          tile_of(it, b, t0);
          if (it > 0) mbar_wait_sleep(x_free, (uint32_t)((it - 1) & 1), 64);
          VV_BWT(0, it);
          mbar_expect_tx(x_full, L::X_BYTES);
#pragma unroll
          for (int c = 0; c < NBOX; ++c) tma_load_3d(sX + c * L::XBOX_BYTES, &tmX, x_full, c * 32, t0 - L::HALO, b);
        }
      }
    } else if (role == 4) {
      // ===== store: finished staging tile -> global =====================================================================
      if (elect_one()) {
        for (int it = 0; it < my_tiles; ++it) {
          int b, t0;
          tile_of(it, b, t0);
          mbar_wait_sleep(out_ready, (uint32_t)(it & 1), 64);
          VV_BWT(13, it);
#pragma unroll
          for (int c = 0; c < NBOX; ++c) tma_store_3d(&tmO, sOut + c * L::OBOX_BYTES, c * 32, t0, b);
          tma_store_commit();
          tma_store_wait_read<0>();
          mbar_arrive(out_free);
          VV_BWT(14, it);
        }
        tma_store_wait_all<0>();
      }
    }
  } else if (warp < WARP_E2) {
    // ===== E1: D1 -> + b1 -> GELU -> bf16 pairs -> H in tensor memory (as block_tma_kernel) =============================
    reg_set<L::REG_LAUNCH, L::REG_E1>();
    const int q = warp & 3;
    const int grp = warp >> 2;
    const uint32_t lane_base = (uint32_t)(q * 32) << 16;
    const uint32_t d1addr = tmem_base + lane_base + (uint32_t)(grp * HC);
    const uint32_t haddr = tmem_base + lane_base + L::H_COL0 + (uint32_t)(grp * (HC / 2));
    for (int g = grp; g < total_g; g += 2) {
      const int j = g % NCH;
      const uint32_t ph = (uint32_t)((g >> 1) & 1);
      const float* bp = p.b1c + j * HC;                  // constant bank (by-value kernel argument)
      if (threadIdx.x == 0 && (j == 0 || j == 2)) VV_BWT(16 + j * 4, g / NCH);
      mbar_wait(&d1_full[grp], ph);
      if (threadIdx.x == 0 && (j == 0 || j == 2)) VV_BWT(17 + j * 4, g / NCH);
      tc_fence_after();
#pragma unroll
      for (int hf = 0; hf < HC / 32; ++hf) {
        uint32_t v[32];
        tmem_ld_32x32(d1addr + (uint32_t)(hf * 32), v);
        tmem_ld_wait();
        if (threadIdx.x == 0 && j == 0) VV_BWT(18 + hf * 2, g / NCH);
        if (hf == HC / 32 - 1) {                     // whole accumulator is in registers: the next G1 may overwrite it
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&d1_empty[grp]);
        }
        uint32_t hv[16];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float* bq = bp + hf * 32 + i * 8;
          hv[i * 4 + 0] = gelu2_bias_pack(__uint_as_float(v[i * 8 + 0]), __uint_as_float(v[i * 8 + 1]), bq[0], bq[1]);
          hv[i * 4 + 1] = gelu2_bias_pack(__uint_as_float(v[i * 8 + 2]), __uint_as_float(v[i * 8 + 3]), bq[2], bq[3]);
          hv[i * 4 + 2] = gelu2_bias_pack(__uint_as_float(v[i * 8 + 4]), __uint_as_float(v[i * 8 + 5]), bq[4], bq[5]);
          hv[i * 4 + 3] = gelu2_bias_pack(__uint_as_float(v[i * 8 + 6]), __uint_as_float(v[i * 8 + 7]), bq[6], bq[7]);
        }
        if (hf == 0) {
          mbar_wait(&h_empty[grp], ph ^ 1);          // GEMM 2 has consumed this group's previous H chunk
          tc_fence_after();
        }
        tmem_st_32x16(haddr + (uint32_t)(hf * 16), hv);   // thread = row: pair (k, k+1) -> one 32-bit column
        if (threadIdx.x == 0 && j == 0) VV_BWT(19 + hf * 2, g / NCH);
      }
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&h_full[grp]);
      if (threadIdx.x == 0 && j == 0) VV_BWT(22, g / NCH);
    }
  } else if (warp < WARP_R) {
    // ===== E2: the accumulator holds out = (y + gb2) + H W2^T: tcgen05.ld -> staging tile (thread = row) -> TMA store ======
    reg_set<L::REG_LAUNCH, L::REG_E2>();
    const int q = warp & 3;
    const int r = q * 32 + lane;
    for (int it = 0; it < my_tiles; ++it) {
      const int d2b = it % ND2;
      int b, t0;
      tile_of(it, b, t0);
      const int t = t0 + r;
      const bool valid = t < p.rows_pb;
      const long long brow = (long long)b * p.ob_rpb + p.ob_off + t;
      mbar_wait(&d2_full[d2b], (uint32_t)((it / ND2) & 1));
      if (threadIdx.x == WARP_E2 * 32) VV_BWT(11, it);
      mbar_wait(out_free, (uint32_t)((it & 1) ^ 1));          // the previous tile's store has read the staging tile
      tc_fence_after();
#pragma unroll 1
      for (int c = 0; c < NBOX; ++c) {
        uint32_t v[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + L::D2_COL0 + (uint32_t)(d2b * C + c * 32), v);
        tmem_ld_wait();
        uint8_t* box = sOut + c * L::OBOX_BYTES;
#pragma unroll
        for (int i = 0; i < 8; ++i)
          *reinterpret_cast<uint4*>(box + xy_cell(r, i)) = make_uint4(v[i * 4], v[i * 4 + 1], v[i * 4 + 2], v[i * 4 + 3]);
        if (valid && p.out_bf16 != nullptr) {
          __nv_bfloat16* bp = p.out_bf16 + brow * C + c * 32;
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            uint4 pk;
            pk.x = pack_bf16x2(__uint_as_float(v[i * 8 + 0]), __uint_as_float(v[i * 8 + 1]));
            pk.y = pack_bf16x2(__uint_as_float(v[i * 8 + 2]), __uint_as_float(v[i * 8 + 3]));
            pk.z = pack_bf16x2(__uint_as_float(v[i * 8 + 4]), __uint_as_float(v[i * 8 + 5]));
            pk.w = pack_bf16x2(__uint_as_float(v[i * 8 + 6]), __uint_as_float(v[i * 8 + 7]));
            *reinterpret_cast<uint4*>(bp + i * 8) = pk;
          }
        }
      }
      tc_fence_before();
      fence_proxy_async_smem();                 // staged tile -> visible to the TMA store
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&d2_empty[d2b]);
        mbar_arrive(out_ready);
      }
      if (threadIdx.x == WARP_E2 * 32) VV_BWT(12, it);
    }
  } else if (warp < WARP_S2) {
    // ===== R: thread = row.  stat1(it): rstd of the raw row (first RMSNorm) + stream context in / out, as block_tma_kernel's
    // S1.  Then per tile: second statistic of y, a = bf16(y * rstd2) -> tensor memory, D2 preload with y + gb2, slot released,
    // and stat1 of the next tile once the loader has refilled the slot.
    reg_set<L::REG_LAUNCH, L::REG_R>();
    const int r = (warp - WARP_R) * 32 + lane;            // tile row 0..127 -> slot row HALO + r
    const float inv_c = 1.0f / (float)C;
    auto row_rstd = [&](int srow) -> float {
      f32x2 acc[4] = {0ull, 0ull, 0ull, 0ull};            // four independent chains
#pragma unroll
      for (int c = 0; c < NBOX; ++c) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const ulonglong2 v = *reinterpret_cast<const ulonglong2*>(sX + c * L::XBOX_BYTES + xy_cell(srow, i));
          acc[(i & 1) * 2] = f2_fma(v.x, v.x, acc[(i & 1) * 2]);
          acc[(i & 1) * 2 + 1] = f2_fma(v.y, v.y, acc[(i & 1) * 2 + 1]);
        }
      }
      float s0, s1, s2, s3, s4, s5, s6, s7;
      f2_unpack(acc[0], s0, s1);
      f2_unpack(acc[1], s2, s3);
      f2_unpack(acc[2], s4, s5);
      f2_unpack(acc[3], s6, s7);
      return rstd_of(((s0 + s1) + (s2 + s3)) + ((s4 + s5) + (s6 + s7)), inv_c, p.eps);
    };
    auto stat1 = [&](int it) {
      int b, t0;
      tile_of(it, b, t0);
      // the stream's state pointers are two dependent global loads: only tiles that read the context (item start) or save it
      // (the item's last 6 rows) need them, and they are fetched while the x tile is still in flight
      const uint8_t* rd8 = nullptr;
      uint8_t* wr8 = nullptr;
      if (STATE && p.st.base != nullptr && (t0 == 0 || t0 + 128 > p.rows_pb - 6)) p.st.resolve(b, rd8, wr8);
      mbar_wait(x_full, (uint32_t)(it & 1));
      if (threadIdx.x == WARP_R * 32) VV_BWT(1, it);
      if (warp == WARP_R + 3 && lane < L::HALO) {
        // halo rows (slot rows 0..7 = t0 - 8 .. t0 - 1): at the item start rows 2..7 are the stream's cached raw rows
        // (TMA delivered zeros = the offline causal pad, :398-403)
        if (STATE && t0 == 0 && rd8 != nullptr && lane >= 2) {
          const float* src = reinterpret_cast<const float*>(rd8) + (lane - 2) * C;
#pragma unroll
          for (int c = 0; c < NBOX; ++c)
#pragma unroll
            for (int i = 0; i < 8; ++i)
              *reinterpret_cast<float4*>(sX + c * L::XBOX_BYTES + xy_cell(lane, i)) =
                  *reinterpret_cast<const float4*>(src + c * 32 + i * 4);
        }
        sRstd[lane + 2] = row_rstd(lane);
        // the stream's new context = last 6 raw rows of [old context, x] (:371-382): rows before the item start count
        if (STATE && wr8 != nullptr && t0 == 0) {
          const int t = lane - L::HALO;                                     // -8 .. -1
          const int dst = t - (p.rows_pb - 6);
          if (dst >= 0) {
            float* o = reinterpret_cast<float*>(wr8) + dst * C;
#pragma unroll
            for (int c = 0; c < NBOX; ++c)
#pragma unroll
              for (int i = 0; i < 8; ++i)
                *reinterpret_cast<float4*>(o + c * 32 + i * 4) =
                    *reinterpret_cast<const float4*>(sX + c * L::XBOX_BYTES + xy_cell(lane, i));
          }
        }
      }
      sRstd[L::HALO + r + 2] = row_rstd(L::HALO + r);
      if (STATE && wr8 != nullptr) {
        const int t = t0 + r;
        const int dst = t - (p.rows_pb - 6);
        if (dst >= 0 && t < p.rows_pb) {
          float* o = reinterpret_cast<float*>(wr8) + dst * C;
#pragma unroll
          for (int c = 0; c < NBOX; ++c)
#pragma unroll
            for (int i = 0; i < 8; ++i)
              *reinterpret_cast<float4*>(o + c * 32 + i * 4) =
                  *reinterpret_cast<const float4*>(sX + c * L::XBOX_BYTES + xy_cell(L::HALO + r, i));
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(st_ready);
      if (threadIdx.x == WARP_R * 32) VV_BWT(2, it);
    };
    // one box of the row (32 channels = cells[0..7]): 16 packed a columns + 32 accumulator columns of y + gb2
    auto emit_box = [&](const ulonglong2 (&cell)[8], int c, f32x2 rr, uint32_t aaddr, uint32_t daddr) {
      uint32_t w[16];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        float a0, a1, a2, a3;
        f2_unpack(f2_mul(cell[i].x, rr), a0, a1);
        f2_unpack(f2_mul(cell[i].y, rr), a2, a3);
        w[i * 2 + 0] = pack_bf16x2(a0, a1);
        w[i * 2 + 1] = pack_bf16x2(a2, a3);
      }
      tmem_st_32x16(aaddr + (uint32_t)(c * 16), w);
#pragma unroll
      for (int h = 0; h < 2; ++h) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          float y0, y1, y2, y3;
          f2_unpack(cell[h * 4 + i].x, y0, y1);
          f2_unpack(cell[h * 4 + i].y, y2, y3);
          const int ch = c * 32 + (h * 4 + i) * 4;
          w[i * 4 + 0] = __float_as_uint(y0 + p.gb2c[ch + 0]);
          w[i * 4 + 1] = __float_as_uint(y1 + p.gb2c[ch + 1]);
          w[i * 4 + 2] = __float_as_uint(y2 + p.gb2c[ch + 2]);
          w[i * 4 + 3] = __float_as_uint(y3 + p.gb2c[ch + 3]);
        }
        tmem_st_32x16(daddr + (uint32_t)(c * 32 + h * 16), w);
      }
    };
    const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
    const uint32_t aaddr = tmem_base + lane_base + L::A_COL0;
    if (my_tiles > 0) stat1(0);
    for (int it = 0; it < my_tiles; ++it) {
      const int d2b = it % ND2;
      mbar_wait(y_full, (uint32_t)(it & 1));
      if (threadIdx.x == WARP_R * 32) VV_BWT(5, it);
      // channels 0..63 (boxes 0, 1) stay in registers; channels 64..127 are read twice (statistic, then conversion)
      ulonglong2 ylo[16];
      f32x2 acc[4] = {0ull, 0ull, 0ull, 0ull};
#pragma unroll
      for (int c = 0; c < 2; ++c)
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          ylo[c * 8 + i] = *reinterpret_cast<const ulonglong2*>(sX + c * L::XBOX_BYTES + xy_cell(L::HALO + r, i));
          acc[(i & 1) * 2] = f2_fma(ylo[c * 8 + i].x, ylo[c * 8 + i].x, acc[(i & 1) * 2]);
          acc[(i & 1) * 2 + 1] = f2_fma(ylo[c * 8 + i].y, ylo[c * 8 + i].y, acc[(i & 1) * 2 + 1]);
        }
#pragma unroll
      for (int c = 2; c < 4; ++c)
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const ulonglong2 v = *reinterpret_cast<const ulonglong2*>(sX + c * L::XBOX_BYTES + xy_cell(L::HALO + r, i));
          acc[(i & 1) * 2] = f2_fma(v.x, v.x, acc[(i & 1) * 2]);
          acc[(i & 1) * 2 + 1] = f2_fma(v.y, v.y, acc[(i & 1) * 2 + 1]);
        }
      float s0, s1, s2, s3, s4, s5, s6, s7;
      f2_unpack(acc[0], s0, s1);
      f2_unpack(acc[1], s2, s3);
      f2_unpack(acc[2], s4, s5);
      f2_unpack(acc[3], s6, s7);
      const float rstd2 = rstd_of(((s0 + s1) + (s2 + s3)) + ((s4 + s5) + (s6 + s7)), inv_c, p.eps);
      const f32x2 rr = f2_pack(rstd2, rstd2);
      mbar_wait(&d2_empty[d2b], (uint32_t)(((it / ND2) & 1) ^ 1));      // E2 drained this accumulator two tiles ago
      mbar_wait(a_empty, (uint32_t)((it & 1) ^ 1));                     // GEMM 1 has read the previous tile's a
      tc_fence_after();
      const uint32_t daddr = tmem_base + lane_base + L::D2_COL0 + (uint32_t)(d2b * C);
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        const ulonglong2 cell[8] = {ylo[c * 8], ylo[c * 8 + 1], ylo[c * 8 + 2], ylo[c * 8 + 3],
                                    ylo[c * 8 + 4], ylo[c * 8 + 5], ylo[c * 8 + 6], ylo[c * 8 + 7]};
        emit_box(cell, c, rr, aaddr, daddr);
      }
#pragma unroll
      for (int c = 2; c < 4; ++c) {
        ulonglong2 cell[8];
#pragma unroll
        for (int i = 0; i < 8; ++i)
          cell[i] = *reinterpret_cast<const ulonglong2*>(sX + c * L::XBOX_BYTES + xy_cell(L::HALO + r, i));
        emit_box(cell, c, rr, aaddr, daddr);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(x_free);       // every y value of this warp's rows is in registers / tensor memory
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(a_full);
        mbar_arrive(&d2_init[d2b]);
      }
      if (threadIdx.x == WARP_R * 32) VV_BWT(7, it);
      if (it + 1 < my_tiles) stat1(it + 1);
    }
  } else {
    // ===== S2: depthwise causal conv k7 over the normalised rows; lane = 4 channels, one warp = one strip of 32 rows ======
    reg_set<L::REG_LAUNCH, L::REG_S2>();
    struct P4 { f32x2 lo, hi; };
    constexpr int SR = L::SROWS, NB = 8;                            // rows are loaded NB at a time: the in-place stores
    const int gi = warp - WARP_S2;                                  // would otherwise order every load behind the last store
    const int c0 = lane * 4;
    const int box = c0 >> 5, qc = (c0 & 31) >> 2;
    // folded taps: y = x + gb + sum_j wq[j] * n[t-6+j],  n = x * rstd,  wq = gamma * w * norm_w,  gb = gamma * conv_bias
    P4 wq[7], gb;
    {
      const float4 nw = __ldg(reinterpret_cast<const float4*>(p.norm_w + c0));
      const float4 gm = __ldg(reinterpret_cast<const float4*>(p.gamma + c0));
      const float4 cb = __ldg(reinterpret_cast<const float4*>(p.dw_b + c0));
      gb.lo = f2_pack(gm.x * cb.x, gm.y * cb.y);
      gb.hi = f2_pack(gm.z * cb.z, gm.w * cb.w);
#pragma unroll
      for (int j = 0; j < 7; ++j) {
        const float4 w = __ldg(reinterpret_cast<const float4*>(p.dw_w + j * C + c0));
        wq[j].lo = f2_pack(gm.x * w.x * nw.x, gm.y * w.y * nw.y);
        wq[j].hi = f2_pack(gm.z * w.z * nw.z, gm.w * w.w * nw.w);
      }
    }
    const int sr0 = L::HALO + gi * SR;                              // slot row of the strip's first output row
    uint8_t* bx = sX + box * L::XBOX_BYTES;
    const float* rs = sRstd + sr0 - 4;                              // rs[i] = rstd of slot row sr0 - 6 + i
    for (int it = 0; it < my_tiles; ++it) {
      mbar_wait(st_ready, (uint32_t)(it & 1));
      if (threadIdx.x == WARP_S2 * 32) VV_BWT(3, it);
      P4 win[7];
#pragma unroll
      for (int i = 0; i < 6; ++i) {                                  // warm-up rows: a load and a multiply each
        const ulonglong2 v = *reinterpret_cast<const ulonglong2*>(bx + xy_cell(sr0 - 6 + i, qc));
        const float rsi = rs[i];
        const f32x2 rr = f2_pack(rsi, rsi);
        win[i].lo = f2_mul(v.x, rr);
        win[i].hi = f2_mul(v.y, rr);
      }
      // y is written in place: every strip must hold its warm-up rows (its neighbour's last rows) before anyone stores
      named_bar_sync(1, L::S2W * 32);
#pragma unroll
      for (int k0 = 0; k0 < SR; k0 += NB) {
        ulonglong2 xv[NB];
        float rv[NB];
#pragma unroll
        for (int k = 0; k < NB; ++k) {
          xv[k] = *reinterpret_cast<const ulonglong2*>(bx + xy_cell(sr0 + k0 + k, qc));
          rv[k] = rs[6 + k0 + k];
        }
#pragma unroll
        for (int k = 0; k < NB; ++k) {
          const int kk = k0 + k;
          const f32x2 rr = f2_pack(rv[k], rv[k]);
          win[(6 + kk) % 7].lo = f2_mul(xv[k].x, rr);
          win[(6 + kk) % 7].hi = f2_mul(xv[k].y, rr);
          P4 acc;
          acc.lo = f2_add(xv[k].x, gb.lo);
          acc.hi = f2_add(xv[k].y, gb.hi);
#pragma unroll
          for (int j = 0; j < 7; ++j) {                              // tap j reads n[t - 6 + j] = win[(kk + j) % 7]
            acc.lo = f2_fma(wq[j].lo, win[(kk + j) % 7].lo, acc.lo);
            acc.hi = f2_fma(wq[j].hi, win[(kk + j) % 7].hi, acc.hi);
          }
          *reinterpret_cast<ulonglong2*>(bx + xy_cell(sr0 + kk, qc)) = make_ulonglong2(acc.lo, acc.hi);
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(y_full);
      if (threadIdx.x == WARP_S2 * 32) VV_BWT(4, it);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == WARP_CTRL + 1) tmem_dealloc(tmem_base, L::TMEM_COLS);
}

}  // namespace vv
