// block_tma.cuh — second-generation whole-Block1D kernel for the narrow stages (C = 32, 64): the x tile is STAGED ONCE by
// TMA into shared memory and every row statistic is computed exactly once, by a thread that owns the row.
//
//   y   = x + gamma * ( dwconv7( RMSNorm(x) * norm_w ) + conv_bias )                      (reference :927-932 / :667-672)
//   out = y + ffn_gamma * ( W2 * gelu_erf( W1 * (RMSNorm(y) * ffn_norm_w) + b1 ) + b2 )   (:936-942 / :675-682)
//
// Why (ncu of block_fused_kernel, profiles/block_fused_s5_r01.txt + the SASS page of the C = 32 capture): that kernel
// moves exactly the algorithmic bytes but runs at 31-36 % of HBM speed because it is bound INSIDE the SM: its mixer warps
// load x rows global -> registers with 6 warm-up rows per 8 produced (1.75x L1 read amplification, L1TEX the busiest unit),
// normalise every row 1.75 times and reduce every statistic with a 16-lane butterfly (8 of ~50 instructions per row and
// lane, twice), at ~0.2 IPC on the critical path.  Per 128-row tile it executes ~10 000 warp instructions at C = 64 of
// which ~4 800 belong to the mixer.  Here:
//   * TMA loads the [8 halo + 128, C] fp32 tile (3-D map, rows before the item start are zero-filled by the hardware =
//     the causal left pad, :398-403) into a 128B-swizzled ring slot; nothing is loaded twice.
//   * S1 warps (thread = row) read their row once, conflict-free, and write rstd[row]: no shuffles, no redundancy.
//   * S2 warps (lane = 4 channels, register sliding window over a 16- or 8-row strip) do the depthwise conv; warm-up rows
//     cost a load and a multiply, not a statistic.  y is written IN PLACE over x in the slot.
//   * S3 warps (thread = row) compute the second statistic and write a = bf16(y * rstd2) into the swizzled GEMM-1 operand
//     tile (ffn_norm_w is folded into W1 at pack time, ffn_gamma into W2, ffn_gamma * b2 is a per-channel constant).
//   * GEMM 1 / E1 (GELU, H back into tensor memory) / GEMM 2 (A from TMEM) as in block_fused_kernel, with HC-column
//     chunks (HC = 128: the a tile is read twice per tile from shared memory instead of four times).
//   * E2 (thread = row): out = y + d2 + gb2, in place in the slot -> TMA store; the slot then goes back to the loader.
// Mixer instruction count per tile: ~2 200 (C = 64) instead of ~4 800; no global load instruction on the hot path.
// Second build, after ncu of the first (profiles/block_tma_s{5,6}_r02.txt: DRAM traffic = algorithmic bytes, the busiest
// unit is the LSU DATA PIPE at 60 / 75 % -- ~1340 shared-memory wavefronts per tile at C = 32, conflict-free, plus ~790
// wavefronts of warp-uniform __ldg bias loads): (1) b1 and ffn_gamma * b2 travel in the kernel's CONSTANT bank (by-value
// arguments), so E1 / S3 read them as constant operands instead of LSU loads; (2) S3 -- thread = row = TMEM lane, y already
// in registers -- writes y + gb2 into the D2 ACCUMULATOR with tcgen05.st and GEMM 2 accumulates on top of it, so E2 no
// longer re-reads y from shared memory (tcgen05.ld -> st.shared -> TMA store); four D2 buffers keep S3 ahead of E2.
// Streaming (STATE): the first tile of an item patches its 6 halo rows from the stream's cached raw rows before S1 runs;
// S1 saves the item's last 6 raw rows (reference :327-382; same state layout as block_fused_kernel / mixer_kernel).
#pragma once
#include <cuda.h>
#include "ptx.cuh"
#include "epilogue.cuh"
#include "mixer.cuh"        // StateRef
#include "block_fused.cuh"  // reg_dec / reg_inc, rstd_of

namespace vv {

#ifndef VV_BT_HC
#define VV_BT_HC 64         // hidden units per GEMM-1 chunk (64, or 128 at C = 64 with two D2 buffers)
#endif
#ifndef VV_BT_NX64
#define VV_BT_NX64 3        // x/y ring slots at C = 64 (34 KB each).  Third build: a slot is released as soon as the S3 warps hold
                            // its y rows (the clock64 timeline of the second build, profiles/block_tma_timeline_r02.txt, showed a
                            // slot's life of 17 500 clocks = 4 tile periods with 4 slots, 9 000 of them spent waiting for the GEMM
                            // phase only to receive the output at its end); the output goes through a separate ring of staging
                            // boxes.  History: 3 slots held to the store 1.371 ms, 4 slots 1.240 ms per 5760 frames
#endif
#ifndef VV_BT_NGRP
#define VV_BT_NGRP 2        // E1 groups (4 warps each; group G owns D1 buffer G and H buffer G, chunks go round-robin).  The clock64
                            // timelines show the chunk chain G1 -> E1 -> G2 as the pace-setter of all fused kernels at ~1 000
                            // clocks per chunk (one group needs ~1 600 per chunk).  A THIRD group fits the tensor memory at C <= 64
                            // and was measured (profiles/block_tma_e1groups_ab_r02.txt): each E1 warp then takes ~2 900 clocks per
                            // chunk instead of 1 600 -- the sub-partition's FMA / MUFU / issue mix is what saturates, at ~60 % of
                            // the pure-GELU rate -- and the kernel is no faster (1.200 / 1.197 ms against 1.213 / 1.178), so 2 stays
#endif
#ifndef VV_BT_NSTG
#define VV_BT_NSTG 2        // output staging boxes of [128 rows, 32 fp32] (16 KB each) at C = 64; 4 at C = 32
#endif
#ifndef VV_BT_NX32
#define VV_BT_NX32 6        // at C = 32 (17 KB each)
#endif

#ifdef VV_BW_TRACE     // timeline instrumentation (A/B builds only): clock64 stamps of CTA 0 for its tiles [8, 24)
__device__ long long vv_bw_trace[16 * 32];
#define VV_BWT(slot, it)                                                                                 \
  do {                                                                                                   \
    if (blockIdx.x == 0 && (it) >= 8 && (it) < 24) vv_bw_trace[((it) - 8) * 32 + (slot)] = clock64();    \
  } while (0)
#else
#define VV_BWT(slot, it) do { } while (0)
#endif

struct BlockTmaArgs {
  int nbatch, rows_pb;
  int tiles_pb;              // ceil(rows_pb / 128)
  float eps;
  const float* x;            // [nbatch, rows_pb, C] fp32: only the STATE form reads it directly (nothing offline)
  const float* norm_w;       // [C]
  const float* dw_w;         // [7][C] tap-major
  const float* dw_b;         // [C]
  const float* gamma;        // [C]
  float b1c[256];            // [4C]  linear1 bias pre-scaled by 0.5 like W1: lives in the constant bank with the other
  float gb2c[64];            // [C]   arguments; ffn_gamma * b2
  __nv_bfloat16* out_bf16;   // optional padded bf16 copy: [nbatch, ob_rpb, C], row t of an item lands at row ob_off + t
  long long ob_rpb;
  int ob_off;
  StateRef st;               // mixer context: fp32 [6][C] raw rows per stream (null base = offline)
};

template <int C>
struct BlockTmaCfg {
  static_assert(C == 32 || C == 64, "block_tma_kernel is laid out for C = 32 / 64");
  static constexpr int HC = (VV_BT_HC == 128 && C == 64) ? 128 : 64;   // 4C = 128 at C = 32: two 64-wide chunks
  static constexpr int NCH = 4 * C / HC;
  static constexpr int ROWB = C * 2;                 // bytes per a / W1 row (64 or 128 = the swizzle span)
  static constexpr int NBOX = C / 32;
  static constexpr int LPR = C / 4;                  // S2 lanes per row
  static constexpr int HALO = 8, XROWS = 128 + HALO;
  static constexpr int XBOX_BYTES = XROWS * 128;     // 17 408 = 17 KB: box bases stay 1 KB aligned
  static constexpr int XY_BYTES = NBOX * XBOX_BYTES;
  static constexpr int NX = (C == 64) ? VV_BT_NX64 : VV_BT_NX32;
  static constexpr int NA = 2;                        // GEMM-1 operand buffers, in TENSOR memory (C / 2 packed columns each)
  static constexpr int NG = VV_BT_NGRP;              // E1 groups
#ifndef VV_BT_EC
#define VV_BT_EC 32
#endif
  static constexpr int EC = VV_BT_EC;                // accumulator columns an E1 warp holds at a time (32, or 16 to save registers)
  static constexpr int E1W = 4 * NG, E2W = 4, S1W = 4, S3W = 4, S2W = 4, CTRL = 4;
  static constexpr int STRIPS = S2W * 32 / LPR;      // 8 (C = 64) or 16 (C = 32)
  static constexpr int R = 128 / STRIPS;             // rows per strip: 16 or 8
  static constexpr int WARPS = E1W + E2W + S1W + S3W + S2W + CTRL;
  static constexpr int THREADS = WARPS * 32;
  static constexpr int W1_BYTES = 4 * C * ROWB;
  static constexpr int W2_BYTES = (4 * C / 64) * C * 128;   // K-blocks of [C rows, 64]
  static constexpr int RSTD_STRIDE = 144;            // floats per slot (row + 2 is the index: 16 B aligned strip windows)
  static constexpr int OFF_W1 = 0;
  static constexpr int OFF_W2 = OFF_W1 + W1_BYTES;
  static constexpr int OFF_XY = OFF_W2 + W2_BYTES;
  static constexpr int NSTG = (C == 64) ? VV_BT_NSTG : 4;     // output staging boxes, used round-robin box by box
  static constexpr int OBOX_BYTES = 128 * 128;
  static constexpr int OFF_STG = OFF_XY + NX * XY_BYTES;
  static constexpr int OFF_RSTD = OFF_STG + NSTG * OBOX_BYTES;
  static constexpr int OFF_BAR = OFF_RSTD + NX * RSTD_STRIDE * 4;
  static constexpr int D2_FIT = (512 - NG * HC - NG * HC / 2 - NA * (C / 2)) / C;
  static constexpr int ND2 = D2_FIT > 4 ? 4 : D2_FIT;              // D2 accumulators (S3 preloads one while E2 drains another)
  static_assert(ND2 >= 2, "tensor memory: two D2 accumulators");
  static constexpr int NBARS = 1 + 4 * NX + 2 * NSTG + 2 * NA + 4 * NG + 3 * ND2;
  static_assert(OFF_STG % 1024 == 0, "swizzled staging boxes need 1 KB aligned bases");
  static constexpr int TOTAL = OFF_BAR + 512 + 1024;
  static_assert(NBARS * 8 + 8 <= 512, "barrier block");
  static_assert(OFF_XY % 1024 == 0 && XBOX_BYTES % 1024 == 0, "swizzled boxes need 1 KB aligned bases");
  static_assert(TOTAL <= 232448, "exceeds the 227 KB dynamic shared memory limit");
  static constexpr uint32_t TMEM_COLS = 512;         // D1: NG x HC at 0; H: NG x HC/2; D2: ND2 x C; a: NA x C/2
  static constexpr uint32_t H_COL0 = NG * HC, D2_COL0 = NG * HC + NG * HC / 2, A_COL0 = D2_COL0 + ND2 * C;
  static_assert(A_COL0 + NA * (C / 2) <= 512, "tensor memory columns (HC = 128 leaves no room for the operand: use HC = 64)");
  // register budgets (setmaxnreg draws from the CTA's launch allocation: threads x launch registers)
  static constexpr int REG_LAUNCH = (65536 / THREADS) / 8 * 8;
  static constexpr int REG_E1 = (EC == 16) ? 48 : (NG >= 3 ? 56 : 64), REG_E2 = (NG >= 3) ? 64 : 72, REG_S1 = 48, REG_S3 = (C == 64) ? 104 : 64,
                       REG_S2 = (NG >= 3) ? 96 : 112, REG_CTRL = 32;
  static_assert(WARPS % 4 == 0, "registers are allocated per SM sub-partition: every one must hold the same warp mix");
  static_assert(32 * (E1W * REG_E1 + E2W * REG_E2 + S1W * REG_S1 + S3W * REG_S3 + S2W * REG_S2 + CTRL * REG_CTRL) <=
                    THREADS * REG_LAUNCH, "setmaxnreg budget exceeds the CTA's launch allocation");
};

template <int FROM, int TO> __device__ __forceinline__ void reg_set() {
  if constexpr (TO < FROM) reg_dec<TO>();
  else if constexpr (TO > FROM) reg_inc<TO>();
}

// address of the 16-byte cell (row, 4-float chunk q of a 32-channel box) inside a 128B-swizzled [rows, 32 fp32] box
__device__ __forceinline__ uint32_t xy_cell(int row, int q) { return (uint32_t)(row * 128 + ((q ^ (row & 7)) << 4)); }

template <int C, bool STATE>
__global__ void __launch_bounds__(BlockTmaCfg<C>::THREADS, 1)
block_tma_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmO,
                 const __grid_constant__ CUtensorMap tmW1, const __grid_constant__ CUtensorMap tmW2,
                 const BlockTmaArgs p) {
  using L = BlockTmaCfg<C>;
  constexpr int NX = L::NX, NA = L::NA, NCH = L::NCH, NBOX = L::NBOX, HC = L::HC, LPR = L::LPR, R = L::R;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_align_1k(smem_raw);
  uint8_t* sW1 = smem + L::OFF_W1;
  uint8_t* sW2 = smem + L::OFF_W2;
  uint8_t* sXY = smem + L::OFF_XY;
  uint8_t* sStg = smem + L::OFF_STG;
  float* sRstd = (float*)(smem + L::OFF_RSTD);
  uint64_t* bars = (uint64_t*)(smem + L::OFF_BAR);
  uint64_t* w_full = bars;                 // [1]
  uint64_t* x_full = bars + 1;             // [NX] TMA landed the x tile
  uint64_t* st_ready = x_full + NX;        // [NX] S1 wrote rstd[] (and patched / saved the stream context)
  uint64_t* y_full = st_ready + NX;        // [NX] S2 wrote y in place
  uint64_t* x_free = y_full + NX;          // [NX] S3 holds the slot's y rows in registers: the loader may refill it
  constexpr int NSTG = L::NSTG;
  uint64_t* out_ready = x_free + NX;       // [NSTG] E2 wrote a staging box
  uint64_t* out_free = out_ready + NSTG;   // [NSTG] the TMA store has read it
  uint64_t* a_full = out_free + NSTG;      // [NA] S3 wrote the a tile
  uint64_t* a_empty = a_full + NA;         // [NA] GEMM 1 has read it
  constexpr int NG = L::NG;
  uint64_t* d1_full = a_empty + NA;        // [NG]
  uint64_t* d1_empty = d1_full + NG;       // [NG]
  uint64_t* h_full = d1_empty + NG;        // [NG]
  uint64_t* h_empty = h_full + NG;         // [NG]
  constexpr int ND2 = L::ND2;
  uint64_t* d2_full = h_empty + NG;        // [ND2] GEMM 2 finished the tile's accumulator
  uint64_t* d2_empty = d2_full + ND2;      // [ND2] E2 has read it
  uint64_t* d2_init = d2_empty + ND2;      // [ND2] S3 preloaded it with y + gb2
  uint32_t* tmem_slot = (uint32_t*)(d2_init + ND2);

  // hardware warps: [0, 8) E1, [8, 12) E2, [12, 16) S1, [16, 20) S3, [20, 24) S2, [24, 28): loader, G1, G2, store
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  constexpr int WARP_E2 = L::E1W, WARP_S1 = WARP_E2 + L::E2W, WARP_S3 = WARP_S1 + L::S1W, WARP_S2 = WARP_S3 + L::S3W,
                WARP_CTRL = WARP_S2 + L::S2W;
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
    mbar_init(w_full, 1);
    for (int i = 0; i < NX; ++i) {
      mbar_init(&x_full[i], 1);
      mbar_init(&st_ready[i], L::S1W);
      mbar_init(&y_full[i], L::S2W);
      mbar_init(&x_free[i], L::S3W);
    }
    for (int i = 0; i < NSTG; ++i) { mbar_init(&out_ready[i], L::E2W); mbar_init(&out_free[i], 1); }
    for (int i = 0; i < NA; ++i) { mbar_init(&a_full[i], L::S3W); mbar_init(&a_empty[i], 1); }
    for (int i = 0; i < NG; ++i) {
      mbar_init(&d1_full[i], 1);
      mbar_init(&d1_empty[i], 4);
      mbar_init(&h_full[i], 4);
      mbar_init(&h_empty[i], 1);
    }
    for (int i = 0; i < ND2; ++i) {
      mbar_init(&d2_full[i], 1);
      mbar_init(&d2_empty[i], L::E2W);
      mbar_init(&d2_init[i], L::S3W);
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
  // weights are constants (written at finalize): their loads go out before this CTA waits for the previous kernel
  if (warp == WARP_CTRL && elect_one()) {
    mbar_expect_tx(w_full, L::W1_BYTES + L::W2_BYTES);
#pragma unroll
    for (int j = 0; j < 4 * C / 128; ++j) tma_load_2d(sW1 + j * 128 * L::ROWB, &tmW1, w_full, 0, j * 128);
#pragma unroll
    for (int kb = 0; kb < 4 * C / 64; ++kb) tma_load_2d(sW2 + kb * (C * 128), &tmW2, w_full, kb * 64, 0);
  }
  pdl_wait();

  if (warp >= WARP_CTRL) {
    reg_set<L::REG_LAUNCH, L::REG_CTRL>();
    const int role = warp - WARP_CTRL;
    if (role == 0) {
      // ===== x loader: one [XROWS, 32] box per 32 channels, rows t0 - 8 .. t0 + 127 (rows outside the item read as zero)
      if (elect_one()) {
        for (int it = 0; it < my_tiles; ++it) {
          const int s = it % NX;
          int b, t0;
          tile_of(it, b, t0);
          mbar_wait_sleep(&x_free[s], (uint32_t)(((it / NX) & 1) ^ 1), 64);
          VV_BWT(0, it);
          mbar_expect_tx(&x_full[s], L::XY_BYTES);
#pragma unroll
          for (int c = 0; c < NBOX; ++c)
            tma_load_3d(sXY + s * L::XY_BYTES + c * L::XBOX_BYTES, &tmX, &x_full[s], c * 32, t0 - L::HALO, b);
        }
      }
    } else if (role == 1) {
      // ===== GEMM 1 issuer: D1[g & 1] = a * W1_chunk^T ========================================================
      if (elect_one()) {
        constexpr uint32_t idesc1 = umma_idesc_bf16(128, HC);
        const uint64_t dw0 = umma_smem_desc<L::ROWB>(smem_u32(sW1));
        mbar_wait(w_full, 0);
        int g = 0;
        for (int it = 0; it < my_tiles; ++it) {
          const int sa = it % NA;
          mbar_wait_sleep(&a_full[sa], (uint32_t)((it / NA) & 1), 64);
          VV_BWT(8, it);
          const uint32_t aa = tmem_base + L::A_COL0 + (uint32_t)(sa * (C / 2));   // A operand from tensor memory
#pragma unroll 1
          for (int j = 0; j < NCH; ++j, ++g) {
            const int db = g % NG;
            mbar_wait(&d1_empty[db], (uint32_t)(((g / NG) & 1) ^ 1));
            tc_fence_after();
            const uint64_t dw = dw0 + (uint64_t)((j * HC * L::ROWB) >> 4);
            const uint32_t d = tmem_base + (uint32_t)(db * HC);
#pragma unroll
            for (int kk = 0; kk < C / 16; ++kk)
              umma_bf16_ts(d, aa + (uint32_t)(kk * 8), dw + (uint64_t)(kk * 2), idesc1, kk != 0);
            umma_commit(&d1_full[db]);
          }
          umma_commit(&a_empty[sa]);
          VV_BWT(9, it);
        }
      }
    } else if (role == 2) {
      // ===== GEMM 2 issuer: D2[it & 1] += H * W2_chunk^T (A operand from tensor memory) ==========================
      if (elect_one()) {
        constexpr uint32_t idesc2 = umma_idesc_bf16(128, C);
        const uint64_t dw0 = umma_smem_desc<128>(smem_u32(sW2));
        mbar_wait(w_full, 0);
        int g = 0;
        for (int it = 0; it < my_tiles; ++it) {
          const int d2b = it % ND2;
          mbar_wait_sleep(&d2_init[d2b], (uint32_t)((it / ND2) & 1), 64);     // the accumulator holds y + gb2
          tc_fence_after();
          const uint32_t d = tmem_base + L::D2_COL0 + (uint32_t)(d2b * C);
#pragma unroll 1
          for (int j = 0; j < NCH; ++j, ++g) {
            const int hb = g % NG;
            mbar_wait(&h_full[hb], (uint32_t)((g / NG) & 1));
            tc_fence_after();
            const uint32_t ah = tmem_base + L::H_COL0 + (uint32_t)(hb * (HC / 2));
#pragma unroll
            for (int kk = 0; kk < HC / 16; ++kk) {
              // hidden units [j HC + 16 kk, +16): K-block (j HC + 16 kk) / 64 of W2, 16-element step (kk & 3) inside it
              const uint64_t dw = dw0 + (uint64_t)((((j * HC + kk * 16) >> 6) * C * 128) >> 4) + (uint64_t)((kk & 3) * 2);
              umma_bf16_ts(d, ah + (uint32_t)(kk * 8), dw, idesc2, 1u);
            }
            umma_commit(&h_empty[hb]);
          }
          umma_commit(&d2_full[d2b]);
          VV_BWT(10, it);
        }
      }
    } else {
      // ===== TMA store issuer: finished staging boxes -> global, box by box ===========================================
      if (elect_one()) {
        int k = 0;
        for (int it = 0; it < my_tiles; ++it) {
          int b, t0;
          tile_of(it, b, t0);
#pragma unroll 1
          for (int c = 0; c < NBOX; ++c, ++k) {
            const int o = k % NSTG;
            mbar_wait_sleep(&out_ready[o], (uint32_t)((k / NSTG) & 1), 64);
            if (c == 0) VV_BWT(13, it);
            tma_store_3d(&tmO, sStg + o * L::OBOX_BYTES, c * 32, t0, b);
            tma_store_commit();
            tma_store_wait_read<0>();
            mbar_arrive(&out_free[o]);
          }
          VV_BWT(14, it);
        }
        tma_store_wait_all<0>();
      }
    }
  } else if (warp < WARP_E2) {
    // ===== E1: D1 -> + b1 -> GELU -> bf16 pairs -> H in tensor memory ===========================================
    // NG groups of 4 warps take chunks round-robin (group G owns D1 buffer G and H buffer G); a warp owns 32 rows and walks
    // the chunk's 64 columns EC at a time.
    reg_set<L::REG_LAUNCH, L::REG_E1>();
    constexpr int EC = L::EC;
    const int q = warp & 3;
    const int grp = warp >> 2;
    const uint32_t lane_base = (uint32_t)(q * 32) << 16;
    const uint32_t d1addr = tmem_base + lane_base + (uint32_t)(grp * HC);
    const uint32_t haddr = tmem_base + lane_base + L::H_COL0 + (uint32_t)(grp * (HC / 2));
    for (int g = grp; g < total_g; g += NG) {
      const int j = g % NCH;
      const uint32_t ph = (uint32_t)((g / NG) & 1);
      const float* bp = p.b1c + j * HC;                  // constant bank (by-value kernel argument)
      if (threadIdx.x == 0) VV_BWT(16, g / NCH);          // (the last chunk of the tile that group 0 handles wins)
      mbar_wait(&d1_full[grp], ph);
      if (threadIdx.x == 0) VV_BWT(17, g / NCH);
      tc_fence_after();
#pragma unroll
      for (int hf = 0; hf < HC / EC; ++hf) {
        uint32_t v[EC];
        if constexpr (EC == 32) tmem_ld_32x32(d1addr + (uint32_t)(hf * EC), v);
        else tmem_ld_32x16(d1addr + (uint32_t)(hf * EC), v);
        tmem_ld_wait();
        if (hf == HC / EC - 1) {                     // whole accumulator is in registers: the next G1 may overwrite it
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&d1_empty[grp]);
        }
        uint32_t hv[EC / 2];
#pragma unroll
        for (int i = 0; i < EC / 8; ++i) {
          const float* bq = bp + hf * EC + i * 8;
          hv[i * 4 + 0] = gelu2_bias_pack(__uint_as_float(v[i * 8 + 0]), __uint_as_float(v[i * 8 + 1]), bq[0], bq[1]);
          hv[i * 4 + 1] = gelu2_bias_pack(__uint_as_float(v[i * 8 + 2]), __uint_as_float(v[i * 8 + 3]), bq[2], bq[3]);
          hv[i * 4 + 2] = gelu2_bias_pack(__uint_as_float(v[i * 8 + 4]), __uint_as_float(v[i * 8 + 5]), bq[4], bq[5]);
          hv[i * 4 + 3] = gelu2_bias_pack(__uint_as_float(v[i * 8 + 6]), __uint_as_float(v[i * 8 + 7]), bq[6], bq[7]);
        }
        if (hf == 0) {
          if (threadIdx.x == 0) VV_BWT(18, g / NCH);
          mbar_wait(&h_empty[grp], ph ^ 1);          // GEMM 2 has consumed this group's previous H chunk
          if (threadIdx.x == 0) VV_BWT(19, g / NCH);
          tc_fence_after();
        }
        // thread = row: pair (k, k+1) -> one 32-bit column
        if constexpr (EC == 32) tmem_st_32x16(haddr + (uint32_t)(hf * (EC / 2)), hv);
        else tmem_st_32x8(haddr + (uint32_t)(hf * (EC / 2)), hv);
      }
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&h_full[grp]);
      if (threadIdx.x == 0) VV_BWT(22, g / NCH);
    }
  } else if (warp < WARP_S1) {
    // ===== E2: the accumulator already holds out = (y + gb2) + H W2^T: tcgen05.ld -> staging box (thread = row) -> TMA store
    reg_set<L::REG_LAUNCH, L::REG_E2>();
    const int q = warp & 3;
    const int r = q * 32 + lane;
    int k = 0;
    for (int it = 0; it < my_tiles; ++it) {
      const int d2b = it % ND2;
      int b, t0;
      tile_of(it, b, t0);
      const int t = t0 + r;
      const bool valid = t < p.rows_pb;
      const long long brow = (long long)b * p.ob_rpb + p.ob_off + t;
      mbar_wait(&d2_full[d2b], (uint32_t)((it / ND2) & 1));
      if (threadIdx.x == WARP_E2 * 32) VV_BWT(11, it);
      tc_fence_after();
#pragma unroll 1
      for (int c = 0; c < NBOX; ++c, ++k) {
        const int o = k % NSTG;
        uint32_t v[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + L::D2_COL0 + (uint32_t)(d2b * C + c * 32), v);
        mbar_wait(&out_free[o], (uint32_t)(((k / NSTG) & 1) ^ 1));        // the store of the box's previous tenant has read it
        tmem_ld_wait();
        uint8_t* box = sStg + o * L::OBOX_BYTES;
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
        if (c == NBOX - 1) tc_fence_before();
        fence_proxy_async_smem();               // staged box -> visible to the TMA store
        __syncwarp();
        if (lane == 0) {
          if (c == NBOX - 1) mbar_arrive(&d2_empty[d2b]);
          mbar_arrive(&out_ready[o]);
        }
      }
      if (threadIdx.x == WARP_E2 * 32) VV_BWT(12, it);
    }
  } else if (warp < WARP_S3) {
    // ===== S1: thread = row: rstd of the raw row (first RMSNorm), stream context in / out ===========================
    reg_set<L::REG_LAUNCH, L::REG_S1>();
    const int r = (warp - WARP_S1) * 32 + lane;           // tile row 0..127 -> slot row HALO + r
    const float inv_c = 1.0f / (float)C;
    auto row_rstd = [&](const uint8_t* slot, int srow) -> float {
      f32x2 acc[4] = {0ull, 0ull, 0ull, 0ull};            // four independent chains
#pragma unroll
      for (int c = 0; c < NBOX; ++c) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const ulonglong2 v = *reinterpret_cast<const ulonglong2*>(slot + c * L::XBOX_BYTES + xy_cell(srow, i));
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
    for (int it = 0; it < my_tiles; ++it) {
      const int s = it % NX;
      int b, t0;
      tile_of(it, b, t0);
      uint8_t* slot = sXY + s * L::XY_BYTES;
      float* rs = sRstd + s * L::RSTD_STRIDE;
      // the stream's state pointers (two dependent global loads) are only needed by tiles that read the context (item start)
      // or save it (the item's last 6 rows); they are fetched before the wait for the x tile
      const uint8_t* rd8 = nullptr;
      uint8_t* wr8 = nullptr;
      if (STATE && p.st.base != nullptr && (t0 == 0 || t0 + 128 > p.rows_pb - 6)) p.st.resolve(b, rd8, wr8);
      mbar_wait(&x_full[s], (uint32_t)((it / NX) & 1));
      if (threadIdx.x == WARP_S1 * 32) VV_BWT(1, it);
      if (warp == WARP_S1 && lane < L::HALO) {
        // halo rows (slot rows 0..7 = t0 - 8 .. t0 - 1): at the item start rows 2..7 are the stream's cached raw rows
        // (TMA delivered zeros = the offline causal pad, :398-403)
        if (STATE && t0 == 0 && rd8 != nullptr && lane >= 2) {
          const float* src = reinterpret_cast<const float*>(rd8) + (lane - 2) * C;
#pragma unroll
          for (int c = 0; c < NBOX; ++c)
#pragma unroll
            for (int i = 0; i < 8; ++i)
              *reinterpret_cast<float4*>(slot + c * L::XBOX_BYTES + xy_cell(lane, i)) =
                  *reinterpret_cast<const float4*>(src + c * 32 + i * 4);
        }
        rs[lane + 2] = row_rstd(slot, lane);
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
                    *reinterpret_cast<const float4*>(slot + c * L::XBOX_BYTES + xy_cell(lane, i));
          }
        }
      }
      rs[L::HALO + r + 2] = row_rstd(slot, L::HALO + r);
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
                  *reinterpret_cast<const float4*>(slot + c * L::XBOX_BYTES + xy_cell(L::HALO + r, i));
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&st_ready[s]);
      if (threadIdx.x == WARP_S1 * 32) VV_BWT(2, it);
    }
  } else if (warp < WARP_S2) {
    // ===== S3: thread = row: second RMSNorm statistic, a = bf16(y * rstd2) -> tensor memory, D2 preload ==================
    reg_set<L::REG_LAUNCH, L::REG_S3>();
    const int r = (warp - WARP_S3) * 32 + lane;
    const float inv_c = 1.0f / (float)C;
    for (int it = 0; it < my_tiles; ++it) {
      const int s = it % NX, sa = it % NA;
      const uint8_t* slot = sXY + s * L::XY_BYTES;
      mbar_wait(&y_full[s], (uint32_t)((it / NX) & 1));
      if (threadIdx.x == WARP_S3 * 32) VV_BWT(5, it);
      ulonglong2 y[C / 4];
      f32x2 acc[4] = {0ull, 0ull, 0ull, 0ull};
#pragma unroll
      for (int c = 0; c < NBOX; ++c)
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          y[c * 8 + i] = *reinterpret_cast<const ulonglong2*>(slot + c * L::XBOX_BYTES + xy_cell(L::HALO + r, i));
          acc[(i & 1) * 2] = f2_fma(y[c * 8 + i].x, y[c * 8 + i].x, acc[(i & 1) * 2]);
          acc[(i & 1) * 2 + 1] = f2_fma(y[c * 8 + i].y, y[c * 8 + i].y, acc[(i & 1) * 2 + 1]);
        }
      float s0, s1, s2, s3, s4, s5, s6, s7;
      f2_unpack(acc[0], s0, s1);
      f2_unpack(acc[1], s2, s3);
      f2_unpack(acc[2], s4, s5);
      f2_unpack(acc[3], s6, s7);
      const float rstd2 = rstd_of(((s0 + s1) + (s2 + s3)) + ((s4 + s5) + (s6 + s7)), inv_c, p.eps);
      const f32x2 rr = f2_pack(rstd2, rstd2);
      // a = bf16(y * rstd2) -> tensor memory (thread = row = TMEM lane; pair (k, k+1) -> one 32-bit column): GEMM 1 takes its
      // A operand from there, so the operand tile costs no shared memory, no swizzled stores and no proxy fence
      mbar_wait(&a_empty[sa], (uint32_t)(((it / NA) & 1) ^ 1));
      tc_fence_after();
      const uint32_t aaddr = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + L::A_COL0 + (uint32_t)(sa * (C / 2));
#pragma unroll
      for (int hcol = 0; hcol < C / 32; ++hcol) {        // 16 packed columns = 32 channels = 8 cells per tcgen05.st
        uint32_t w[16];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          float a0, a1, a2, a3;
          f2_unpack(f2_mul(y[hcol * 8 + i].x, rr), a0, a1);
          f2_unpack(f2_mul(y[hcol * 8 + i].y, rr), a2, a3);
          w[i * 2 + 0] = pack_bf16x2(a0, a1);
          w[i * 2 + 1] = pack_bf16x2(a2, a3);
        }
        tmem_st_32x16(aaddr + (uint32_t)(hcol * 16), w);
      }
      tmem_st_wait();                           // GEMM 1 must not wait for the D2 preload below (which may have to wait for E2)
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&a_full[sa]);
        mbar_arrive(&x_free[s]);                // this warp's y rows are in registers: nothing reads the slot after S3
      }
      if (threadIdx.x == WARP_S3 * 32) VV_BWT(6, it);
      // D2[it % ND2] = y + gb2 (thread = row = TMEM lane; this warp's lane quarter is warp % 4): GEMM 2 accumulates on it
      const int d2b = it % ND2;
      mbar_wait(&d2_empty[d2b], (uint32_t)(((it / ND2) & 1) ^ 1));
      tc_fence_after();
      const uint32_t daddr = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + L::D2_COL0 + (uint32_t)(d2b * C);
#pragma unroll
      for (int hcol = 0; hcol < C / 16; ++hcol) {        // 16 columns = 4 cells of 4 floats per tcgen05.st
        uint32_t w[16];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int cell = hcol * 4 + i;                 // channels [4 cell, 4 cell + 4)
          float a0, a1, a2, a3;
          f2_unpack(y[cell].x, a0, a1);
          f2_unpack(y[cell].y, a2, a3);
          w[i * 4 + 0] = __float_as_uint(a0 + p.gb2c[cell * 4 + 0]);
          w[i * 4 + 1] = __float_as_uint(a1 + p.gb2c[cell * 4 + 1]);
          w[i * 4 + 2] = __float_as_uint(a2 + p.gb2c[cell * 4 + 2]);
          w[i * 4 + 3] = __float_as_uint(a3 + p.gb2c[cell * 4 + 3]);
        }
        tmem_st_32x16(daddr + (uint32_t)(hcol * 16), w);
      }
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&d2_init[d2b]);
      if (threadIdx.x == WARP_S3 * 32) VV_BWT(7, it);
    }
  } else {
    // ===== S2: depthwise causal conv k7 over the normalised rows; lane = 4 channels, strip = R consecutive rows =======
    reg_set<L::REG_LAUNCH, L::REG_S2>();
    struct P4 { f32x2 lo, hi; };
    const int gi = (warp - WARP_S2) * (32 / LPR) + lane / LPR;      // strip 0..STRIPS-1: tile rows [R gi, R gi + R)
    const int c0 = (lane % LPR) * 4;
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
    const int sr0 = L::HALO + gi * R;                               // slot row of the strip's first output row
    for (int it = 0; it < my_tiles; ++it) {
      const int s = it % NX;
      uint8_t* bx = sXY + s * L::XY_BYTES + box * L::XBOX_BYTES;
      const float* rs = sRstd + s * L::RSTD_STRIDE + sr0 - 4;      // rs[i] = rstd of slot row sr0 - 6 + i  (16 B aligned)
      mbar_wait(&st_ready[s], (uint32_t)((it / NX) & 1));
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
      // rows are LOADED NB at a time: a load may not be moved above the previous row's in-place store (same buffer), so a
      // row-by-row loop serialises load -> math -> store per row (clock64 timeline of block_ws_kernel: 4 500 clocks per tile)
      constexpr int NB = (L::NG >= 3) ? 4 : 8;
#pragma unroll
      for (int k0 = 0; k0 < R; k0 += NB) {
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
      if (lane == 0) mbar_arrive(&y_full[s]);
      if (threadIdx.x == WARP_S2 * 32) VV_BWT(4, it);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == WARP_CTRL + 1) tmem_dealloc(tmem_base, L::TMEM_COLS);
}

}  // namespace vv
