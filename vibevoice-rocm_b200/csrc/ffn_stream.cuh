// ffn_stream.cuh — kernel class (b) for C = 128 (when block_ws is off) and C = 256: the whole Block1D FFN half in ONE
// persistent tcgen05 kernel with the hidden activation kept on chip and the weights STREAMED through shared memory.
// (C = 256, added after the clock64 timelines of the whole-block kernels: the two-GEMM form there moves 26 bytes per
// element and its second GEMM runs at 90 % of HBM speed on the hidden tensor's round trip; here the tile's budget is one
// 64 KB a tile, two 32 KB stages per weight ring, two staging boxes and ONE 256-column D2 accumulator.)
//
//   out = resid + ffn_gamma * ( W2 * gelu_erf( W1 * a + b1 ) + b2 )         (reference :592-596, :937-942)
//
// At C = 128 the two-GEMM form is HBM-bound by the [rows, 4C] hidden tensor the reference materialises
// (FFN.forward :594-596): 16 B per activation element out and back in, against 10 B for everything else.  W1 and
// W2 (128 KB each) do not fit in shared memory next to the tiles, so per 128-row tile the hidden dimension is
// walked in NCH chunks of HC = 64 units and the weight chunks arrive through two TMA rings (they are L2-resident:
// 256 KB per block):
//   role 0  TMA : W1 chunk [64, C] (ring S1); W2 chunk [C, 64] (ring S2)
//   role 1  G1  : D1[128, 64] = a * W1_chunk^T                  (tcgen05.mma, 2 accumulators in TMEM)
//   role 2  G2  : D2[128, C] += H * W2_chunk^T                  (tcgen05.mma; 2 accumulators at C = 128, 1 at C = 256)
//   role 3  TMA : residual boxes [128 rows, 32 fp32] in (ring NY), finished boxes out; the reload of a slot is
//                 issued by the same thread right after its store has read the slot.  In-place mode (out == resid):
//                 nothing is loaded, a finished box is ADDED to the residual in global memory by a TMA reduction
//   role 4  TMA : a tile [128, C] bf16 (ring of NA; ONE slot at C = 256: its own thread, so that waiting for GEMM 1 to
//                 release the slot cannot stall the weight rings)
//   E1, 2 x 8 warps alternating chunks: tcgen05.ld D1 -> + b1 -> GELU -> bf16 pairs -> tcgen05.st H (TENSOR memory,
//       2 buffers): GEMM 2 reads its A operand from TMEM, so H costs no shared memory, no swizzled stores and no
//       generic->async proxy fence (clock64 timeline: the st.shared + fence + arrive tail was ~500 cycles per chunk)
//   E2, 4 warps, box by box: tcgen05.ld D2 -> + b2, * gamma, + residual (in the staged box, in place)
// Two separate issuing threads for the two GEMMs: measured on B200, ONE thread issuing 12 MMAs + 6 barrier
// waits + 6 commits per chunk was the bottleneck of this kernel (0.8 us per chunk with every other stage removed;
// the descriptor and barrier bookkeeping of a single warp issues at ~6 cycles per instruction).  All hand-offs are
// mbarriers; there is no __syncthreads in the loop.  Loop counters are 32-bit and all ring sizes powers of two.
#pragma once
#include <cuda.h>
#include "ptx.cuh"
#include "epilogue.cuh"
#include "ffn_fused.cuh"   // FfnFusedArgs

#ifndef VV_FS_CONST
#define VV_FS_CONST 1     // 1: b1 / b2 / ffn_gamma from the constant bank (by-value kernel argument); 0: plain global loads
#endif
namespace vv {

#ifdef VV_FS_TRACE     // timeline instrumentation (A/B builds only): clock64 stamps of CTA 0 for chunks [TR_G0, TR_G0+32)
__device__ long long vv_fs_trace[32 * 16];
#define VV_TR(slot, g)                                                                              \
  do {                                                                                              \
    if (blockIdx.x == 0 && (g) >= 256 && (g) < 288) vv_fs_trace[((g) - 256) * 16 + (slot)] = clock64(); \
  } while (0)
#else
#define VV_TR(slot, g) do { } while (0)
#endif

#ifndef VV_FS_GROUPS
#define VV_FS_GROUPS 2      // E1 groups taking chunks round-robin: 2 (8 warps each, 32 columns per warp) or 4 (4 warps
#endif                      // each, the whole 64-column chunk per warp).  Measured at C = 128, 5760 frames: 1.45 ms with 2,
                            // 1.70 ms with 4 — a warp's own 64-column walk costs more than the extra overlap returns
constexpr int FS_GROUPS = VV_FS_GROUPS;
constexpr int FS_E1_WARPS = 16;
constexpr int FS_E2_WARPS = 4;
constexpr int FS_CTRL = 8;                         // control warps: weights, G1, G2, residual / store ring, a tiles (+ 3 idle: the
constexpr int FS_THREADS = (FS_CTRL + FS_E1_WARPS + FS_E2_WARPS) * 32;   // warp count stays a multiple of 4)

template <int C>
struct FfnStreamCfg {
  static_assert(C == 128 || C == 256, "ffn_stream_kernel is laid out for C = 128 / 256");
  static constexpr int HC = 64;                    // hidden units per chunk
  static constexpr int NCH = 4 * C / HC;           // chunks per tile
  static constexpr int KB = C / 64;                // 64-wide K blocks of the a tile / a W1 chunk
  static constexpr int NBOX = C / 32;              // residual / output boxes per tile
  static constexpr int NA = (C == 128) ? 2 : 1, S1 = (C == 128) ? 4 : 2, S2 = 2, NY = (C == 128) ? 4 : 2, ND1 = 2;
  static constexpr int ND2 = (C == 128) ? 2 : 1;   // D2 accumulators of C columns at tensor-memory column 256
  static constexpr int WLAG = 2;                   // the producer issues W2 chunk g - WLAG after W1 chunk g
  static constexpr int A_BYTES = 128 * C * 2;      // KB blocks of [128 rows x 128 B]
  static constexpr int W1_BYTES = HC * C * 2;      // KB blocks of [64 rows x 128 B]
  static constexpr int W2_BYTES = C * HC * 2;      // [C rows x 128 B]
  static constexpr int H_BYTES = 128 * HC * 2;     // [128 rows x 128 B]
  static constexpr int Y_BYTES = 128 * 128;        // [128 rows x 32 fp32]
  static constexpr int OFF_A = 0;
  static constexpr int OFF_W1 = OFF_A + NA * A_BYTES;
  static constexpr int OFF_W2 = OFF_W1 + S1 * W1_BYTES;
  static constexpr int OFF_H = OFF_W2 + S2 * W2_BYTES;
  static constexpr int OFF_Y = OFF_H;              // H lives in tensor memory (no shared-memory copy)
  static constexpr int OFF_BAR = OFF_Y + NY * Y_BYTES;
  static constexpr int NBARS = 2 * NA + 2 * S1 + 2 * S2 + ND1 + 3 * FS_GROUPS + 4 + 2 * NY;
  static_assert(256 + ND2 * C <= 512, "tensor memory columns");
  static constexpr int TOTAL = OFF_BAR + 512 + 1024;   // + alignment slack
  static_assert(NBARS * 8 + 8 <= 512, "barrier block");
  static_assert(TOTAL <= 232448, "exceeds the 227 KB dynamic shared memory limit");
  static constexpr uint32_t TMEM_COLS = 512;       // D1: 2 x 64 at column 0; H: FS_GROUPS x 32 at 128; D2: 2 x C at 256
  static_assert(FS_GROUPS == 2 || FS_GROUPS == 4, "E1 groups");
  static constexpr uint32_t H_COL0 = 128;          // H chunk as packed bf16 pairs: 64 hidden units = 32 columns
  static constexpr uint32_t D2_COL0 = 256;
};

template <int C>
__global__ void __launch_bounds__(FS_THREADS, 1)
ffn_stream_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmR,
                  const __grid_constant__ CUtensorMap tmO, const __grid_constant__ CUtensorMap tmW1,
                  const __grid_constant__ CUtensorMap tmW2, const FfnFusedArgs p) {
  using L = FfnStreamCfg<C>;
  constexpr int NA = L::NA, S1 = L::S1, S2 = L::S2, NY = L::NY, NCH = L::NCH, KB = L::KB, NBOX = L::NBOX;
  constexpr int ND1 = L::ND1, ND2 = L::ND2;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_align_1k(smem_raw);
  uint8_t* sA = smem + L::OFF_A;
  uint8_t* sW1 = smem + L::OFF_W1;
  uint8_t* sW2 = smem + L::OFF_W2;
  uint8_t* sY = smem + L::OFF_Y;
  uint64_t* bars = (uint64_t*)(smem + L::OFF_BAR);
  uint64_t* a_full = bars;                 // [NA]
  uint64_t* a_empty = a_full + NA;         // [NA]
  uint64_t* w1_full = a_empty + NA;        // [S1]
  uint64_t* w1_empty = w1_full + S1;       // [S1]
  uint64_t* w2_full = w1_empty + S1;       // [S2]
  uint64_t* w2_empty = w2_full + S2;       // [S2]
  // d1_full is per E1 GROUP, not per accumulator: a waiter must see the phases of its barrier in order, and with more
  // groups than accumulators a group would otherwise first wait on a later phase of a buffer it never used (parity alias)
  uint64_t* d1_full = w2_empty + S2;       // [FS_GROUPS]
  uint64_t* d1_empty = d1_full + FS_GROUPS; // [ND1]
  uint64_t* h_full = d1_empty + ND1;       // [FS_GROUPS]
  uint64_t* h_empty = h_full + FS_GROUPS;  // [FS_GROUPS]
  uint64_t* d2_full = h_empty + FS_GROUPS; // [2]
  uint64_t* d2_empty = d2_full + 2;        // [2]
  uint64_t* y_full = d2_empty + 2;         // [NY] residual box landed
  uint64_t* y_ready = y_full + NY;         // [NY] E2 finished the box: ready to store
  uint32_t* tmem_slot = (uint32_t*)(y_ready + NY);

  // Role index: the four single-thread roles (TMA, G1, G2, residual ring) sit in the HIGHEST warps of the CTA.
  // The SM's warp arbiter favours higher warp ids; with the issuers in warps 0..3 they lost most issue slots to
  // the E1 warps of their scheduler and needed ~950 cycles per chunk for ~100 instructions (clock64 timeline).
  const int hw_warp = threadIdx.x >> 5;
  const int warp = (hw_warp >= FS_E1_WARPS + FS_E2_WARPS) ? hw_warp - (FS_E1_WARPS + FS_E2_WARPS) : hw_warp + FS_CTRL;
  const int lane = threadIdx.x & 31;
  const long long ntiles = (p.R + 127) / 128;
  const int my_tiles = (ntiles > blockIdx.x) ? (int)((ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x) : 0;
  const int total_g = my_tiles * NCH;
  const int tile_step = (int)gridDim.x * 128;      // row distance between this CTA's consecutive tiles
  const int row0 = (int)blockIdx.x * 128;          // (rows fit 31 bits: the host checks R < 2^31)

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmR);
    tma_prefetch_desc(&tmO);
    tma_prefetch_desc(&tmW1);
    tma_prefetch_desc(&tmW2);
    for (int i = 0; i < NA; ++i) { mbar_init(&a_full[i], 1); mbar_init(&a_empty[i], 1); }
    for (int i = 0; i < S1; ++i) { mbar_init(&w1_full[i], 1); mbar_init(&w1_empty[i], 1); }
    for (int i = 0; i < S2; ++i) { mbar_init(&w2_full[i], 1); mbar_init(&w2_empty[i], 1); }
    for (int i = 0; i < ND1; ++i) mbar_init(&d1_empty[i], FS_E1_WARPS / FS_GROUPS);
    for (int i = 0; i < FS_GROUPS; ++i) {
      mbar_init(&d1_full[i], 1);
      mbar_init(&h_full[i], FS_E1_WARPS / FS_GROUPS);
      mbar_init(&h_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&d2_full[i], 1);
      mbar_init(&d2_empty[i], FS_E2_WARPS);
    }
    for (int i = 0; i < NY; ++i) { mbar_init(&y_full[i], 1); mbar_init(&y_ready[i], FS_E2_WARPS); }
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, L::TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch();
  pdl_wait();

  if (warp == 0) {
    // ===== TMA producer: weight chunks (the a tiles have their own thread, role 4) ============================
    // W2 chunks trail W1 by WLAG chunks (GEMM 2 runs behind GEMM 1); the single in-order producer cannot
    // deadlock: every wait below depends only on loads issued earlier in this same order.
    if (elect_one()) {
      auto load_w2 = [&](int gp) {
        const int s2 = gp & (S2 - 1);
        mbar_wait(&w2_empty[s2], (uint32_t)(((gp / S2) & 1) ^ 1));
        mbar_expect_tx(&w2_full[s2], L::W2_BYTES);
        tma_load_2d(sW2 + s2 * L::W2_BYTES, &tmW2, &w2_full[s2], (gp & (NCH - 1)) * L::HC, 0);
      };
      int g = 0;
      for (int it = 0; it < my_tiles; ++it) {
        for (int j = 0; j < NCH; ++j, ++g) {
          const int s1 = g & (S1 - 1);
          mbar_wait(&w1_empty[s1], (uint32_t)(((g / S1) & 1) ^ 1));
          mbar_expect_tx(&w1_full[s1], L::W1_BYTES);
#pragma unroll
          for (int kb = 0; kb < KB; ++kb)
            tma_load_2d(sW1 + s1 * L::W1_BYTES + kb * (L::HC * 128), &tmW1, &w1_full[s1], kb * 64, j * L::HC);
          if (g >= L::WLAG) load_w2(g - L::WLAG);
        }
      }
      for (int gp = (total_g > L::WLAG ? total_g - L::WLAG : 0); gp < total_g; ++gp) load_w2(gp);
    }
  } else if (warp == 1) {
    // ===== GEMM 1 issuer: D1[g % ND1] = a * W1_chunk^T ==================================================
    if (elect_one()) {
      constexpr uint32_t idesc1 = umma_idesc_bf16(128, L::HC);
      const uint64_t da0 = umma_smem_desc<128>(smem_u32(sA));
      const uint64_t dw0 = umma_smem_desc<128>(smem_u32(sW1));
      int g = 0;
      for (int it = 0; it < my_tiles; ++it) {
        const int sa = it & (NA - 1);
        mbar_wait(&a_full[sa], (uint32_t)((it / NA) & 1));
        const uint64_t da = da0 + (uint64_t)((sa * L::A_BYTES) >> 4);
        for (int j = 0; j < NCH; ++j, ++g) {
          const int db = g & (ND1 - 1), s1 = g & (S1 - 1);
          mbar_wait2(&d1_empty[db], (uint32_t)(((g / ND1) & 1) ^ 1), &w1_full[s1], (uint32_t)((g / S1) & 1));
          VV_TR(1, g);
          tc_fence_after();
          const uint64_t dw = dw0 + (uint64_t)((s1 * L::W1_BYTES) >> 4);
          const uint32_t d = tmem_base + (uint32_t)(db * L::HC);
#pragma unroll
          for (int kb = 0; kb < KB; ++kb) {
#pragma unroll
            for (int kk = 0; kk < 4; ++kk)
              umma_bf16(d, da + (uint64_t)(((kb * 128 * 128) >> 4) + kk * 2),
                        dw + (uint64_t)(((kb * L::HC * 128) >> 4) + kk * 2), idesc1, (kb | kk) != 0);
          }
          umma_commit(&d1_full[g & (FS_GROUPS - 1)]);
          umma_commit(&w1_empty[s1]);
          VV_TR(2, g);
        }
        umma_commit(&a_empty[sa]);
      }
    }
  } else if (warp == 2) {
    // ===== GEMM 2 issuer: D2[it & 1] += H * W2_chunk^T ==================================================
    if (elect_one()) {
      constexpr uint32_t idesc2 = umma_idesc_bf16(128, C);
      const uint64_t dw0 = umma_smem_desc<128>(smem_u32(sW2));
      int g = 0;
      for (int it = 0; it < my_tiles; ++it) {
        const int d2b = it % ND2;
        mbar_wait(&d2_empty[d2b], (uint32_t)(((it / ND2) & 1) ^ 1));
        const uint32_t d = tmem_base + L::D2_COL0 + (uint32_t)(d2b * C);
        for (int j = 0; j < NCH; ++j, ++g) {
          const int hb = g & (FS_GROUPS - 1), s2 = g & (S2 - 1);
          mbar_wait2(&w2_full[s2], (uint32_t)((g / S2) & 1), &h_full[hb], (uint32_t)((g / FS_GROUPS) & 1));
          VV_TR(4, g);
          tc_fence_after();
          const uint32_t ah = tmem_base + L::H_COL0 + (uint32_t)(hb * 32);   // A operand from tensor memory
          const uint64_t dw = dw0 + (uint64_t)((s2 * L::W2_BYTES) >> 4);
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)
            umma_bf16_ts(d, ah + (uint32_t)(kk * 8), dw + (uint64_t)(kk * 2), idesc2, (j | kk) != 0);
          umma_commit(&h_empty[hb]);
          umma_commit(&w2_empty[s2]);
          VV_TR(5, g);
        }
        umma_commit(&d2_full[d2b]);
      }
    }
  } else if (warp == 3) {
    // ===== residual boxes in, finished boxes out (one thread owns the Y ring) =================================
    // p.inplace_add (out == resid): nothing is loaded; a finished box holds gamma * (acc + b2) and is ADDED to the residual
    // in global memory by a TMA reduction, y_full then means "the reduction has read the box: it is free".  Why: the
    // chunk-level timeline of the first C = 256 build showed ~11 000 idle clocks per tile behind the single D2 accumulator,
    // the chain store -> read wait -> residual reload (DRAM latency) -> E2 running once per pair of staging boxes.
    if (elect_one()) {
      const int total_k = my_tiles * NBOX;
      if (p.inplace_add) {
        for (int k = 0; k < total_k; ++k) {
          const int sy = k & (NY - 1);
          mbar_wait(&y_ready[sy], (uint32_t)((k / NY) & 1));
          tma_reduce_add_2d(&tmO, sY + sy * L::Y_BYTES, (k & (NBOX - 1)) * 32, row0 + (k / NBOX) * tile_step);
          tma_store_commit();
          tma_store_wait_read<0>();
          mbar_arrive(&y_full[sy]);
        }
      } else {
        auto load_box = [&](int k) {
          const int sy = k & (NY - 1);
          mbar_expect_tx(&y_full[sy], L::Y_BYTES);
          tma_load_2d(sY + sy * L::Y_BYTES, &tmR, &y_full[sy], (k & (NBOX - 1)) * 32, row0 + (k / NBOX) * tile_step);
        };
        for (int k = 0; k < NY && k < total_k; ++k) load_box(k);
        for (int k = 0; k < total_k; ++k) {
          const int sy = k & (NY - 1);
          mbar_wait(&y_ready[sy], (uint32_t)((k / NY) & 1));
          tma_store_2d(&tmO, sY + sy * L::Y_BYTES, (k & (NBOX - 1)) * 32, row0 + (k / NBOX) * tile_step);
          tma_store_commit();
          tma_store_wait_read<0>();               // the store has read the slot: refill it
          if (k + NY < total_k) load_box(k + NY);
        }
      }
      tma_store_wait_all<0>();
    }
  } else if (warp == 4) {
    // ===== a tiles: [128, C] bf16 in KB blocks, requested as soon as GEMM 1 has read the slot's previous tenant ===========
    if (elect_one()) {
      for (int it = 0; it < my_tiles; ++it) {
        const int sa = it & (NA - 1);
        mbar_wait(&a_empty[sa], (uint32_t)(((it / NA) & 1) ^ 1));
        mbar_expect_tx(&a_full[sa], L::A_BYTES);
#pragma unroll
        for (int kb = 0; kb < KB; ++kb)
          tma_load_2d(sA + sa * L::A_BYTES + kb * (128 * 128), &tmA, &a_full[sa], kb * 64, row0 + it * tile_step);
      }
    }
  } else if (warp < FS_CTRL) {
    // idle control warps (keep the warp count a multiple of 4)
  } else if (warp < FS_CTRL + FS_E1_WARPS) {
    // ===== E1: D1 -> bias + GELU -> bf16 -> swizzled smem (A operand of GEMM 2) ============================
    // Two groups of 8 warps alternate chunks (group G owns H buffer G), so one group's TMEM-load / MUFU /
    // barrier latency overlaps the other's math; within a chunk a warp owns 32 rows x 32 columns.
    const int q = warp & 3;                   // TMEM lane quarter
    constexpr int WPG = FS_E1_WARPS / FS_GROUPS;          // warps per group: 8 (two column halves) or 4
    constexpr int NHF = 8 / WPG;                          // 32-column halves a warp walks per chunk: 1 or 2
    const int grp = (warp - FS_CTRL) / WPG;               // chunks g = grp (mod FS_GROUPS)
    const int ch0 = (WPG == 8) ? (((warp - FS_CTRL) >> 2) & 1) : 0;
    const uint32_t haddr = tmem_base + ((uint32_t)(q * 32) << 16) + L::H_COL0 + (uint32_t)(grp * 32);
    const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16);
    for (int g = grp; g < total_g; g += FS_GROUPS) {
      const int j = g & (NCH - 1), db = g & (ND1 - 1);
      if (q == 0 && ch0 == 0 && lane == 0) VV_TR(6, g);
      mbar_wait(&d1_full[grp], (uint32_t)((g / FS_GROUPS) & 1));
      if (q == 0 && ch0 == 0 && lane == 0) VV_TR(7, g);
      tc_fence_after();
#pragma unroll
      for (int hf = 0; hf < NHF; ++hf) {
        const int ch = ch0 + hf;                          // 32-column half of the 64-column chunk
#if VV_FS_CONST
        const float* bp = p.b1c + j * L::HC + ch * 32;    // constant bank
#else
        const float* bp = p.b1 + j * L::HC + ch * 32;     // global (A/B build: -DVV_FS_CONST=0)
#endif
        uint32_t v[32];
        tmem_ld_32x32(taddr + (uint32_t)(db * L::HC + ch * 32), v);
        tmem_ld_wait();
        if (hf == NHF - 1) {
          if (q == 0 && ch0 == 0 && lane == 0) VV_TR(8, g);
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&d1_empty[db]);      // accumulator is in registers: a later G1 may overwrite it
        }
        uint32_t hv[16];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float* bq = bp + i * 8;
          hv[i * 4 + 0] = gelu2_bias_pack(__uint_as_float(v[i * 8 + 0]), __uint_as_float(v[i * 8 + 1]), bq[0], bq[1]);
          hv[i * 4 + 1] = gelu2_bias_pack(__uint_as_float(v[i * 8 + 2]), __uint_as_float(v[i * 8 + 3]), bq[2], bq[3]);
          hv[i * 4 + 2] = gelu2_bias_pack(__uint_as_float(v[i * 8 + 4]), __uint_as_float(v[i * 8 + 5]), bq[4], bq[5]);
          hv[i * 4 + 3] = gelu2_bias_pack(__uint_as_float(v[i * 8 + 6]), __uint_as_float(v[i * 8 + 7]), bq[6], bq[7]);
        }
        if (hf == 0) {
          if (q == 0 && ch0 == 0 && lane == 0) VV_TR(9, g);
          mbar_wait(&h_empty[grp], (uint32_t)(((g / FS_GROUPS) & 1) ^ 1));
          if (q == 0 && ch0 == 0 && lane == 0) VV_TR(10, g);
          tc_fence_after();
        }
        tmem_st_32x16(haddr + (uint32_t)(ch * 16), hv);   // thread = row: packed pairs (k, k+1) -> one 32-bit column
      }
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&h_full[grp]);
      if (q == 0 && ch0 == 0 && lane == 0) VV_TR(11, g);
    }
  } else {
    // ===== E2: D2 -> + b2, * gamma, + residual (TMA-staged box, swizzled) -> in place -> TMA store ===========
    const int q = warp & 3;
    const int r = q * 32 + lane;
    int k = 0;
    for (int it = 0; it < my_tiles; ++it) {
      const int d2b = it % ND2;
      const long long m = (long long)row0 + (long long)it * tile_step + r;
      const bool valid = m < p.R;
      long long brow = 0;
      if (p.out_bf16 != nullptr && valid) {
        const long long b = m / p.rows_pb;
        brow = b * p.ob_rpb + p.ob_off + (m - b * p.rows_pb);
      }
      mbar_wait(&d2_full[d2b], (uint32_t)((it / ND2) & 1));
      tc_fence_after();
#pragma unroll 1
      for (int c = 0; c < NBOX; ++c, ++k) {
        const int sy = k & (NY - 1);
        uint32_t v[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + L::D2_COL0 + (uint32_t)(d2b * C + c * 32), v);
        // (in-place mode: wait for the reduction of the box's previous tenant; the first NY waits pass on fresh barriers)
        mbar_wait(&y_full[sy], (uint32_t)(((k / NY) & 1) ^ (p.inplace_add ? 1 : 0)));
        tmem_ld_wait();
        uint8_t* rrow = sY + sy * L::Y_BYTES + r * 128;
        float x[32];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int col = c * 32 + i * 4;
          float4* cell = reinterpret_cast<float4*>(rrow + ((i ^ (r & 7)) * 16));
          const float4 rv = p.inplace_add ? make_float4(0.f, 0.f, 0.f, 0.f) : *cell;
#if VV_FS_CONST
          const float* bv = p.b2c + col;                  // constant bank, same arithmetic as before
          const float* gv = p.gammac + col;
#else
          const float* bv = p.b2 + col;
          const float* gv = p.gamma + col;
#endif
          x[i * 4 + 0] = fmaf(gv[0], __uint_as_float(v[i * 4 + 0]) + bv[0], rv.x);
          x[i * 4 + 1] = fmaf(gv[1], __uint_as_float(v[i * 4 + 1]) + bv[1], rv.y);
          x[i * 4 + 2] = fmaf(gv[2], __uint_as_float(v[i * 4 + 2]) + bv[2], rv.z);
          x[i * 4 + 3] = fmaf(gv[3], __uint_as_float(v[i * 4 + 3]) + bv[3], rv.w);
          *cell = make_float4(x[i * 4], x[i * 4 + 1], x[i * 4 + 2], x[i * 4 + 3]);
        }
        if (valid && p.out_bf16 != nullptr) {
          __nv_bfloat16* bp = p.out_bf16 + brow * C + c * 32;
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            uint4 pk;
            pk.x = pack_bf16x2(x[i * 8 + 0], x[i * 8 + 1]);
            pk.y = pack_bf16x2(x[i * 8 + 2], x[i * 8 + 3]);
            pk.z = pack_bf16x2(x[i * 8 + 4], x[i * 8 + 5]);
            pk.w = pack_bf16x2(x[i * 8 + 6], x[i * 8 + 7]);
            *reinterpret_cast<uint4*>(bp + i * 8) = pk;
          }
        }
        fence_proxy_async_smem();               // staged box -> visible to the TMA store
        __syncwarp();
        if (lane == 0) mbar_arrive(&y_ready[sy]);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&d2_empty[d2b]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, L::TMEM_COLS);
}

}  // namespace vv
