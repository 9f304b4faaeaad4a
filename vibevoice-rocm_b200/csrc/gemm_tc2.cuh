// gemm_tc2.cuh — persistent version of the tap GEMM (gemm_tc.cuh) for 128 x {64, 128, 256} x 64 tiles, i.e. every hot
// dense layer with N % 64 == 0: FFN linear1/linear2 at C >= 256, the polyphase transposed convs, the decoder stem, the
// encoder's strided convs.
//
// What changed against gemm_tc_kernel, and why (ncu on the first round-1 build, profiles/):
//   * row-per-thread global loads/stores in the epilogue cost one L1TEX tag cycle per ROW per instruction;
//     L1TEX was the busiest unit.  Now the epilogue works on a shared-memory staging tile in the
//     128B-swizzled layout TMA uses: the residual tile is TMA-LOADED into it, the epilogue updates it in
//     place (thread = row, conflict-free in 4 wavefronts), and a dedicated warp TMA-STORES it.
//   * one CTA per SM loops over tiles (N fastest, so CTAs running together share the A tile in L2);
//     two TMEM accumulators and two staging slots let tile i's epilogue overlap tile i+1's TMA + MMA.
//   * 8 epilogue warps (two per TMEM lane quarter, splitting the columns) instead of 4.
//   * BN = 256 instantiation (N % 256 == 0 and enough tiles): one 128 x 256 accumulator per MMA tile halves the
//     shared-memory operand traffic per flop (A 16 KB + B 32 KB per 512 MMA clocks = 96 B/clk against 128 B/clk at
//     BN = 128, which sits exactly on the TMA/smem feed limit); the epilogue walks it in 128-column (bf16 out) or
//     64-column (fp32 out) sub-tiles through the same staging slots, which leaves room for a 4-stage A/B ring.
//     K2 step: 236 -> 216 ms; FFN GEMMs 650-930 -> 800-1235 TFLOP/s.
// Roles: warp 0 TMA loads (A/B ring), warp 1 MMA issue + TMEM owner, warps 2..9 (2..17 for GELU) epilogue, last warp =
// staging manager: TMA-stores finished sub-tiles and TMA-loads the residual of the sub-tile that takes the slot next.
#pragma once
#include <cuda.h>
#include "ptx.cuh"
#include "epilogue.cuh"
#include "gemm_tc.cuh"

namespace vv {

// epilogue warps: 8 (two per TMEM lane quarter, 64 columns each) — or 16 (32 columns each) for the GELU epilogue,
// which is issue-bound and needs more warps in flight to fill the issue slots (measured on the fused FFN: IPC 1.9 -> 2.2+)
template <int EPI> struct G2Warps {
  static constexpr int EPI_WARPS = (EPI == EPI_GELU_BF16 || EPI == EPI_SWIGLU_BF16) ? 16 : 8;
  static constexpr int THREADS = (2 + EPI_WARPS + 1) * 32;
};
constexpr int G2_BN = 128, G2_BK = 64;            // G2_BN: the narrow instantiation (and the epilogue sub-tile width)
constexpr int G2_A_BYTES = GEMM_BM * G2_BK * 2;   // 16 KB
constexpr int G2_BOX_BYTES = 128 * 128;           // one staged box: 128 rows x 128 bytes

// DEEP = long K loop (>= 32 k-blocks): the mainloop dominates, so shared memory goes to a deeper A/B ring
// and the epilogue staging is single-buffered; otherwise the epilogue dominates and staging is double-buffered.
template <int EPI, bool DEEP, int BN>
struct Gemm2Cfg {
  static_assert(BN == 64 || BN == 128 || BN == 256, "tile width");
  static constexpr bool OUT_BF16 = (EPI == EPI_GELU_BF16 || EPI == EPI_SWIGLU_BF16);
  static_assert(BN != 64 || !OUT_BF16, "the 64-wide tile has fp32 epilogues only");
  static_assert(EPI != EPI_SWIGLU_BF16 || BN == 256, "the SwiGLU epilogue pairs the two halves of a 256-wide tile");
  static constexpr int OUTW = (EPI == EPI_SWIGLU_BF16) ? BN / 2 : BN;   // output columns one MMA tile produces
  static constexpr int COLS_PER_BOX = OUT_BF16 ? 64 : 32;
  // epilogue sub-tile: 128 columns, or 64 for the fp32 epilogues of the wide tile (32 KB of staging instead of 64 KB
  // buys a 4th ring stage for long K loops, or a second staging slot for short ones)
#ifndef VV_G2_SUBW128_NARROW
#define VV_G2_SUBW128_NARROW 0     // 1: 128-column fp32 sub-tiles on the 128-wide tile (the earlier layout; A/B: the
#endif                             //    64 -> 128-channel upsampler runs 4.2 ms with it, 3.15 ms without)
  static constexpr int SUBW = (!OUT_BF16 && (BN != 128 || !VV_G2_SUBW128_NARROW)) ? 64 : 128;
  static constexpr int NSUB = OUTW / SUBW;                           // sub-tiles per MMA tile: 1, 2 or 4
  static constexpr int BOXES = SUBW / COLS_PER_BOX;                  // per sub-tile: 2 or 4
  static constexpr int STG_BYTES = BOXES * G2_BOX_BYTES;             // 32 KB or 64 KB
  static constexpr int B_BYTES = BN * G2_BK * 2;                     // 16 KB or 32 KB
  // BN = 64 (the last upsampler, the encoder's first strided conv and its head): a 24 KB ring stage; the short-K
  // layers are store-bound, so three staging slots keep TMA stores in flight behind the epilogue
  static constexpr int NSTG = (BN == 64) ? (DEEP ? 2 : 3) : (DEEP ? 1 : 2);
  static constexpr int STAGES = (BN == 64)  ? (DEEP ? 6 : 4)
                              : (BN == 128) ? (SUBW == 64 ? (DEEP ? 6 : 5) : (DEEP ? (OUT_BF16 ? 6 : 5) : (OUT_BF16 ? 5 : 3)))
                                            : (DEEP ? 4 : 3);                                        // A/B ring depth
  static constexpr int OFF_RING = 0;
  static constexpr int OFF_STG = STAGES * (G2_A_BYTES + B_BYTES);
  static constexpr int OFF_BAR = OFF_STG + NSTG * STG_BYTES;
  static constexpr int TOTAL = OFF_BAR + 512 + 1024;
  static_assert(TOTAL <= 232448, "shared memory budget");
};

struct Gemm2Args {
  GemmArgs g;          // shapes, bias / gamma, optional haloed bf16 copy (out_bf16, ob_rpb, ob_off) for EPI_RESID
};

template <int EPI, bool DEEP, int BN>
__global__ void __launch_bounds__(G2Warps<EPI>::THREADS, 1)
gemm_tc2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                const __grid_constant__ CUtensorMap tmR, const __grid_constant__ CUtensorMap tmO,
                const GemmArgs g) {
  using Cfg = Gemm2Cfg<EPI, DEEP, BN>;
  constexpr int S = Cfg::STAGES;
  constexpr int NSTG = Cfg::NSTG;
  constexpr int BK = G2_BK, NSUB = Cfg::NSUB, SUBW = Cfg::SUBW;
  constexpr int G2_B_BYTES = Cfg::B_BYTES;
  constexpr int EW = G2Warps<EPI>::EPI_WARPS;
  constexpr int CPW = SUBW / (EW / 4);      // columns per epilogue warp inside a sub-tile: 64 or 32
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_align_1k(smem_raw);
  uint8_t* sA = smem + Cfg::OFF_RING;
  uint8_t* sB = sA + S * G2_A_BYTES;
  uint8_t* sStg = smem + Cfg::OFF_STG;
  uint64_t* bars = (uint64_t*)(smem + Cfg::OFF_BAR);
  uint64_t* full = bars;                   // [S]
  uint64_t* empty = full + S;              // [S]
  uint64_t* acc_full = empty + S;          // [2]
  uint64_t* acc_empty = acc_full + 2;      // [2]
  uint64_t* stg_full = acc_empty + 2;      // [NSTG] residual tile landed (EPI_RESID)
  uint64_t* stg_ready = stg_full + NSTG;   // [NSTG] epilogue finished writing the staged output
  uint64_t* stg_empty = stg_ready + NSTG;  // [NSTG] TMA store has read the slot
  uint32_t* tmem_slot = (uint32_t*)(stg_empty + NSTG);
  static_assert((2 * S + 4 + 3 * NSTG + 1) * 8 <= 512, "barrier block");

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int bt = 1 << g.bt_log2;
  const int bb = GEMM_BM >> g.bt_log2;
  const int tiles_t = (g.rows + bt - 1) >> g.bt_log2;
  const int tiles_b = (g.nbatch + bb - 1) / bb;
  const int tiles_n = g.N / BN;
  const int KS = g.k_slices > 1 ? g.k_slices : 1;        // K slices as separate tiles (GemmArgs::k_slices)
  const long long ntiles = (long long)tiles_t * tiles_b * tiles_n * KS;
  const long long my_tiles = (ntiles > blockIdx.x) ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  const int nk = g.n_taps * g.kb_per_tap / KS;           // k-blocks per tile (the host checks divisibility)

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmO);
    if (EPI == EPI_RESID) tma_prefetch_desc(&tmR);
    for (int s = 0; s < S; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], EW); }
    for (int i = 0; i < NSTG; ++i) {
      mbar_init(&stg_full[i], 1);
      mbar_init(&stg_ready[i], EW);
      mbar_init(&stg_empty[i], 1);
    }
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, 2 * BN);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch();
  pdl_wait();

  // tile index -> (n, t, b) coordinates
  auto tile_coords = [&](long long it, int& tile_n, int& tile_t, int& tile_b, int& slice) {
    long long tile = blockIdx.x + it * gridDim.x;
    slice = (int)(tile % KS);
    tile /= KS;
    tile_n = (int)(tile % tiles_n);
    const long long tm = tile / tiles_n;
    tile_t = (int)(tm % tiles_t);
    tile_b = (int)(tm / tiles_t);
  };

  if (warp == 0) {
    // ===== TMA loads ===============================================================================
    if (elect_one()) {
      long long kbg = 0;                                  // global k-block counter (ring position)
      for (long long it = 0; it < my_tiles; ++it) {
        int tile_n, tile_t, tile_b, slice;
        tile_coords(it, tile_n, tile_t, tile_b, slice);
        for (int kl = 0; kl < nk; ++kl, ++kbg) {
          const int kb = slice * nk + kl;
          const int s = (int)(kbg % S);
          mbar_wait(&empty[s], (uint32_t)(((kbg / S) & 1) ^ 1));
          const int tap = kb / g.kb_per_tap;
          const int kin = (kb - tap * g.kb_per_tap) * BK;
          mbar_expect_tx(&full[s], G2_A_BYTES + G2_B_BYTES);
          tma_load_3d(sA + s * G2_A_BYTES, &tmA, &full[s], kin, tile_t * bt + tap, tile_b * bb);
          tma_load_2d(sB + s * G2_B_BYTES, &tmB, &full[s], kb * BK, tile_n * BN);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer ================================================================================
    if (elect_one()) {
      constexpr uint32_t idesc = umma_idesc_bf16(GEMM_BM, BN);
      long long kbg = 0;
      for (long long it = 0; it < my_tiles; ++it) {
        const int ab = (int)(it & 1);
        mbar_wait(&acc_empty[ab], (uint32_t)(((it >> 1) & 1) ^ 1));
        tc_fence_after();
        for (int kb = 0; kb < nk; ++kb, ++kbg) {
          const int s = (int)(kbg % S);
          mbar_wait(&full[s], (uint32_t)((kbg / S) & 1));
          tc_fence_after();
          const uint64_t da = umma_smem_desc<128>(smem_u32(sA + s * G2_A_BYTES));
          const uint64_t db = umma_smem_desc<128>(smem_u32(sB + s * G2_B_BYTES));
#pragma unroll
          for (int kk = 0; kk < BK / 16; ++kk)
            umma_bf16(tmem_base + (uint32_t)(ab * BN), da + (uint64_t)(kk * 2), db + (uint64_t)(kk * 2), idesc,
                      (kb | kk) != 0);
          umma_commit(&empty[s]);
        }
        umma_commit(&acc_full[ab]);
      }
    }
  } else if (warp < 2 + EW) {
    // ===== epilogue: TMEM -> (bias | GELU | layer-scale + residual) -> staged tile in shared memory ==========
    const int q = warp & 3;
    const int colgrp = (warp - 2) >> 2;                   // columns [colgrp*CPW, +CPW) of the tile
    const int r = q * 32 + lane;
    for (long long it = 0; it < my_tiles; ++it) {
      int tile_n, tile_t, tile_b, slice;
      tile_coords(it, tile_n, tile_t, tile_b, slice);
      const int ab = (int)(it & 1);
      mbar_wait(&acc_full[ab], (uint32_t)((it >> 1) & 1));
#pragma unroll 1
      for (int h = 0; h < NSUB; ++h) {                    // 128-column sub-tiles of the accumulator
      const long long e = it * NSUB + h;
      const int ss = (int)(e % NSTG);
      const int n0 = tile_n * Cfg::OUTW + h * SUBW + colgrp * CPW;
      if (EPI == EPI_RESID) mbar_wait(&stg_full[ss], (uint32_t)((e / NSTG) & 1));
      else mbar_wait(&stg_empty[ss], (uint32_t)(((e / NSTG) & 1) ^ 1));
      tc_fence_after();
      uint8_t* stg = sStg + ss * Cfg::STG_BYTES + r * 128;
#pragma unroll
      for (int c = 0; c < CPW / 32; ++c) {                // 32-column groups
        uint32_t v[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(ab * BN + h * SUBW + colgrp * CPW + c * 32), v);
        tmem_ld_wait();
        const int n = n0 + c * 32;                        // global column of v[0]
        const int nl = colgrp * CPW + c * 32;             // column inside the sub-tile
        if (EPI == EPI_SWIGLU_BF16) {
          // v = gate columns; the matching up columns sit 128 accumulator columns further.  silu(g) * u -> bf16
          uint32_t u[32];
          tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(ab * BN + 128 + colgrp * CPW + c * 32), u);
          tmem_ld_wait();
          uint8_t* box = stg + (nl >> 6) * G2_BOX_BYTES;
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            uint32_t pkv[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const float g0 = __uint_as_float(v[i * 8 + 2 * j]), g1 = __uint_as_float(v[i * 8 + 2 * j + 1]);
              const float h0 = g0 / (1.0f + __expf(-g0)) * __uint_as_float(u[i * 8 + 2 * j]);
              const float h1 = g1 / (1.0f + __expf(-g1)) * __uint_as_float(u[i * 8 + 2 * j + 1]);
              pkv[j] = pack_bf16x2(h0, h1);
            }
            const int chunk = (((nl & 63) >> 3) + i) ^ (r & 7);
            *reinterpret_cast<uint4*>(box + chunk * 16) = make_uint4(pkv[0], pkv[1], pkv[2], pkv[3]);
          }
        } else if (EPI == EPI_GELU_BF16) {
          uint8_t* box = stg + (nl >> 6) * G2_BOX_BYTES;  // 64 bf16 columns per box
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float4 b0 = __ldg(reinterpret_cast<const float4*>(g.bias + n + i * 8));
            const float4 b1 = __ldg(reinterpret_cast<const float4*>(g.bias + n + i * 8 + 4));
            uint4 pk;
            pk.x = gelu2_bias_pack(__uint_as_float(v[i * 8 + 0]), __uint_as_float(v[i * 8 + 1]), b0.x, b0.y);
            pk.y = gelu2_bias_pack(__uint_as_float(v[i * 8 + 2]), __uint_as_float(v[i * 8 + 3]), b0.z, b0.w);
            pk.z = gelu2_bias_pack(__uint_as_float(v[i * 8 + 4]), __uint_as_float(v[i * 8 + 5]), b1.x, b1.y);
            pk.w = gelu2_bias_pack(__uint_as_float(v[i * 8 + 6]), __uint_as_float(v[i * 8 + 7]), b1.z, b1.w);
            const int chunk = (((nl & 63) >> 3) + i) ^ (r & 7);
            *reinterpret_cast<uint4*>(box + chunk * 16) = pk;
          }
        } else {
          uint8_t* box = stg + (nl >> 5) * G2_BOX_BYTES;  // 32 fp32 columns per box
          float x[32];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            float4 bv = __ldg(reinterpret_cast<const float4*>(g.bias + n + i * 4));
            if (slice != 0) bv = make_float4(0.f, 0.f, 0.f, 0.f);          // partial-sum planes: plane 0 carries the bias
            float4* cell = reinterpret_cast<float4*>(box + ((i ^ (r & 7)) * 16));
            float4 o;
            o.x = __uint_as_float(v[i * 4 + 0]) + bv.x;
            o.y = __uint_as_float(v[i * 4 + 1]) + bv.y;
            o.z = __uint_as_float(v[i * 4 + 2]) + bv.z;
            o.w = __uint_as_float(v[i * 4 + 3]) + bv.w;
            if (EPI == EPI_RESID) {
              const float4 gv = __ldg(reinterpret_cast<const float4*>(g.gamma + n + i * 4));
              const float4 rv = *cell;
              o.x = fmaf(gv.x, o.x, rv.x);
              o.y = fmaf(gv.y, o.y, rv.y);
              o.z = fmaf(gv.z, o.z, rv.z);
              o.w = fmaf(gv.w, o.w, rv.w);
            }
            *cell = o;
            x[i * 4 + 0] = o.x; x[i * 4 + 1] = o.y; x[i * 4 + 2] = o.z; x[i * 4 + 3] = o.w;
          }
          if (EPI == EPI_RESID && g.out_bf16 != nullptr) {  // haloed bf16 copy for the upsampler (last block only)
            const int b = tile_b * bb + (r >> g.bt_log2);
            const int t = tile_t * bt + (r & (bt - 1));
            if (b < g.nbatch && t < g.rows) {
              long long brow = (long long)b * g.ob_rpb + g.ob_off + t;
              if (g.ob_rows_pb > 0) {                      // flat row range of a [items, rows_pb] tensor
                const long long m = g.ob_row0 + t;
                const long long item = m / g.ob_rows_pb;
                brow = item * g.ob_rpb + g.ob_off + (m - item * g.ob_rows_pb);
              }
              __nv_bfloat16* op = g.out_bf16 + brow * g.N + n;
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                uint4 pk;
                pk.x = pack_bf16x2(x[i * 8 + 0], x[i * 8 + 1]);
                pk.y = pack_bf16x2(x[i * 8 + 2], x[i * 8 + 3]);
                pk.z = pack_bf16x2(x[i * 8 + 4], x[i * 8 + 5]);
                pk.w = pack_bf16x2(x[i * 8 + 6], x[i * 8 + 7]);
                *reinterpret_cast<uint4*>(op + i * 8) = pk;
              }
            }
          }
        }
      }
      tc_fence_before();
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        if (h == NSUB - 1) mbar_arrive(&acc_empty[ab]);
        mbar_arrive(&stg_ready[ss]);
      }
      }
    }
  } else {
    // ===== staging manager: TMA stores + residual TMA loads ===========================================
    if (elect_one()) {
      const long long total_e = my_tiles * NSUB;
      auto sub_coords = [&](long long e, int& col, int& t0, int& b0) {
        int tile_n, tile_t, tile_b, slice;
        tile_coords(e / NSUB, tile_n, tile_t, tile_b, slice);
        col = tile_n * Cfg::OUTW + (int)(e % NSUB) * SUBW;
        t0 = tile_t * bt;
        b0 = tile_b * bb + slice;          // k_slices > 1 only with nbatch == 1: the output's third coordinate is the plane
      };
      auto load_resid = [&](long long e) {                // residual sub-tile -> its staging slot
        const int ss = (int)(e % NSTG);
        int col, t0, b0;
        sub_coords(e, col, t0, b0);
        mbar_expect_tx(&stg_full[ss], Cfg::STG_BYTES);
#pragma unroll
        for (int c = 0; c < Cfg::BOXES; ++c)
          tma_load_3d(sStg + ss * Cfg::STG_BYTES + c * G2_BOX_BYTES, &tmR, &stg_full[ss], col + c * Cfg::COLS_PER_BOX, t0, b0);
      };
      if (EPI == EPI_RESID)
        for (long long e = 0; e < NSTG && e < total_e; ++e) load_resid(e);
      for (long long e = 0; e < total_e; ++e) {
        const int ss = (int)(e % NSTG);
        int col, t0, b0;
        sub_coords(e, col, t0, b0);
        mbar_wait(&stg_ready[ss], (uint32_t)((e / NSTG) & 1));
#pragma unroll
        for (int c = 0; c < Cfg::BOXES; ++c)
          tma_store_3d(&tmO, sStg + ss * Cfg::STG_BYTES + c * G2_BOX_BYTES, col + c * Cfg::COLS_PER_BOX, t0, b0);
        tma_store_commit();
        tma_store_wait_read<0>();
        if (EPI == EPI_RESID) { if (e + NSTG < total_e) load_resid(e + NSTG); }   // the slot's next tenant
        else mbar_arrive(&stg_empty[ss]);
      }
      tma_store_wait_all<0>();
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 2 * BN);
}

}  // namespace vv
