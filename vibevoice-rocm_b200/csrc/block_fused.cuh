// block_fused.cuh — kernel classes (a) + (b) for the narrow stages (C = 32, 64): a WHOLE Block1D in ONE persistent
// kernel.  x (fp32) is read once, the new x is written once: 8 B per activation element — exactly SURVEY's
// algorithmic bytes for a mixer half plus an FFN half (4 + 4), against 20 B for the two-kernel form
// (mixer: x in, y + a out; FFN: a + y in, x out).
//
//   y   = x + gamma * ( dwconv7( RMSNorm(x) * norm_w ) + conv_bias )               (reference :927-932 / :667-672)
//   out = y + ffn_gamma * ( W2 * gelu_erf( W1 * (RMSNorm(y) * ffn_norm_w) + b1 ) + b2 )   (:936-942 / :675-682)
//
// Per tile of 128 time steps of one batch item (tiles never straddle items):
//   mixer warps (C/8 of them): 16 strips of 8 rows; a group of C/4 lanes owns a strip, each lane 4 channels.  x rows
//       come straight from global memory into registers (coalesced float4, the NEXT 7-row batch is always in
//       flight), 6 warm-up rows per strip rebuild the normalised sliding window (causal context: zeros offline,
//       the stream's cached raw rows when a state is given, reference :327-382).  Writes y (fp32) into a staging
//       slot in the layout E2 / the TMA store use and a = bf16(RMSNorm(y) * w) into the 128B/64B-swizzled K-major
//       tile tcgen05.mma reads as GEMM 1's A operand.  Neither y nor a ever touches HBM.
//   G1 warp : D1[128, 64] = a * W1_chunk^T      (W1, W2 resident in shared memory; 2 TMEM accumulators)
//   E1, 2 x 4 warps alternating chunks: tcgen05.ld D1 -> + b1 -> GELU -> bf16 pairs -> tcgen05.st H (tensor memory)
//   G2 warp : D2[128, C] += H * W2_chunk^T      (A operand from TENSOR memory, 2 accumulators)
//   E2, 4 warps: tcgen05.ld D2 -> + b2, * ffn_gamma, + y (staged slot, in place) -> TMA store warp
// Register file: the mixer warps need ~140 registers (two 7-row batches, the window, the taps), everything else
// <= 64, so the warpgroups re-balance with setmaxnreg.  All hand-offs are mbarriers.
#pragma once
#include <cuda.h>
#include "ptx.cuh"
#include "epilogue.cuh"
#include "mixer.cuh"       // StateRef

namespace vv {

struct BlockFusedArgs {
  const float* x;            // [nbatch, rows_pb, C] fp32 (NOT aliased with the output: tiles read 6 rows of their neighbour)
  int nbatch, rows_pb;
  int tiles_pb;              // ceil(rows_pb / 128)
  float eps;
  const float* norm_w;       // [C]
  const float* dw_w;         // [7][C] tap-major
  const float* dw_b;         // [C]
  const float* gamma;        // [C]
  const float* ffn_norm_w;   // [C]
  const float* b1;           // [4C]  (pre-scaled by 0.5 like W1, see epilogue.cuh)
  const float* b2;           // [C]
  const float* ffn_gamma;    // [C]
  __nv_bfloat16* out_bf16;   // optional padded bf16 copy: [nbatch, ob_rpb, C], row t of an item lands at row ob_off + t
  long long ob_rpb;          //   (decoder: ob_rpb = rows_pb + 1, ob_off = 1 = the upsampler halo row)
  int ob_off;
  StateRef st;               // mixer context: fp32 [6][C] raw rows per stream (null base = offline)
};

// VV_BF_MXSETS32 = 2: at C = 32 two sets of mixer warps work on alternate tiles (set s takes this CTA's tiles s, s + 2,
// ...).  ncu (profiles/block_fused_s6_r01.txt + source page): the 4 mixer warps -- one per scheduler -- are never idle
// and issue at ~0.2 IPC (dependency, shuffle and load latency), everything downstream waits for them; a second set doubles
// the latency-bound work in flight without adding warm-up rows.  All slots and barriers stay per tile.
// MEASURED SLOWER on B200 (tools/kbench.py --kind block --stage 6, 5760 frames): 1.72 ms -> 2.06 ms with 2 sets (parity
// tests green); the 8 mixer warps get 128 registers instead of 168 (256 B of spills) and share the issue slots and the
// L1 with E1 / E2 -- so the default stays 1 and the finding stands next to DESIGN.md section 4 (5).
#ifndef VV_BF_MXSETS32
#define VV_BF_MXSETS32 1
#endif
constexpr int BF_E1_WARPS = 8;
constexpr int BF_E2_WARPS = 4;

template <int C>
struct BlockFusedCfg {
  static_assert(C == 32 || C == 64, "block_fused_kernel is laid out for C = 32 / 64");
  static constexpr int MXW_TILE = C / 8;            // mixer warps that build one tile
  static constexpr int MXSETS = (C == 32) ? VV_BF_MXSETS32 : 1;   // sets of them, working on alternate tiles
  static constexpr int MXW = MXW_TILE * MXSETS;     // mixer warps in the CTA
  static constexpr int LPR = C / 4;                 // lanes per row
  static constexpr int WARPS = BF_E1_WARPS + BF_E2_WARPS + MXW + 4;
  static constexpr int THREADS = WARPS * 32;
  static constexpr int HC = 64;                     // hidden units per chunk
  static constexpr int NCH = 4 * C / HC;            // chunks per tile
  static constexpr int ROWB = C * 2;                // bytes per a / W1 row (64 or 128 = the swizzle span)
  static constexpr int NBOX = C / 32;
  static constexpr int NA = (C == 64) ? 2 : 4;
  static constexpr int NY = (C == 64) ? 3 : 4;
  static constexpr int W1_BYTES = 4 * C * ROWB;     // [4C rows, C]
  static constexpr int W2_BYTES = NCH * C * 128;    // NCH K-blocks of [C rows, 64]
  static constexpr int A_BYTES = 128 * ROWB;
  static constexpr int Y_BYTES = NBOX * 128 * 128;  // NBOX boxes of [128 rows, 32 fp32]
  static constexpr int OFF_W1 = 0;
  static constexpr int OFF_W2 = OFF_W1 + W1_BYTES;
  static constexpr int OFF_A = OFF_W2 + W2_BYTES;
  static constexpr int OFF_Y = OFF_A + NA * A_BYTES;
  static constexpr int OFF_BAR = OFF_Y + NY * Y_BYTES;
  static constexpr int NBARS = 1 + 2 * NA + 3 * NY + 12;
  static constexpr int TOTAL = OFF_BAR + 512 + 1024;
  static_assert(NBARS * 8 + 8 <= 512, "barrier block");
  static_assert(TOTAL <= 232448, "exceeds the 227 KB dynamic shared memory limit");
  static constexpr uint32_t TMEM_COLS = 512;        // D1: 2 x 64 at 0; H: 2 x 32 at 128; D2: 2 x C at 256
  static constexpr uint32_t H_COL0 = 128, D2_COL0 = 256;
  // Register budgets per role after setmaxnreg.  The pool setmaxnreg.inc draws from is what the CTA was LAUNCHED
  // with (threads x launch registers), not the whole register file: with 768 threads x 80 registers a 144-register
  // mixer (total 65536) hung on B200, 128 (total 61440) runs.
  static constexpr int REG_LAUNCH = (65536 / THREADS) / 8 * 8;      // what __launch_bounds__(THREADS, 1) yields
  static constexpr int REG_E = 64, REG_E2 = 64, REG_CTRL = 32, REG_MX = (C == 64 || MXSETS > 1) ? 128 : 168;   // (136 / 56 / 24 at C = 64: spill-free mixer, not faster)
  static_assert(32 * (BF_E1_WARPS * REG_E + BF_E2_WARPS * REG_E2 + 4 * REG_CTRL + MXW * REG_MX) <= THREADS * REG_LAUNCH,
                "setmaxnreg budget exceeds the CTA's launch allocation");
};

template <int N> __device__ __forceinline__ void reg_dec() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N)); }
template <int N> __device__ __forceinline__ void reg_inc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N)); }

// sum over the LPR lanes of a group (butterfly), 7 values at a time so the shuffle latencies overlap
template <int LPR>
__device__ __forceinline__ void group_sum7(float (&s)[7]) {
#pragma unroll
  for (int m = LPR / 2; m >= 1; m >>= 1) {
#pragma unroll
    for (int i = 0; i < 7; ++i) s[i] += __shfl_xor_sync(0xffffffffu, s[i], m);
  }
}
// 1/sqrt(mean + eps): one MUFU op (rsqrt.approx, 2 ulp; the statistic feeds a bf16 operand and a fp32 residual update)
__device__ __forceinline__ float rstd_of(float ss, float inv_c, float eps) {
  float r;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(fmaf(ss, inv_c, eps)));
  return r;
}
template <int LPR>
__device__ __forceinline__ float group_sum1(float s) {
#pragma unroll
  for (int m = LPR / 2; m >= 1; m >>= 1) s += __shfl_xor_sync(0xffffffffu, s, m);
  return s;
}

// STATE = false is the offline form (no stream context): the context pointers and their per-row checks are compiled
// out of the mixer loop (they cost registers and tie their spill reloads to the scoreboards of the prefetch loads).
template <int C, bool STATE>
__global__ void __launch_bounds__(BlockFusedCfg<C>::THREADS, 1)
block_fused_kernel(const __grid_constant__ CUtensorMap tmO, const __grid_constant__ CUtensorMap tmW1,
                   const __grid_constant__ CUtensorMap tmW2, const BlockFusedArgs p) {
  using L = BlockFusedCfg<C>;
  constexpr int NA = L::NA, NY = L::NY, NCH = L::NCH, NBOX = L::NBOX, MXW = L::MXW, LPR = L::LPR;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_align_1k(smem_raw);
  uint8_t* sW1 = smem + L::OFF_W1;
  uint8_t* sW2 = smem + L::OFF_W2;
  uint8_t* sA = smem + L::OFF_A;
  uint8_t* sY = smem + L::OFF_Y;
  uint64_t* bars = (uint64_t*)(smem + L::OFF_BAR);
  uint64_t* w_full = bars;                 // [1]
  uint64_t* a_full = bars + 1;             // [NA]  mixer wrote the a tile
  uint64_t* a_empty = a_full + NA;         // [NA]  GEMM 1 has read it
  uint64_t* y_full = a_empty + NA;         // [NY]  mixer wrote the y tile
  uint64_t* y_ready = y_full + NY;         // [NY]  E2 turned it into the block output
  uint64_t* y_empty = y_ready + NY;        // [NY]  the TMA store has read it
  uint64_t* d1_full = y_empty + NY;        // [2]
  uint64_t* d1_empty = d1_full + 2;        // [2]
  uint64_t* h_full = d1_empty + 2;         // [2]
  uint64_t* h_empty = h_full + 2;          // [2]
  uint64_t* d2_full = h_empty + 2;         // [2]
  uint64_t* d2_empty = d2_full + 2;        // [2]
  uint32_t* tmem_slot = (uint32_t*)(d2_empty + 2);

  // hardware warps: [0, 8) E1, [8, 12) E2, [12, 12 + MXW) mixer, last 4: W loader, G1 issuer, G2 issuer, TMA store
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  constexpr int WARP_E2 = BF_E1_WARPS, WARP_MX = BF_E1_WARPS + BF_E2_WARPS, WARP_CTRL = WARP_MX + MXW;
  const int ntiles = p.nbatch * p.tiles_pb;
  const int my_tiles = (ntiles > (int)blockIdx.x) ? (ntiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  const int total_g = my_tiles * NCH;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmO);
    tma_prefetch_desc(&tmW1);
    tma_prefetch_desc(&tmW2);
    mbar_init(w_full, 1);
    for (int i = 0; i < NA; ++i) { mbar_init(&a_full[i], L::MXW_TILE); mbar_init(&a_empty[i], 1); }
    for (int i = 0; i < NY; ++i) {
      mbar_init(&y_full[i], L::MXW_TILE);
      mbar_init(&y_ready[i], BF_E2_WARPS);
      mbar_init(&y_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&d1_full[i], 1);
      mbar_init(&d1_empty[i], BF_E1_WARPS / 2);
      mbar_init(&h_full[i], BF_E1_WARPS / 2);
      mbar_init(&h_empty[i], 1);
      mbar_init(&d2_full[i], 1);
      mbar_init(&d2_empty[i], BF_E2_WARPS);
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
  // ===== weights: loaded once, resident for the whole launch.  They are constants (written at finalize), so under
  // programmatic dependent launch their TMA loads go out BEFORE this CTA waits for the previous kernel to drain. =====
  if (warp == BF_E1_WARPS + BF_E2_WARPS + MXW && elect_one()) {
    mbar_expect_tx(w_full, L::W1_BYTES + L::W2_BYTES);
#pragma unroll
    for (int j = 0; j < 4 * C / 128; ++j) tma_load_2d(sW1 + j * 128 * L::ROWB, &tmW1, w_full, 0, j * 128);
#pragma unroll
    for (int kb = 0; kb < NCH; ++kb) tma_load_2d(sW2 + kb * (C * 128), &tmW2, w_full, kb * 64, 0);
  }
  pdl_wait();

  if (warp >= WARP_CTRL) {
    reg_dec<L::REG_CTRL>();
    const int role = warp - WARP_CTRL;
    if (role == 0) {
      // (weights loader: its TMA loads were issued above, before pdl_wait)
    } else if (role == 1) {
      // ===== GEMM 1 issuer: D1[g & 1] = a * W1_chunk^T =====================================================
      if (elect_one()) {
        constexpr uint32_t idesc1 = umma_idesc_bf16(128, L::HC);
        const uint64_t da0 = umma_smem_desc<L::ROWB>(smem_u32(sA));
        const uint64_t dw0 = umma_smem_desc<L::ROWB>(smem_u32(sW1));
        mbar_wait(w_full, 0);
        int g = 0;
        for (int it = 0; it < my_tiles; ++it) {
          const int sa = it % NA;
          mbar_wait_sleep(&a_full[sa], (uint32_t)((it / NA) & 1), 64);
          const uint64_t da = da0 + (uint64_t)((sa * L::A_BYTES) >> 4);
#pragma unroll 1
          for (int j = 0; j < NCH; ++j, ++g) {
            const int db = g & 1;
            mbar_wait(&d1_empty[db], (uint32_t)(((g >> 1) & 1) ^ 1));
            tc_fence_after();
            const uint64_t dw = dw0 + (uint64_t)((j * L::HC * L::ROWB) >> 4);
            const uint32_t d = tmem_base + (uint32_t)(db * L::HC);
#pragma unroll
            for (int kk = 0; kk < C / 16; ++kk)
              umma_bf16(d, da + (uint64_t)(kk * 2), dw + (uint64_t)(kk * 2), idesc1, kk != 0);
            umma_commit(&d1_full[db]);
          }
          umma_commit(&a_empty[sa]);
        }
      }
    } else if (role == 2) {
      // ===== GEMM 2 issuer: D2[it & 1] += H * W2_chunk^T (A operand from tensor memory) ========================
      if (elect_one()) {
        constexpr uint32_t idesc2 = umma_idesc_bf16(128, C);
        const uint64_t dw0 = umma_smem_desc<128>(smem_u32(sW2));
        mbar_wait(w_full, 0);
        int g = 0;
        for (int it = 0; it < my_tiles; ++it) {
          const int d2b = it & 1;
          mbar_wait_sleep(&d2_empty[d2b], (uint32_t)(((it >> 1) & 1) ^ 1), 64);
          const uint32_t d = tmem_base + L::D2_COL0 + (uint32_t)(d2b * C);
#pragma unroll 1
          for (int j = 0; j < NCH; ++j, ++g) {
            const int hb = g & 1;
            mbar_wait(&h_full[hb], (uint32_t)((g >> 1) & 1));
            tc_fence_after();
            const uint32_t ah = tmem_base + L::H_COL0 + (uint32_t)(hb * 32);
            const uint64_t dw = dw0 + (uint64_t)((j * C * 128) >> 4);
#pragma unroll
            for (int kk = 0; kk < 4; ++kk)
              umma_bf16_ts(d, ah + (uint32_t)(kk * 8), dw + (uint64_t)(kk * 2), idesc2, (j | kk) != 0);
            umma_commit(&h_empty[hb]);
          }
          umma_commit(&d2_full[d2b]);
        }
      }
    } else {
      // ===== TMA store issuer: finished y slot -> global, then the slot goes back to the mixer =================
      if (elect_one()) {
        for (int it = 0; it < my_tiles; ++it) {
          const int sy = it % NY;
          const int tile = (int)blockIdx.x + it * (int)gridDim.x;
          const int b = tile / p.tiles_pb, t0 = (tile - b * p.tiles_pb) * 128;
          mbar_wait_sleep(&y_ready[sy], (uint32_t)((it / NY) & 1), 128);
#pragma unroll
          for (int c = 0; c < NBOX; ++c) tma_store_3d(&tmO, sY + sy * L::Y_BYTES + c * (128 * 128), c * 32, t0, b);
          tma_store_commit();
          tma_store_wait_read<0>();
          mbar_arrive(&y_empty[sy]);
        }
        tma_store_wait_all<0>();
      }
    }
  } else if (warp < WARP_E2) {
    // ===== E1: D1 -> + b1 -> GELU -> bf16 pairs -> H in tensor memory ===========================================
    // Two groups of 4 warps alternate chunks (group G owns D1 buffer G and H buffer G); a warp owns 32 rows and
    // walks the 64 columns of its chunk in two halves.
    reg_dec<L::REG_E>();
    const int q = warp & 3;
    const int grp = warp >> 2;
    const uint32_t lane_base = (uint32_t)(q * 32) << 16;
    const uint32_t d1addr = tmem_base + lane_base + (uint32_t)(grp * L::HC);
    const uint32_t haddr = tmem_base + lane_base + L::H_COL0 + (uint32_t)(grp * 32);
    for (int g = grp; g < total_g; g += 2) {
      const int j = g & (NCH - 1);
      const uint32_t ph = (uint32_t)((g >> 1) & 1);
      const float* bp = p.b1 + j * L::HC;
      mbar_wait(&d1_full[grp], ph);
      tc_fence_after();
#pragma unroll
      for (int hf = 0; hf < 2; ++hf) {
        uint32_t v[32];
        tmem_ld_32x32(d1addr + (uint32_t)(hf * 32), v);
        tmem_ld_wait();
        if (hf == 1) {                               // whole accumulator is in registers: the next G1 may overwrite it
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&d1_empty[grp]);
        }
        uint32_t hv[16];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float4 b0 = __ldg(reinterpret_cast<const float4*>(bp + hf * 32 + i * 8));
          const float4 b1v = __ldg(reinterpret_cast<const float4*>(bp + hf * 32 + i * 8 + 4));
          hv[i * 4 + 0] = gelu2_bias_pack(__uint_as_float(v[i * 8 + 0]), __uint_as_float(v[i * 8 + 1]), b0.x, b0.y);
          hv[i * 4 + 1] = gelu2_bias_pack(__uint_as_float(v[i * 8 + 2]), __uint_as_float(v[i * 8 + 3]), b0.z, b0.w);
          hv[i * 4 + 2] = gelu2_bias_pack(__uint_as_float(v[i * 8 + 4]), __uint_as_float(v[i * 8 + 5]), b1v.x, b1v.y);
          hv[i * 4 + 3] = gelu2_bias_pack(__uint_as_float(v[i * 8 + 6]), __uint_as_float(v[i * 8 + 7]), b1v.z, b1v.w);
        }
        if (hf == 0) {
          mbar_wait(&h_empty[grp], ph ^ 1);         // GEMM 2 has consumed this group's previous H chunk
          tc_fence_after();
        }
        tmem_st_32x16(haddr + (uint32_t)(hf * 16), hv);   // thread = row: pair (k, k+1) -> one 32-bit column
      }
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&h_full[grp]);
    }
  } else if (warp < WARP_MX) {
    // ===== E2: D2 -> + b2, * ffn_gamma, + y (staged by the mixer warps, swizzled) -> in place -> TMA store ========
    reg_dec<L::REG_E2>();
    const int q = warp & 3;
    const int r = q * 32 + lane;
    for (int it = 0; it < my_tiles; ++it) {
      const int d2b = it & 1, sy = it % NY;
      const int tile = (int)blockIdx.x + it * (int)gridDim.x;
      const int b = tile / p.tiles_pb, t = (tile - b * p.tiles_pb) * 128 + r;
      const bool valid = t < p.rows_pb;
      const long long brow = (long long)b * p.ob_rpb + p.ob_off + t;
      mbar_wait(&y_full[sy], (uint32_t)((it / NY) & 1));
      mbar_wait(&d2_full[d2b], (uint32_t)((it >> 1) & 1));
      tc_fence_after();
      uint8_t* rrow = sY + sy * L::Y_BYTES + r * 128;
#pragma unroll
      for (int c = 0; c < NBOX; ++c) {
        uint32_t v[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + L::D2_COL0 + (uint32_t)(d2b * C + c * 32), v);
        tmem_ld_wait();
        float x[32];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int col = c * 32 + i * 4;
          float4* cell = reinterpret_cast<float4*>(rrow + c * (128 * 128) + ((i ^ (r & 7)) * 16));
          const float4 rv = *cell;
          const float4 bv = __ldg(reinterpret_cast<const float4*>(p.b2 + col));
          const float4 gv = __ldg(reinterpret_cast<const float4*>(p.ffn_gamma + col));
          x[i * 4 + 0] = fmaf(gv.x, __uint_as_float(v[i * 4 + 0]) + bv.x, rv.x);
          x[i * 4 + 1] = fmaf(gv.y, __uint_as_float(v[i * 4 + 1]) + bv.y, rv.y);
          x[i * 4 + 2] = fmaf(gv.z, __uint_as_float(v[i * 4 + 2]) + bv.z, rv.z);
          x[i * 4 + 3] = fmaf(gv.w, __uint_as_float(v[i * 4 + 3]) + bv.w, rv.w);
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
      }
      tc_fence_before();
      fence_proxy_async_smem();                 // staged tile -> visible to the TMA store
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&d2_empty[d2b]);
        mbar_arrive(&y_ready[sy]);
      }
    }
  } else {
    // ===== mixer: x (global) -> y (staging slot) + a (GEMM 1 operand tile) ====================================
    reg_inc<L::REG_MX>();
    struct P4 { f32x2 lo, hi; };               // 4 adjacent channels as two packed fp32 pairs (FFMA2)
    constexpr int MXSETS = L::MXSETS;
    // this warp's set: tiles mset, mset + MXSETS, ...  (a compile-time 0 with one set: at C = 64 the mixer sits on its
    // register limit, and a run-time mset cost one more live register = 20 % on the kernel, measured)
    const int mset = (MXSETS > 1) ? (warp - WARP_MX) / L::MXW_TILE : 0;
    const int mwarp = (MXSETS > 1) ? (warp - WARP_MX) % L::MXW_TILE : (warp - WARP_MX);
    const int gi = mwarp * (32 / LPR) + lane / LPR;                  // strip 0..15: tile rows [8 gi, 8 gi + 8)
    const int c0 = (lane % LPR) * 4;
    const float inv_c = 1.0f / (float)C;
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
    // this thread's cells in a y slot / an a slot (row r = 8 gi + k, k added at the store)
    const uint32_t y_off = (uint32_t)((c0 >> 5) * (128 * 128) + gi * 1024);
    const int y_chunk = (c0 & 31) >> 2;
    const uint32_t a_off = (uint32_t)(gi * 8 * L::ROWB + (c0 & 7) * 2);
    const int a_chunk = c0 >> 3;

    const float* st_rd = nullptr;
    float* st_wr = nullptr;
    auto tile_of = [&](int it, int& b, int& t0) {
      const int tile = (int)blockIdx.x + it * (int)gridDim.x;
      b = tile / p.tiles_pb;
      t0 = (tile - b * p.tiles_pb) * 128;
    };
    // batch 0 of a tile = rows 8gi-6 .. 8gi (6 warm-up rows + the first output row); batch 1 = rows 8gi+1 .. 8gi+7
    auto load_rows = [&](int it, int b, int t0, int bt, float4 (&dst)[7]) {
      if (it >= my_tiles) return;                                    // (nothing will consume dst)
      const int ts = t0 + gi * 8 + (bt ? 1 : -6);
      const float* xb = p.x + ((long long)b * p.rows_pb + ts) * C + c0;
      if (ts >= 0 && ts + 7 <= p.rows_pb) {                          // interior: 7 unpredicated loads off one pointer
#pragma unroll
        for (int i = 0; i < 7; ++i) dst[i] = __ldg(reinterpret_cast<const float4*>(xb + i * C));
        return;
      }
      const float* srd = nullptr;
      if (STATE && ts < 0 && p.st.base != nullptr) {
        const uint8_t* rd8;
        uint8_t* wr8;
        p.st.resolve(b, rd8, wr8);
        srd = rd8 ? reinterpret_cast<const float*>(rd8) + c0 : nullptr;
      }
#pragma unroll
      for (int i = 0; i < 7; ++i) {
        const int t = ts + i;
        dst[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (t < p.rows_pb) {
          if (t >= 0) dst[i] = __ldg(reinterpret_cast<const float4*>(xb + i * C));
          else if (srd != nullptr) dst[i] = *reinterpret_cast<const float4*>(srd + (6 + t) * C);
        }
      }
    };
    auto dot4 = [](const P4& v) -> float {
      float s0, s1;
      f2_unpack(f2_fma(v.hi, v.hi, f2_mul(v.lo, v.lo)), s0, s1);
      return s0 + s1;
    };
    P4 win[7];
#pragma unroll
    for (int j = 0; j < 7; ++j) { win[j].lo = 0ull; win[j].hi = 0ull; }

    // one output row: window slot i holds n[t]; taps read slots (i+1+j) % 7 = n[t-6+j].  Returns y in `row`.
    auto conv_row = [&](int i, P4& row, float rstd) {
      const f32x2 rr = f2_pack(rstd, rstd);
      win[i].lo = f2_mul(row.lo, rr);
      win[i].hi = f2_mul(row.hi, rr);
      P4 acc;
      acc.lo = f2_add(row.lo, gb.lo);
      acc.hi = f2_add(row.hi, gb.hi);
#pragma unroll
      for (int j = 0; j < 7; ++j) {
        acc.lo = f2_fma(wq[j].lo, win[(i + 1 + j) % 7].lo, acc.lo);
        acc.hi = f2_fma(wq[j].hi, win[(i + 1 + j) % 7].hi, acc.hi);
      }
      row = acc;
    };
    // y row -> staging slot (fp32, swizzled float4 cell);  a = bf16(y * rstd2 * ffn_norm_w) -> operand tile
    auto store_row = [&](int k, const P4& y, float rstd2, const P4& fw, uint8_t* ys, uint8_t* as) {
      *reinterpret_cast<ulonglong2*>(ys + y_off + k * 128 + ((y_chunk ^ k) << 4)) = make_ulonglong2(y.lo, y.hi);
      const f32x2 rr = f2_pack(rstd2, rstd2);
      float a0, a1, a2, a3;
      f2_unpack(f2_mul(f2_mul(y.lo, rr), fw.lo), a0, a1);
      f2_unpack(f2_mul(f2_mul(y.hi, rr), fw.hi), a2, a3);
      uint2 pk;
      pk.x = pack_bf16x2(a0, a1);
      pk.y = pack_bf16x2(a2, a3);
      const int sw = (L::ROWB == 128) ? (a_chunk ^ k) : (a_chunk ^ ((k >> 1) & 3));
      *reinterpret_cast<uint2*>(as + a_off + k * L::ROWB + (sw << 4)) = pk;
    };
    // the stream's new context = last 6 raw rows of [old context, x]  (reference :371-382)
    auto state_row = [&](int t, const float4& raw) {
      const int dst = t - (p.rows_pb - 6);
      if (dst >= 0 && t < p.rows_pb) *reinterpret_cast<float4*>(st_wr + dst * C) = raw;
    };

    // (tried and measured slower on B200: the folded taps read back from shared memory per use, 1.87 -> 2.21 ms at
    //  C = 64; the next-tile prefetch staged through shared memory with cp.async instead of registers, 1.87 -> 2.03 ms)
    float4 bufA[7], bufB[7];
    int b, t0, nb = 0, nt0 = 0;
    tile_of(mset, b, t0);
    load_rows(mset, b, t0, 0, bufA);
    for (int it = mset; it < my_tiles; it += MXSETS) {
      const int sa = it % NA, sy = it % NY;
      if (it > mset) { b = nb; t0 = nt0; }
      tile_of(it + MXSETS, nb, nt0);
      if (STATE && p.st.base != nullptr) {
        const uint8_t* rd8;
        uint8_t* wr8;
        p.st.resolve(b, rd8, wr8);
        st_rd = reinterpret_cast<const float*>(rd8);
        st_wr = wr8 ? reinterpret_cast<float*>(wr8) + c0 : nullptr;
      }
      uint8_t* ys = sY + sy * L::Y_BYTES;
      uint8_t* as = sA + sa * L::A_BYTES;
      const float4 fw4 = __ldg(reinterpret_cast<const float4*>(p.ffn_norm_w + c0));
      P4 fw;
      fw.lo = f2_pack(fw4.x, fw4.y);
      fw.hi = f2_pack(fw4.z, fw4.w);
      load_rows(it, b, t0, 1, bufB);                       // second batch of this tile: in flight during batch 0

      // ---- batch 0: six warm-up rows rebuild the window, then output row k = 0 ---------------------------------
      {
        const int ts = t0 + gi * 8 - 6;
        P4 rows[7];
        float ss[7];
#pragma unroll
        for (int i = 0; i < 7; ++i) {
          if (STATE && st_wr != nullptr && (i == 6 || ts + i < 0)) state_row(ts + i, bufA[i]);
          rows[i].lo = f2_pack(bufA[i].x, bufA[i].y);
          rows[i].hi = f2_pack(bufA[i].z, bufA[i].w);
          ss[i] = dot4(rows[i]);
        }
        group_sum7<LPR>(ss);
#pragma unroll
        for (int i = 0; i < 6; ++i) {
          const float rstd = rstd_of(ss[i], inv_c, p.eps);
          const f32x2 rr = f2_pack(rstd, rstd);
          win[i].lo = f2_mul(rows[i].lo, rr);
          win[i].hi = f2_mul(rows[i].hi, rr);
        }
        conv_row(6, rows[6], rstd_of(ss[6], inv_c, p.eps));
        const float s2 = group_sum1<LPR>(dot4(rows[6]));
        // the slots of this tile must be free before the first store
        mbar_wait(&a_empty[sa], (uint32_t)(((it / NA) & 1) ^ 1));
        mbar_wait(&y_empty[sy], (uint32_t)(((it / NY) & 1) ^ 1));
        store_row(0, rows[6], rstd_of(s2, inv_c, p.eps), fw, ys, as);
      }
      load_rows(it + MXSETS, nb, nt0, 0, bufA);            // first batch of this set's next tile: in flight during batch 1

      // ---- batch 1: output rows k = 1..7 ------------------------------------------------------------------------
      {
        const int ts = t0 + gi * 8 + 1;
        P4 rows[7];
        float ss[7];
#pragma unroll
        for (int i = 0; i < 7; ++i) {
          if (STATE && st_wr != nullptr) state_row(ts + i, bufB[i]);
          rows[i].lo = f2_pack(bufB[i].x, bufB[i].y);
          rows[i].hi = f2_pack(bufB[i].z, bufB[i].w);
          ss[i] = dot4(rows[i]);
        }
        group_sum7<LPR>(ss);
#pragma unroll
        for (int i = 0; i < 7; ++i) {
          conv_row(i, rows[i], rstd_of(ss[i], inv_c, p.eps));
          ss[i] = dot4(rows[i]);
        }
        group_sum7<LPR>(ss);
#pragma unroll
        for (int i = 0; i < 7; ++i) store_row(i + 1, rows[i], rstd_of(ss[i], inv_c, p.eps), fw, ys, as);
      }
      fence_proxy_async_smem();                 // a tile (generic-proxy stores) -> visible to tcgen05.mma
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&a_full[sa]);
        mbar_arrive(&y_full[sy]);
      }
    }
    (void)st_rd;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == WARP_CTRL + 1) tmem_dealloc(tmem_base, L::TMEM_COLS);
}

}  // namespace vv
