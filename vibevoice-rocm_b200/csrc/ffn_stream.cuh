// ffn_stream.cuh — kernel class (b) for the middle stage (C = 128): the whole Block1D FFN half in ONE persistent
// tcgen05 kernel with the hidden activation kept on chip and the weights STREAMED through shared memory.
//
//   out = resid + ffn_gamma * ( W2 * gelu_erf( W1 * a + b1 ) + b2 )         (reference :592-596, :937-942)
//
// At C = 128 the two-GEMM form is HBM-bound by the [rows, 4C] hidden tensor the reference materialises
// (FFN.forward :594-596): 16 B per activation element out and back in, against 10 B for everything else.  W1 and
// W2 (128 KB each) do not fit in shared memory next to the tiles, so per 128-row tile the hidden dimension is
// walked in NCH chunks of HC = 64 units and the weight chunks arrive through two TMA rings (they are L2-resident:
// 256 KB per block):
//   TMA :  a tile [128, C] bf16 (ring of NA);  W1 chunk [64, C] (ring S1);  W2 chunk [C, 64] (ring S2);
//          residual in boxes of [128 rows, 32 fp32] (ring NY) — fine-grained so 2-3 boxes cover the HBM latency
//   MMA :  D1[128, 64] = a * W1_chunk^T  (2 TMEM buffers);   D2[128, C] += H * W2_chunk^T  (2 TMEM buffers)
//   E1  :  16 warps: tcgen05.ld D1 -> + b1 -> GELU -> bf16 -> 128B-swizzled smem H (A operand of GEMM 2), 2 buffers
//   E2  :  4 warps, box by box: tcgen05.ld D2 -> + b2, * gamma, + residual (in the staged box, in place) -> TMA store
// The single MMA thread issues G1(g+1) before G2(g) so the tensor pipe has work while E1 converts chunk g.
// All hand-offs are mbarriers; there is no __syncthreads in the loop.
#pragma once
#include <cuda.h>
#include "ptx.cuh"
#include "epilogue.cuh"
#include "ffn_fused.cuh"   // FfnFusedArgs

namespace vv {

constexpr int FS_E1_WARPS = 16;
constexpr int FS_E2_WARPS = 4;
constexpr int FS_THREADS = (4 + FS_E1_WARPS + FS_E2_WARPS) * 32;   // warps 0..3: TMA a/W, MMA, TMA residual, TMA store

template <int C>
struct FfnStreamCfg {
  static_assert(C == 128, "ffn_stream_kernel is laid out for C = 128");
  static constexpr int HC = 64;                    // hidden units per chunk
  static constexpr int NCH = 4 * C / HC;           // chunks per tile
  static constexpr int KB = C / 64;                // 64-wide K blocks of the a tile / a W1 chunk
  static constexpr int NBOX = C / 32;              // residual / output boxes per tile
  static constexpr int NA = 2, S1 = 3, S2 = 3, NY = 2;
  static constexpr int A_BYTES = 128 * C * 2;      // KB blocks of [128 rows x 128 B]
  static constexpr int W1_BYTES = HC * C * 2;      // KB blocks of [64 rows x 128 B]
  static constexpr int W2_BYTES = C * HC * 2;      // [C rows x 128 B]
  static constexpr int H_BYTES = 128 * HC * 2;     // [128 rows x 128 B]
  static constexpr int Y_BYTES = 128 * 128;        // [128 rows x 32 fp32]
  static constexpr int OFF_A = 0;
  static constexpr int OFF_W1 = OFF_A + NA * A_BYTES;
  static constexpr int OFF_W2 = OFF_W1 + S1 * W1_BYTES;
  static constexpr int OFF_H = OFF_W2 + S2 * W2_BYTES;
  static constexpr int OFF_Y = OFF_H + 2 * H_BYTES;
  static constexpr int OFF_BAR = OFF_Y + NY * Y_BYTES;
  static constexpr int NBARS = 2 * NA + 2 * S1 + 2 * S2 + 12 + 3 * NY;
  static constexpr int TOTAL = OFF_BAR + 512 + 1024;   // + alignment slack
  static_assert(NBARS * 8 + 8 <= 512, "barrier block");
  static_assert(TOTAL <= 232448, "exceeds the 227 KB dynamic shared memory limit");
  static constexpr uint32_t TMEM_COLS = 512;       // D1: 2 x 64 at column 0; D2: 2 x C at column 128
  static constexpr uint32_t D2_COL0 = 128;
};

template <int C>
__global__ void __launch_bounds__(FS_THREADS, 1)
ffn_stream_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmR,
                  const __grid_constant__ CUtensorMap tmO, const __grid_constant__ CUtensorMap tmW1,
                  const __grid_constant__ CUtensorMap tmW2, const FfnFusedArgs p) {
  using L = FfnStreamCfg<C>;
  constexpr int NA = L::NA, S1 = L::S1, S2 = L::S2, NY = L::NY, NCH = L::NCH, KB = L::KB, NBOX = L::NBOX;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* sA = smem + L::OFF_A;
  uint8_t* sW1 = smem + L::OFF_W1;
  uint8_t* sW2 = smem + L::OFF_W2;
  uint8_t* sH = smem + L::OFF_H;
  uint8_t* sY = smem + L::OFF_Y;
  uint64_t* bars = (uint64_t*)(smem + L::OFF_BAR);
  uint64_t* a_full = bars;                 // [NA]
  uint64_t* a_empty = a_full + NA;         // [NA]
  uint64_t* w1_full = a_empty + NA;        // [S1]
  uint64_t* w1_empty = w1_full + S1;       // [S1]
  uint64_t* w2_full = w1_empty + S1;       // [S2]
  uint64_t* w2_empty = w2_full + S2;       // [S2]
  uint64_t* d1_full = w2_empty + S2;       // [2]
  uint64_t* d1_empty = d1_full + 2;        // [2]
  uint64_t* h_full = d1_empty + 2;         // [2]
  uint64_t* h_empty = h_full + 2;          // [2]
  uint64_t* d2_full = h_empty + 2;         // [2]
  uint64_t* d2_empty = d2_full + 2;        // [2]
  uint64_t* y_full = d2_empty + 2;         // [NY] residual box landed
  uint64_t* y_ready = y_full + NY;         // [NY] E2 finished the box: ready to store
  uint64_t* y_empty = y_ready + NY;        // [NY] the TMA store has read the box
  uint32_t* tmem_slot = (uint32_t*)(y_empty + NY);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const long long ntiles = (p.R + 127) / 128;
  const long long my_tiles = (ntiles > blockIdx.x) ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  const long long total_g = my_tiles * NCH;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmR);
    tma_prefetch_desc(&tmO);
    tma_prefetch_desc(&tmW1);
    tma_prefetch_desc(&tmW2);
    for (int i = 0; i < NA; ++i) { mbar_init(&a_full[i], 1); mbar_init(&a_empty[i], 1); }
    for (int i = 0; i < S1; ++i) { mbar_init(&w1_full[i], 1); mbar_init(&w1_empty[i], 1); }
    for (int i = 0; i < S2; ++i) { mbar_init(&w2_full[i], 1); mbar_init(&w2_empty[i], 1); }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&d1_full[i], 1);
      mbar_init(&d1_empty[i], FS_E1_WARPS);
      mbar_init(&h_full[i], FS_E1_WARPS);
      mbar_init(&h_empty[i], 1);
      mbar_init(&d2_full[i], 1);
      mbar_init(&d2_empty[i], FS_E2_WARPS);
    }
    for (int i = 0; i < NY; ++i) {
      mbar_init(&y_full[i], 1);
      mbar_init(&y_ready[i], FS_E2_WARPS);
      mbar_init(&y_empty[i], 1);
    }
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

  if (warp == 0) {
    // ===== TMA producer: a tiles and weight chunks ===================================================
    if (elect_one()) {
      long long g = 0;
      for (long long it = 0; it < my_tiles; ++it) {
        const long long tile = blockIdx.x + it * gridDim.x;
        const int sa = (int)(it % NA);
        mbar_wait(&a_empty[sa], (uint32_t)(((it / NA) & 1) ^ 1));
        mbar_expect_tx(&a_full[sa], L::A_BYTES);
#pragma unroll
        for (int kb = 0; kb < KB; ++kb)
          tma_load_2d(sA + sa * L::A_BYTES + kb * (128 * 128), &tmA, &a_full[sa], kb * 64, (int)(tile * 128));
        for (int j = 0; j < NCH; ++j, ++g) {
          const int s1 = (int)(g % S1), s2 = (int)(g % S2);
          mbar_wait(&w1_empty[s1], (uint32_t)(((g / S1) & 1) ^ 1));
          mbar_expect_tx(&w1_full[s1], L::W1_BYTES);
#pragma unroll
          for (int kb = 0; kb < KB; ++kb)
            tma_load_2d(sW1 + s1 * L::W1_BYTES + kb * (L::HC * 128), &tmW1, &w1_full[s1], kb * 64, j * L::HC);
          mbar_wait(&w2_empty[s2], (uint32_t)(((g / S2) & 1) ^ 1));
          mbar_expect_tx(&w2_full[s2], L::W2_BYTES);
          tma_load_2d(sW2 + s2 * L::W2_BYTES, &tmW2, &w2_full[s2], j * L::HC, 0);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer ===============================================================================
    if (elect_one()) {
      constexpr uint32_t idesc1 = umma_idesc_bf16(128, L::HC);
      constexpr uint32_t idesc2 = umma_idesc_bf16(128, C);
      for (long long g = 0; g <= total_g; ++g) {
        if (g < total_g) {
          const long long it = g / NCH;
          const int j = (int)(g % NCH), sa = (int)(it % NA), db = (int)(g & 1), s1 = (int)(g % S1);
          mbar_wait(&d1_empty[db], (uint32_t)(((g >> 1) & 1) ^ 1));
          if (j == 0) mbar_wait(&a_full[sa], (uint32_t)((it / NA) & 1));
          mbar_wait(&w1_full[s1], (uint32_t)((g / S1) & 1));
          tc_fence_after();
#pragma unroll
          for (int kb = 0; kb < KB; ++kb) {
            const uint64_t da = umma_smem_desc<128>(smem_u32(sA + sa * L::A_BYTES + kb * (128 * 128)));
            const uint64_t dw = umma_smem_desc<128>(smem_u32(sW1 + s1 * L::W1_BYTES + kb * (L::HC * 128)));
#pragma unroll
            for (int kk = 0; kk < 4; ++kk)
              umma_bf16(tmem_base + (uint32_t)(db * L::HC), da + (uint64_t)(kk * 2), dw + (uint64_t)(kk * 2), idesc1,
                        (kb | kk) != 0);
          }
          umma_commit(&d1_full[db]);
          umma_commit(&w1_empty[s1]);
          if (j == NCH - 1) umma_commit(&a_empty[sa]);
        }
        if (g >= 1) {
          const long long gp = g - 1, it = gp / NCH;
          const int j = (int)(gp % NCH), hb = (int)(gp & 1), d2b = (int)(it & 1), s2 = (int)(gp % S2);
          mbar_wait(&h_full[hb], (uint32_t)((gp >> 1) & 1));
          mbar_wait(&w2_full[s2], (uint32_t)((gp / S2) & 1));
          if (j == 0) mbar_wait(&d2_empty[d2b], (uint32_t)(((it >> 1) & 1) ^ 1));
          tc_fence_after();
          const uint64_t dh = umma_smem_desc<128>(smem_u32(sH + hb * L::H_BYTES));
          const uint64_t dw = umma_smem_desc<128>(smem_u32(sW2 + s2 * L::W2_BYTES));
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)
            umma_bf16(tmem_base + L::D2_COL0 + (uint32_t)(d2b * C), dh + (uint64_t)(kk * 2), dw + (uint64_t)(kk * 2),
                      idesc2, (j | kk) != 0);
          umma_commit(&h_empty[hb]);
          umma_commit(&w2_empty[s2]);
          if (j == NCH - 1) umma_commit(&d2_full[d2b]);
        }
      }
    }
  } else if (warp == 2) {
    // ===== TMA producer: residual boxes ================================================================
    if (elect_one()) {
      long long k = 0;
      for (long long it = 0; it < my_tiles; ++it) {
        const long long tile = blockIdx.x + it * gridDim.x;
        for (int c = 0; c < NBOX; ++c, ++k) {
          const int sy = (int)(k % NY);
          mbar_wait(&y_empty[sy], (uint32_t)(((k / NY) & 1) ^ 1));
          mbar_expect_tx(&y_full[sy], L::Y_BYTES);
          tma_load_2d(sY + sy * L::Y_BYTES, &tmR, &y_full[sy], c * 32, (int)(tile * 128));
        }
      }
    }
  } else if (warp == 3) {
    // ===== TMA store issuer: finished box -> global, then the slot goes back to the residual ring =========
    if (elect_one()) {
      long long k = 0;
      for (long long it = 0; it < my_tiles; ++it) {
        const long long tile = blockIdx.x + it * gridDim.x;
        for (int c = 0; c < NBOX; ++c, ++k) {
          const int sy = (int)(k % NY);
          mbar_wait(&y_ready[sy], (uint32_t)((k / NY) & 1));
          tma_store_2d(&tmO, sY + sy * L::Y_BYTES, c * 32, (int)(tile * 128));
          tma_store_commit();
          tma_store_wait_read<0>();
          mbar_arrive(&y_empty[sy]);
        }
      }
      tma_store_wait_all<0>();
    }
  } else if (warp < 4 + FS_E1_WARPS) {
    // ===== E1: D1 -> bias + GELU -> bf16 -> swizzled smem (A operand of GEMM 2) ============================
    const int q = warp & 3;                   // TMEM lane quarter
    const int cg = (warp - 4) >> 2;           // 16-column group of the 64-column chunk (0..3)
    const int r = q * 32 + lane;              // row in the tile
    uint8_t* hrow_base = sH + (r >> 3) * 1024 + (r & 7) * 128;
    for (long long g = 0; g < total_g; ++g) {
      const int j = (int)(g % NCH), db = (int)(g & 1);
      const float* bp = p.b1 + j * L::HC + cg * 16;
      const float4 b0 = __ldg(reinterpret_cast<const float4*>(bp));
      const float4 b1v = __ldg(reinterpret_cast<const float4*>(bp + 4));
      const float4 b2v = __ldg(reinterpret_cast<const float4*>(bp + 8));
      const float4 b3v = __ldg(reinterpret_cast<const float4*>(bp + 12));
      mbar_wait(&d1_full[db], (uint32_t)((g >> 1) & 1));
      tc_fence_after();
      uint32_t v[16];
      tmem_ld_32x16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(db * L::HC + cg * 16), v);
      tmem_ld_wait();
      tc_fence_before();
      uint4 pk0, pk1;
      pk0.x = gelu2_bias_pack(__uint_as_float(v[0]), __uint_as_float(v[1]), b0.x, b0.y);
      pk0.y = gelu2_bias_pack(__uint_as_float(v[2]), __uint_as_float(v[3]), b0.z, b0.w);
      pk0.z = gelu2_bias_pack(__uint_as_float(v[4]), __uint_as_float(v[5]), b1v.x, b1v.y);
      pk0.w = gelu2_bias_pack(__uint_as_float(v[6]), __uint_as_float(v[7]), b1v.z, b1v.w);
      pk1.x = gelu2_bias_pack(__uint_as_float(v[8]), __uint_as_float(v[9]), b2v.x, b2v.y);
      pk1.y = gelu2_bias_pack(__uint_as_float(v[10]), __uint_as_float(v[11]), b2v.z, b2v.w);
      pk1.z = gelu2_bias_pack(__uint_as_float(v[12]), __uint_as_float(v[13]), b3v.x, b3v.y);
      pk1.w = gelu2_bias_pack(__uint_as_float(v[14]), __uint_as_float(v[15]), b3v.z, b3v.w);
      mbar_wait(&h_empty[db], (uint32_t)(((g >> 1) & 1) ^ 1));
      uint8_t* hrow = hrow_base + db * L::H_BYTES;
      *reinterpret_cast<uint4*>(hrow + (((cg * 2 + 0) ^ (r & 7)) * 16)) = pk0;   // 128B swizzle: chunk XOR row-in-atom
      *reinterpret_cast<uint4*>(hrow + (((cg * 2 + 1) ^ (r & 7)) * 16)) = pk1;
      fence_proxy_async_smem();                 // generic-proxy smem writes -> visible to tcgen05.mma
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&d1_empty[db]);
        mbar_arrive(&h_full[db]);
      }
    }
  } else {
    // ===== E2: D2 -> + b2, * gamma, + residual (TMA-staged box, swizzled) -> in place -> TMA store ===========
    const int q = warp & 3;
    const int r = q * 32 + lane;
    long long k = 0;
    for (long long it = 0; it < my_tiles; ++it) {
      const int d2b = (int)(it & 1);
      const long long tile = blockIdx.x + it * gridDim.x;
      const long long m = tile * 128 + r;
      const bool valid = m < p.R;
      long long brow = 0;
      if (p.out_bf16 != nullptr && valid) {
        const long long b = m / p.rows_pb;
        brow = b * (p.rows_pb + 1) + 1 + (m - b * p.rows_pb);
      }
      mbar_wait(&d2_full[d2b], (uint32_t)((it >> 1) & 1));
      tc_fence_after();
#pragma unroll 1
      for (int c = 0; c < NBOX; ++c, ++k) {
        const int sy = (int)(k % NY);
        uint32_t v[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + L::D2_COL0 + (uint32_t)(d2b * C + c * 32), v);
        mbar_wait(&y_full[sy], (uint32_t)((k / NY) & 1));
        tmem_ld_wait();
        uint8_t* rrow = sY + sy * L::Y_BYTES + r * 128;
        float x[32];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int col = c * 32 + i * 4;
          float4* cell = reinterpret_cast<float4*>(rrow + ((i ^ (r & 7)) * 16));
          const float4 rv = *cell;
          const float4 bv = __ldg(reinterpret_cast<const float4*>(p.b2 + col));
          const float4 gv = __ldg(reinterpret_cast<const float4*>(p.gamma + col));
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
