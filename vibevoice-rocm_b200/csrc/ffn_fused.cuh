// ffn_fused.cuh — kernel class (b) for the narrow stages (C = 32, 64): the whole Block1D FFN half in ONE
// persistent tcgen05 kernel, hidden activation never leaves the SM.
//
//   out = resid + ffn_gamma * ( W2 * gelu_erf( W1 * a + b1 ) + b2 )         (reference :592-596, :937-942)
//
// The reference materialises the [rows, 4C] hidden tensor in HBM (FFN.forward :594-596); at C <= 64 that
// traffic (16 B per activation element) is larger than everything else the block moves, so these stages
// are HBM-bound by it.  Here, per 128-row tile:
//   TMA:  a tile [128, C] bf16  -> smem (2-deep ring);  W1, W2 are loaded once per CTA and stay resident
//   MMA:  D1[128, 128-col chunk] = a * W1_chunk^T          (tcgen05.mma, accumulator in TMEM, 2 buffers)
//   E1 : 16 warps: tcgen05.ld D1 -> +b1 -> GELU -> bf16 -> st.shared into the 128B-swizzled K-major layout
//         that tcgen05.mma reads as an A operand (H ring, 2 x [128, 128])
//   MMA:  D2[128, C] += H * W2_chunk^T                      (TMEM, 2 buffers across tiles)
//   E2 :  4 warps: tcgen05.ld D2 -> +b2, * gamma, + resid (fp32 tile staged by TMA) -> in place in smem ->
//         TMA store of the fp32 tile (+ haloed bf16 copy for the upsampler when this is the last block of a stage)
// The single MMA thread issues G1(g+1) before G2(g) (g = global 128-column chunk index), so the tensor
// pipe always has work while E1 converts chunk g; all hand-offs are mbarriers, no __syncthreads in the loop.
#pragma once
#include <cuda.h>
#include "ptx.cuh"
#include "epilogue.cuh"

namespace vv {

constexpr int FF_THREADS = 736;      // warp 0 TMA loads, warp 1 MMA, warps 2..17 E1, warps 18..21 E2, warp 22 TMA stores
constexpr int FF_E1_WARPS = 16;
constexpr int FF_E2_WARPS = 4;

struct FfnFusedArgs {
  long long R;             // total rows (batch * rows_per_batch)
  int rows_pb;             // rows per batch item (for the haloed bf16 copy)
  const float* b1;         // [4C]
  const float* b2;         // [C]
  const float* gamma;      // [C]
  const float* resid;      // [R, C] fp32
  float* out;              // [R, C] fp32 (may alias resid)
  int inplace_add;         // ffn_stream_kernel only: out == resid and the kernel ADDS gamma * (acc + b2) to it with TMA
                           // reductions (no residual is read by the SM; out_bf16 must be null)
  __nv_bfloat16* out_bf16; // optional padded bf16 copy: [batch, ob_rpb, C], row t of an item lands at row ob_off + t
  long long ob_rpb;        //   (decoder: ob_rpb = rows_pb + 1, ob_off = 1 = the upsampler halo row)
  int ob_off;
  // ffn_stream_kernel (C = 128) reads the three per-channel vectors from the CONSTANT bank (by-value argument) instead of
  // warp-uniform __ldg loads: ncu of block_tma_kernel showed such loads costing ~4 LSU data-pipe wavefronts each, a third
  // of that kernel's LSU traffic (profiles/block_tma_s6_r02.txt).  Filled by the host when C == 128.
  float b1c[1024];         // = b1 (pre-scaled by 0.5); sized for C = 256
  float b2c[256];
  float gammac[256];
};

template <int C>
struct FfnFusedSmem {
  static constexpr int NA = (C == 32) ? 4 : 2;             // a-tile ring depth (covers TMA latency per tile time)
  static constexpr int NR = (C == 32) ? 4 : 2;             // residual-tile ring depth
  static constexpr int ROWB = C * 2;                       // bytes per a / W1 row (64 or 128)
  static constexpr int W1_BYTES = 4 * C * ROWB;            // [4C, C]
  static constexpr int W2_KB = 4 * C / 64;                 // K-blocks of 64 hidden units
  static constexpr int W2_KB_BYTES = C * 128;              // [C, 64] per K-block
  static constexpr int W2_BYTES = W2_KB * W2_KB_BYTES;
  static constexpr int A_BYTES = 128 * ROWB;
  static constexpr int H_BYTES = 128 * 128 * 2;            // two [128, 64] K-blocks
  static constexpr int R_BOX_BYTES = 128 * 128;            // [128 rows, 32 floats], 128B-swizzled
  static constexpr int R_BYTES = (C / 32) * R_BOX_BYTES;
  static constexpr int OFF_W1 = 0;
  static constexpr int OFF_W2 = OFF_W1 + W1_BYTES;
  static constexpr int OFF_A = OFF_W2 + W2_BYTES;
  static constexpr int OFF_H = OFF_A + NA * A_BYTES;
  static constexpr int OFF_R = OFF_H + 2 * H_BYTES;
  static constexpr int OFF_B1 = OFF_R + NR * R_BYTES;      // b1 staged once: [4C] floats
  static constexpr int OFF_BAR = OFF_B1 + 4 * C * 4;
  static constexpr int TOTAL = OFF_BAR + 512 + 1024;       // + alignment slack
  static_assert(TOTAL <= 232448, "exceeds the 227 KB dynamic shared memory limit");
};

template <int C>
__global__ void __launch_bounds__(FF_THREADS, 1)
ffn_fused_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmR,
                 const __grid_constant__ CUtensorMap tmO, const __grid_constant__ CUtensorMap tmW1,
                 const __grid_constant__ CUtensorMap tmW2, const FfnFusedArgs p) {
  using L = FfnFusedSmem<C>;
  constexpr int NA = L::NA, NR = L::NR;
  constexpr int NCHUNK = 4 * C / 128;                      // 128-column chunks of the hidden dim per tile
  constexpr uint32_t TMEM_COLS = 512;
  constexpr uint32_t D2_COL0 = 256;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_align_1k(smem_raw);
  uint8_t* sW1 = smem + L::OFF_W1;
  uint8_t* sW2 = smem + L::OFF_W2;
  uint8_t* sA = smem + L::OFF_A;
  uint8_t* sH = smem + L::OFF_H;
  uint8_t* sR = smem + L::OFF_R;
  float* sB1 = (float*)(smem + L::OFF_B1);
  uint64_t* bars = (uint64_t*)(smem + L::OFF_BAR);
  uint64_t* w_full = bars + 0;
  uint64_t* a_full = bars + 1;            // [NA]
  uint64_t* a_empty = a_full + NA;        // [NA]
  uint64_t* r_full = a_empty + NA;        // [NR]
  uint64_t* r_empty = r_full + NR;        // [NR]
  uint64_t* o_full = r_empty + NR;        // [NR]  output staged in the residual slot, ready to store
  uint64_t* d1_full = o_full + NR;        // [2]
  uint64_t* d1_empty = d1_full + 2;       // [2]
  uint64_t* h_full = d1_empty + 2;        // [2]
  uint64_t* h_empty = h_full + 2;         // [2]
  uint64_t* d2_full = h_empty + 2;        // [2]
  uint64_t* d2_empty = d2_full + 2;       // [2]
  uint32_t* tmem_slot = (uint32_t*)(d2_empty + 2);
  static_assert((1 + 2 * NA + 3 * NR + 12 + 1) * 8 <= 512, "barrier block");

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const long long ntiles = (p.R + 127) / 128;
  // tiles of this CTA: blockIdx.x, blockIdx.x + gridDim.x, ...
  const long long my_tiles = (ntiles > blockIdx.x) ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  const long long total_g = my_tiles * NCHUNK;

  for (int i = threadIdx.x; i < 4 * C; i += FF_THREADS) sB1[i] = p.b1[i];
  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmR);
    tma_prefetch_desc(&tmO);
    tma_prefetch_desc(&tmW1);
    tma_prefetch_desc(&tmW2);
    mbar_init(w_full, 1);
    for (int i = 0; i < NA; ++i) { mbar_init(&a_full[i], 1); mbar_init(&a_empty[i], 1); }
    for (int i = 0; i < NR; ++i) {
      mbar_init(&r_full[i], 1);
      mbar_init(&r_empty[i], 1);
      mbar_init(&o_full[i], FF_E2_WARPS);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&d1_full[i], 1);
      mbar_init(&d1_empty[i], FF_E1_WARPS);
      mbar_init(&h_full[i], FF_E1_WARPS);
      mbar_init(&h_empty[i], 1);
      mbar_init(&d2_full[i], 1);
      mbar_init(&d2_empty[i], FF_E2_WARPS);
    }
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch();
  pdl_wait();

  if (warp == 0) {
    // ===== TMA producer ==========================================================================
    if (elect_one()) {
      mbar_expect_tx(w_full, L::W1_BYTES + L::W2_BYTES);
#pragma unroll
      for (int j = 0; j < NCHUNK; ++j) tma_load_2d(sW1 + j * 128 * L::ROWB, &tmW1, w_full, 0, j * 128);
#pragma unroll
      for (int kb = 0; kb < L::W2_KB; ++kb) tma_load_2d(sW2 + kb * L::W2_KB_BYTES, &tmW2, w_full, kb * 64, 0);
      for (long long it = 0; it < my_tiles; ++it) {
        const long long tile = blockIdx.x + it * gridDim.x;
        const int sa = (int)(it % NA), sr = (int)(it % NR);
        mbar_wait(&a_empty[sa], (uint32_t)(((it / NA) & 1) ^ 1));
        mbar_expect_tx(&a_full[sa], L::A_BYTES);
        tma_load_2d(sA + sa * L::A_BYTES, &tmA, &a_full[sa], 0, (int)(tile * 128));
        mbar_wait(&r_empty[sr], (uint32_t)(((it / NR) & 1) ^ 1));
        mbar_expect_tx(&r_full[sr], L::R_BYTES);
#pragma unroll
        for (int c = 0; c < C / 32; ++c)
          tma_load_2d(sR + sr * L::R_BYTES + c * L::R_BOX_BYTES, &tmR, &r_full[sr], c * 32, (int)(tile * 128));
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer ===============================================================================
    if (elect_one()) {
      constexpr uint32_t idesc1 = umma_idesc_bf16(128, 128);
      constexpr uint32_t idesc2 = umma_idesc_bf16(128, C);
      mbar_wait(w_full, 0);
      for (long long g = 0; g <= total_g; ++g) {
        if (g < total_g) {
          const long long it = g / NCHUNK;
          const int j = (int)(g % NCHUNK), sa = (int)(it % NA), db = (int)(g & 1);
          mbar_wait(&d1_empty[db], (uint32_t)(((g >> 1) & 1) ^ 1));
          if (j == 0) mbar_wait(&a_full[sa], (uint32_t)((it / NA) & 1));
          tc_fence_after();
          const uint64_t da = umma_smem_desc<L::ROWB>(smem_u32(sA + sa * L::A_BYTES));
          const uint64_t dw = umma_smem_desc<L::ROWB>(smem_u32(sW1 + j * 128 * L::ROWB));
#pragma unroll
          for (int kk = 0; kk < C / 16; ++kk)
            umma_bf16(tmem_base + (uint32_t)(db * 128), da + (uint64_t)(kk * 2), dw + (uint64_t)(kk * 2), idesc1,
                      kk != 0);
          umma_commit(&d1_full[db]);
          if (j == NCHUNK - 1) umma_commit(&a_empty[sa]);
        }
        if (g >= 1) {
          const long long gp = g - 1, it = gp / NCHUNK;
          const int j = (int)(gp % NCHUNK), hb = (int)(gp & 1), d2b = (int)(it & 1);
          mbar_wait(&h_full[hb], (uint32_t)((gp >> 1) & 1));
          if (j == 0) mbar_wait(&d2_empty[d2b], (uint32_t)(((it >> 1) & 1) ^ 1));
          tc_fence_after();
#pragma unroll
          for (int hk = 0; hk < 2; ++hk) {
            const uint64_t dh = umma_smem_desc<128>(smem_u32(sH + hb * L::H_BYTES + hk * (128 * 128)));
            const uint64_t dw = umma_smem_desc<128>(smem_u32(sW2 + (j * 2 + hk) * L::W2_KB_BYTES));
#pragma unroll
            for (int kk = 0; kk < 4; ++kk)
              umma_bf16(tmem_base + D2_COL0 + (uint32_t)(d2b * 64), dh + (uint64_t)(kk * 2), dw + (uint64_t)(kk * 2),
                        idesc2, (j | hk | kk) != 0);
          }
          umma_commit(&h_empty[hb]);
          if (j == NCHUNK - 1) umma_commit(&d2_full[d2b]);
        }
      }
    }
  } else if (warp < 2 + FF_E1_WARPS) {
    // ===== E1: D1 -> bias + GELU -> bf16 -> swizzled smem (A operand of GEMM 2) ============================
    const int q = warp & 3;                   // TMEM lane quarter
    const int colblk = (warp - 2) >> 2;       // which 32-column block of the 128-column chunk (0..3)
    const int half = colblk >> 1;             // its 64-column K-block in the H buffer
    const int c = colblk & 1;                 // position inside the K-block
    const int r = q * 32 + lane;              // row in the tile
    uint8_t* hrow_base = sH + half * (128 * 128) + (r >> 3) * 1024 + (r & 7) * 128;
    for (long long g = 0; g < total_g; ++g) {
      const int j = (int)(g % NCHUNK), db = (int)(g & 1);
      mbar_wait(&d1_full[db], (uint32_t)((g >> 1) & 1));
      mbar_wait(&h_empty[db], (uint32_t)(((g >> 1) & 1) ^ 1));
      tc_fence_after();
      uint8_t* hrow = hrow_base + db * L::H_BYTES;
      uint32_t v[32];
      tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(db * 128 + colblk * 32), v);
      tmem_ld_wait();
      tc_fence_before();
      const float* bp = sB1 + j * 128 + colblk * 32;
#pragma unroll
      for (int i = 0; i < 4; ++i) {             // 8 columns -> one 16-byte chunk
        const float4 b0 = *reinterpret_cast<const float4*>(bp + i * 8);
        const float4 b1v = *reinterpret_cast<const float4*>(bp + i * 8 + 4);
        uint4 pk;
        pk.x = gelu2_bias_pack(__uint_as_float(v[i * 8 + 0]), __uint_as_float(v[i * 8 + 1]), b0.x, b0.y);
        pk.y = gelu2_bias_pack(__uint_as_float(v[i * 8 + 2]), __uint_as_float(v[i * 8 + 3]), b0.z, b0.w);
        pk.z = gelu2_bias_pack(__uint_as_float(v[i * 8 + 4]), __uint_as_float(v[i * 8 + 5]), b1v.x, b1v.y);
        pk.w = gelu2_bias_pack(__uint_as_float(v[i * 8 + 6]), __uint_as_float(v[i * 8 + 7]), b1v.z, b1v.w);
        const int chunk = (c * 4 + i) ^ (r & 7);            // 128B swizzle: 16-byte chunk index XOR row-in-atom
        *reinterpret_cast<uint4*>(hrow + chunk * 16) = pk;
      }
      fence_proxy_async_smem();                 // generic-proxy smem writes -> visible to tcgen05.mma
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&d1_empty[db]);
        mbar_arrive(&h_full[db]);
      }
    }
  } else if (warp == 2 + FF_E1_WARPS + FF_E2_WARPS) {
    // ===== TMA store issuer: staged output tile -> global, then the slot goes back to the load ring =========
    if (elect_one()) {
      for (long long it = 0; it < my_tiles; ++it) {
        const int sr = (int)(it % NR);
        const long long tile = blockIdx.x + it * gridDim.x;
        mbar_wait(&o_full[sr], (uint32_t)((it / NR) & 1));
#pragma unroll
        for (int c = 0; c < C / 32; ++c)
          tma_store_2d(&tmO, sR + sr * L::R_BYTES + c * L::R_BOX_BYTES, c * 32, (int)(tile * 128));
        tma_store_commit();
        tma_store_wait_read<0>();               // smem has been read: the slot may be refilled
        mbar_arrive(&r_empty[sr]);
      }
      tma_store_wait_all<0>();
    }
  } else {
    // ===== E2: D2 -> + b2, * gamma, + resid (TMA-staged, swizzled smem) -> smem in place -> TMA store ========
    // Row-per-thread global stores cost one L1TEX tag cycle per row per instruction (32 per warp store) and
    // were the kernel's busiest unit; the TMA store reads the staged tile straight from shared memory.
    const int q = warp & 3;
    const int r = q * 32 + lane;
    for (long long it = 0; it < my_tiles; ++it) {
      const int d2b = (int)(it & 1), sr = (int)(it % NR);
      const long long tile = blockIdx.x + it * gridDim.x;
      const long long m = tile * 128 + r;
      const bool valid = m < p.R;
      long long brow = 0;
      if (p.out_bf16 != nullptr && valid) {
        const long long b = m / p.rows_pb;
        brow = b * p.ob_rpb + p.ob_off + (m - b * p.rows_pb);
      }
      mbar_wait(&r_full[sr], (uint32_t)((it / NR) & 1));
      mbar_wait(&d2_full[d2b], (uint32_t)((it >> 1) & 1));
      tc_fence_after();
      uint8_t* rrow = sR + sr * L::R_BYTES + r * 128;
#pragma unroll
      for (int c = 0; c < C / 16; ++c) {        // 16 columns at a time keeps E2 under the 88-register budget
        uint32_t v[16];
        tmem_ld_32x16(tmem_base + ((uint32_t)(q * 32) << 16) + D2_COL0 + (uint32_t)(d2b * 64 + c * 16), v);
        tmem_ld_wait();
        float x[16];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int col = c * 16 + i * 4;       // column in the tile; box = col / 32, 16-byte chunk = (col % 32) / 4
          float4* cell = reinterpret_cast<float4*>(rrow + (col >> 5) * L::R_BOX_BYTES +
                                                   ((((col & 31) >> 2) ^ (r & 7)) * 16));
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
          __nv_bfloat16* bp = p.out_bf16 + brow * C + c * 16;
#pragma unroll
          for (int i = 0; i < 2; ++i) {
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
        mbar_arrive(&o_full[sr]);               // this warp's 32 rows are staged
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, TMEM_COLS);
}

}  // namespace vv
