// gemm_tc.cuh — the pointwise GEMM of the decode path on tcgen05 tensor cores (sm_100a).
//
//   D[m, n] = sum_tap sum_k A[row(m) + tap, k] * Wp[n, tap*Ktap + k]        (bf16 x bf16 -> fp32 in TMEM)
//
// One kernel serves every dense layer of the reference decoder
// (vibevoice/modular/modular_vibevoice_tokenizer.py):
//   * FFN linear1 / linear2 (:588-596)               taps = 1
//   * SConvTranspose1d in polyphase form (:875-878)  taps = 2  (x[q-1], x[q]),  N = stride*Cout
//   * stem SConv1d 64->2048 k7 (:864-867)            taps = 7  (implicit GEMM over the padded latents)
// A is a channels-last bf16 activation [batch, rows(+halo), C] read through a 3-D TMA tensor map; a
// "tap" is simply a +1 row shift of the TMA coordinate, so causal context comes from halo rows that
// hold zeros (offline, :398-403) or the streaming state (:327-382, :478-549).  Out-of-range rows and
// batches are zero-filled by TMA and masked at the store.
//
// CTA = 128 x BN output tile, 192 threads: warp 0 TMA producer, warp 1 TMEM owner + single-thread
// tcgen05.mma issuer, warps 2..5 epilogue (tcgen05.ld -> bias / erf-GELU / layer-scale + residual -> global).
// Several CTAs are resident per SM (TMEM: BN <= 128 of 512 columns), so one tile's epilogue overlaps
// another tile's TMA + MMA without a persistent scheduler.
#pragma once
#include <cuda.h>
#include "ptx.cuh"
#include "epilogue.cuh"

namespace vv {

enum GemmEpilogue : int {
  EPI_BIAS_F32 = 0,   // out_f32 = acc + bias                                  (stem, upsample)
  EPI_GELU_BF16 = 1,  // out_bf16 = bf16(gelu_erf(acc + bias))                  (FFN linear1, :594-595)
  EPI_RESID = 2,      // v = resid + gamma*(acc + bias); out_f32 = v; out_bf16 = bf16(v) if given (:941-942)
  EPI_SWIGLU_BF16 = 3 // gemm_tc2, 256-wide tile only: columns [0,128) of a tile = gate, [128,256) = up of the same 128 hidden
                      // units (weights packed interleaved); out_bf16[:, tile*128 + j] = bf16(silu(gate_j) * up_j), no bias
                      // (diffusion head FeedForwardNetwork, modular_vibevoice_diffusion_head.py:117-124)
};

constexpr int GEMM_BM = 128;
constexpr int GEMM_MAX_STAGES = 8;
constexpr int GEMM_THREADS = 192;

struct GemmArgs {
  int rows;          // valid rows per batch item (time steps at this layer's input rate)
  int nbatch;        // batch items
  int bt_log2;       // tile = (1 << bt_log2) rows x (128 >> bt_log2) batch items
  int n_taps;        // 1, 2 or 7
  int kb_per_tap;    // (channels per tap) / BK
  int N;             // output columns
  int stages;        // smem pipeline depth
  int epi;           // GemmEpilogue
  const float* bias;    // [N]
  const float* gamma;   // [N]           (EPI_RESID)
  const float* resid;   // [nbatch*rows, N]  (EPI_RESID)
  float* out_f32;       // row(b, t) = b*of_rpb + of_off + t, leading dimension N
  long long of_rpb;
  int of_off;
  __nv_bfloat16* out_bf16;
  long long ob_rpb;
  int ob_off;
  // When ob_rows_pb > 0 the launch covers a FLAT row range [ob_row0, ob_row0 + rows) of a [items, ob_rows_pb] tensor
  // (nbatch = 1): the padded bf16 copy then addresses item = m / ob_rows_pb, t = m % ob_rows_pb (gemm_tc2 only).
  int ob_rows_pb;
  long long ob_row0;
  // gemm_tc2 only, nbatch == 1, fp32-out bias epilogue: the K loop is cut into k_slices equal parts that run as separate
  // tiles (more CTAs streaming weights when there are few output tiles) and write PARTIAL sums into k_slices output planes
  // (the output map's third coordinate); plane 0 carries the bias.  The consumer adds the planes in a fixed order.
  int k_slices;         // 0 / 1 = off
};

template <int BN, int BK>
constexpr size_t gemm_smem_bytes(int stages) {
  return 1024 + (size_t)stages * (GEMM_BM * BK * 2 + BN * BK * 2) + 256;
}

template <int BN, int BK>
__global__ void __launch_bounds__(GEMM_THREADS)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const GemmArgs g) {
  constexpr int A_BYTES = GEMM_BM * BK * 2;
  constexpr int B_BYTES = BN * BK * 2;
  constexpr int ROW_BYTES = BK * 2;            // 128 (SWIZZLE_128B) or 64 (SWIZZLE_64B)
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_align_1k(smem_raw);
  uint8_t* sA = smem;
  uint8_t* sB = smem + (size_t)g.stages * A_BYTES;
  uint64_t* bars = (uint64_t*)(sB + (size_t)g.stages * B_BYTES);
  uint64_t* full = bars;
  uint64_t* empty = bars + GEMM_MAX_STAGES;
  uint64_t* acc_full = bars + 2 * GEMM_MAX_STAGES;
  uint32_t* tmem_slot = (uint32_t*)(bars + 2 * GEMM_MAX_STAGES + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int bt = 1 << g.bt_log2;
  const int bb = GEMM_BM >> g.bt_log2;
  const int tiles_t = (g.rows + bt - 1) >> g.bt_log2;
  const int tiles_n = g.N / BN;                 // 1-D grid, N fastest: CTAs sharing an A tile run together
  const int tile_n = blockIdx.x % tiles_n;
  const int tile_m = blockIdx.x / tiles_n;
  const int tile_t = tile_m % tiles_t;
  const int tile_b = tile_m / tiles_t;
  const int nk = g.n_taps * g.kb_per_tap;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < g.stages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    mbar_init(acc_full, 1);
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, BN < 32 ? 32 : BN);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch();
  pdl_wait();

  if (warp == 0) {
    // ===== TMA producer ======================================================================
    if (elect_one()) {
      for (int kb = 0; kb < nk; ++kb) {
        const int s = kb % g.stages;
        const uint32_t ph = (kb / g.stages) & 1;
        mbar_wait(&empty[s], ph ^ 1);
        const int tap = kb / g.kb_per_tap;
        const int kin = (kb - tap * g.kb_per_tap) * BK;
        mbar_expect_tx(&full[s], A_BYTES + B_BYTES);
        tma_load_3d(sA + (size_t)s * A_BYTES, &tmA, &full[s], kin, tile_t * bt + tap, tile_b * bb);
        tma_load_2d(sB + (size_t)s * B_BYTES, &tmB, &full[s], kb * BK, tile_n * BN);
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (one elected thread) ======================================================
    constexpr uint32_t idesc = umma_idesc_bf16(GEMM_BM, BN);
    if (elect_one()) {
      for (int kb = 0; kb < nk; ++kb) {
        const int s = kb % g.stages;
        const uint32_t ph = (kb / g.stages) & 1;
        mbar_wait(&full[s], ph);
        tc_fence_after();
        const uint64_t da = umma_smem_desc<ROW_BYTES>(smem_u32(sA + (size_t)s * A_BYTES));
        const uint64_t db = umma_smem_desc<ROW_BYTES>(smem_u32(sB + (size_t)s * B_BYTES));
#pragma unroll
        for (int kk = 0; kk < BK / 16; ++kk) {
          // advance 16 elements (32 bytes) along K inside the swizzle span: +2 in the >>4 address field
          umma_bf16(tmem_base, da + (uint64_t)(kk * 2), db + (uint64_t)(kk * 2), idesc, (kb | kk) != 0);
        }
        umma_commit(&empty[s]);            // frees the smem stage once these MMAs have read it
      }
      umma_commit(acc_full);               // accumulator complete
    }
  } else {
    // ===== epilogue warps 2..5 ================================================================
    const int q = warp & 3;                // TMEM lane quarter this warp may read
    const int r = q * 32 + lane;           // row inside the tile
    const int b = tile_b * bb + (r >> g.bt_log2);
    const int t = tile_t * bt + (r & (bt - 1));
    const bool valid = (b < g.nbatch) && (t < g.rows);
    mbar_wait(acc_full, 0);
    tc_fence_after();
    const int n0 = tile_n * BN;
#pragma unroll 1
    for (int c = 0; c < BN / 32; ++c) {
      uint32_t v[32];
      tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(c * 32), v);
      tmem_ld_wait();
      const int n = n0 + c * 32;
      if (valid) epilogue_store_32(g.epi, v, n, g.N, g.bias, g.gamma, g.resid, (long long)b * g.rows + t,
                                   g.out_f32, (long long)b * g.of_rpb + g.of_off + t, g.out_bf16,
                                   (long long)b * g.ob_rpb + g.ob_off + t);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, BN < 32 ? 32 : BN);
}

}  // namespace vv
