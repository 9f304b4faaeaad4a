// rowops.cuh — the small HBM-bound kernels around the GEMMs: latent staging (stem input), the upsampler
// halo row, the head FIR, streaming-state bookkeeping, and the CUDA-core validation GEMM.
#pragma once
#include "mixer.cuh"
#include "gemm_tc.cuh"

namespace vv {

// ---- stem input ------------------------------------------------------------------------------------
// Stage the latents as bf16 [nbatch, T+6, D] with 6 leading context rows per item, so the stem SConv1d
// (64->2048, k7, reference :864-867) becomes a 7-tap implicit GEMM.  Context rows are zeros offline
// (left zero pad, :398-403) or the stream's cached last 6 latent frames (:327-382).  Accepts either
// reference layout ([B,T,D] or [B,D,T], :1102-1105) through strides, fp32 or bf16.
struct PrepArgs {
  const void* lat;
  int lat_is_bf16;
  long long sb, st, sd;      // element strides of (batch, time, dim)
  __nv_bfloat16* out;        // [nbatch, T+6, D]
  int nbatch, T, D;
  StateRef st_ref;           // state rows: bf16 [6][D]
};

__global__ void prep_latents_kernel(const PrepArgs p) {
  pdl_launch();
  pdl_wait();
  const long long per_b = (long long)(p.T + 6) * p.D;
  const long long total = per_b * p.nbatch;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int b = (int)(i / per_b);
    const int rem = (int)(i - (long long)b * per_b);
    const int r = rem / p.D;          // padded row: 0..5 context, 6.. = frames
    const int d = rem - r * p.D;
    const uint8_t* rd8;
    uint8_t* wr8;
    p.st_ref.resolve(b, rd8, wr8);
    __nv_bfloat16 v = __float2bfloat16(0.f);
    if (r >= 6) {
      const long long off = (long long)b * p.sb + (long long)(r - 6) * p.st + (long long)d * p.sd;
      v = p.lat_is_bf16 ? reinterpret_cast<const __nv_bfloat16*>(p.lat)[off]
                        : __float2bfloat16(reinterpret_cast<const float*>(p.lat)[off]);
    } else if (rd8 != nullptr) {
      v = reinterpret_cast<const __nv_bfloat16*>(rd8)[r * p.D + d];
    }
    p.out[i] = v;
    if (wr8 != nullptr && r >= p.T) reinterpret_cast<__nv_bfloat16*>(wr8)[(r - p.T) * p.D + d] = v;
  }
}

// ---- upsampler halo row ------------------------------------------------------------------------------
// xb is the bf16 copy of a stage's output, [nbatch, rows+1, C]; rows 1..rows were written by the last FFN
// epilogue of the stage.  Row 0 is x[-1] of the polyphase transposed conv: zero offline (the trimmed
// causal form never sees earlier input, reference :563-571) or the stream's previous last row (:539-547,
// of whose k-1 cached inputs only the newest one reaches the kept outputs).
struct HaloArgs {
  __nv_bfloat16* xb;
  int nbatch, rows, C;
  StateRef st_ref;           // state: bf16 [C]
};

__global__ void upsample_halo_kernel(const HaloArgs p) {
  pdl_launch();
  pdl_wait();
  const int total = p.nbatch * p.C;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int b = i / p.C, c = i - b * p.C;
    const uint8_t* rd8;
    uint8_t* wr8;
    p.st_ref.resolve(b, rd8, wr8);
    __nv_bfloat16* base = p.xb + (long long)b * (p.rows + 1) * p.C;
    base[c] = rd8 ? reinterpret_cast<const __nv_bfloat16*>(rd8)[c] : __float2bfloat16(0.f);
    if (wr8) reinterpret_cast<__nv_bfloat16*>(wr8)[c] = base[(long long)p.rows * p.C + c];
  }
}

// ---- head: SConv1d(32 -> 1, k7), reference :912, :950 ---------------------------------------------------
// audio[t] = bias + sum_j sum_c w[c][j] * x[t-6+j][c].  8 threads x float4 cover the 32 channels of a row;
// each 8-thread group walks a strip of rows with a 7-row register ring, 7 rows per step.
struct HeadArgs {
  const float* x;        // [nbatch, rows, 32]
  const float* w;        // [7][32] tap-major
  float bias;
  void* out;             // [nbatch, rows] fp32 or bf16
  int out_is_bf16;
  long long out_batch_stride;
  int nbatch, rows;
  int strip, strips_per_batch;
  StateRef st;           // state: fp32 [6][32]
};

__global__ void __launch_bounds__(128) head_kernel(const HeadArgs p) {
  pdl_launch();
  pdl_wait();
  constexpr int C = 32, LPR = 8;
  const int gtid = blockIdx.x * blockDim.x + threadIdx.x;
  const int group = gtid / LPR, lig = gtid % LPR;
  const int total_groups = p.nbatch * p.strips_per_batch;
  const bool live = group < total_groups;
  const int gsafe = live ? group : 0;
  const int b = gsafe / p.strips_per_batch;
  const int t_begin = (gsafe % p.strips_per_batch) * p.strip;
  const int t_end = live ? min(p.rows, t_begin + p.strip) : t_begin;
  const int c0 = lig * 4;
  const uint8_t* rd8;
  uint8_t* wr8;
  p.st.resolve(b, rd8, wr8);
  const float* st_rd = reinterpret_cast<const float*>(rd8);
  float* st_wr = reinterpret_cast<float*>(wr8);
  float4 w[7];
#pragma unroll
  for (int j = 0; j < 7; ++j) w[j] = *reinterpret_cast<const float4*>(p.w + j * C + c0);
  const float* xb = p.x + (long long)b * p.rows * C + c0;
  float4 win[7];
#pragma unroll
  for (int j = 0; j < 7; ++j) win[j] = make_float4(0.f, 0.f, 0.f, 0.f);
  const int ngroups7 = (p.strip + 12) / 7;
  for (int g7 = 0; g7 < ngroups7; ++g7) {
    const int tbase = t_begin - 6 + g7 * 7;
    float4 xr[7];
#pragma unroll
    for (int i = 0; i < 7; ++i) {
      const int t = tbase + i;
      xr[i] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (t < t_end) {
        if (t >= 0) xr[i] = *reinterpret_cast<const float4*>(xb + (long long)t * C);
        else if (st_rd != nullptr) xr[i] = *reinterpret_cast<const float4*>(st_rd + (6 + t) * C + c0);
      }
    }
    if (st_wr != nullptr) {
#pragma unroll
      for (int i = 0; i < 7; ++i) {
        const int t = tbase + i;
        const int dst = t - (p.rows - 6);
        if (t < t_end && dst >= 0 && (t >= t_begin || (t < 0 && t_begin == 0)))
          *reinterpret_cast<float4*>(st_wr + dst * C + c0) = xr[i];
      }
    }
    float s[7];
#pragma unroll
    for (int i = 0; i < 7; ++i) {
      win[i] = xr[i];
      float acc = 0.f;
#pragma unroll
      for (int j = 0; j < 7; ++j) {
        const float4 v = win[(i + 1 + j) % 7];
        acc = fmaf(w[j].x, v.x, acc);
        acc = fmaf(w[j].y, v.y, acc);
        acc = fmaf(w[j].z, v.z, acc);
        acc = fmaf(w[j].w, v.w, acc);
      }
      s[i] = acc;
    }
#pragma unroll
    for (int m = LPR / 2; m >= 1; m >>= 1) {
#pragma unroll
      for (int i = 0; i < 7; ++i) s[i] += __shfl_xor_sync(0xffffffffu, s[i], m);
    }
    // lane i of the group stores row i of this step: a 7-sample contiguous run per group
    float mine = 0.f;
#pragma unroll
    for (int i = 0; i < 7; ++i) mine = (lig == i) ? s[i] : mine;
    const int t = tbase + lig;
    if (lig < 7 && t >= t_begin && t < t_end) {
      const float v = mine + p.bias;
      const long long o = (long long)b * p.out_batch_stride + t;
      if (p.out_is_bf16) reinterpret_cast<__nv_bfloat16*>(p.out)[o] = __float2bfloat16(v);
      else reinterpret_cast<float*>(p.out)[o] = v;
    }
  }
}

// ---- streaming-state bookkeeping -----------------------------------------------------------------------
// After every stateful layer of a pass has written the inactive bank, the pass flips its slots' parity.
// A slot id outside [0, n_slots) was treated as "no state" by every layer (StateRef::resolve); here it is reported
// through the state's error word (host-mapped: the next vv_* call on the state returns VV_ERR_SLOT).
__global__ void state_flip_kernel(const int* slot_ids, int n, int* parity, int n_slots, int* err) {
  pdl_launch();
  pdl_wait();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int slot = slot_ids[i];
  if ((unsigned)slot < (unsigned)n_slots) parity[slot] ^= 1;
  else if (slot != -1 && err != nullptr) *err = (slot < 0) ? slot : slot + 1;   // nonzero either way; -1 = padding row
}
// set_to_zero (reference :234-241): clear the CURRENT bank of the listed slots (n < 0: every slot).
__global__ void state_zero_kernel(uint8_t* base, long long bank_bytes, long long slot_bytes, const int* slot_ids,
                                  int n, const int* parity, int n_slots, int* err) {
  const int slot = (slot_ids != nullptr) ? slot_ids[blockIdx.y] : blockIdx.y;
  if ((unsigned)slot >= (unsigned)n_slots) {
    if (slot != -1 && err != nullptr && blockIdx.x == 0 && threadIdx.x == 0) *err = (slot < 0) ? slot : slot + 1;
    return;
  }
  uint4* p = reinterpret_cast<uint4*>(base + (long long)(parity[slot] & 1) * bank_bytes +
                                      (long long)slot * slot_bytes);
  const long long n16 = slot_bytes / 16;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n16;
       i += (long long)gridDim.x * blockDim.x)
    p[i] = make_uint4(0, 0, 0, 0);
}

// ---- CUDA-core validation GEMM (gemm_path = 1) -------------------------------------------------------------
// Same operands, tap addressing and epilogue as gemm_tc_kernel, no tensor cores, no TMA.  It exists so the
// tcgen05 path can be checked against an independent implementation ON THE GPU and bisected; it is never
// selected automatically.  One thread = one row x 32 columns.
struct GemmRefArgs {
  const __nv_bfloat16* A;   // [nbatch, a_rpb, Ktap]
  long long a_rpb;          // rows per batch item in the A buffer (incl. halo)
  const __nv_bfloat16* W;   // [N, n_taps*Ktap]
  int Ktap;
  GemmArgs g;
};

__global__ void gemm_ref_kernel(const GemmRefArgs p) {
  const long long total_rows = (long long)p.g.nbatch * p.g.rows;
  const int chunks = p.g.N / 32;
  const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (idx >= total_rows * chunks) return;
  const int chunk = (int)(idx % chunks);
  const long long row = idx / chunks;
  const int b = (int)(row / p.g.rows), t = (int)(row - (long long)b * p.g.rows);
  const int n = chunk * 32;
  const int K = p.g.n_taps * p.Ktap;
  float acc[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) acc[j] = 0.f;
  for (int tap = 0; tap < p.g.n_taps; ++tap) {
    const __nv_bfloat16* a = p.A + ((long long)b * p.a_rpb + t + tap) * p.Ktap;
    for (int k = 0; k < p.Ktap; ++k) {
      const float av = __bfloat162float(a[k]);
#pragma unroll
      for (int j = 0; j < 32; ++j)
        acc[j] = fmaf(av, __bfloat162float(p.W[(long long)(n + j) * K + tap * p.Ktap + k]), acc[j]);
    }
  }
  uint32_t v[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = __float_as_uint(acc[j]);
  epilogue_store_32(p.g.epi, v, n, p.g.N, p.g.bias, p.g.gamma, p.g.resid, row, p.g.out_f32,
                    (long long)b * p.g.of_rpb + p.g.of_off + t, p.g.out_bf16,
                    (long long)b * p.g.ob_rpb + p.g.ob_off + t);
}

}  // namespace vv
