// mixer.cuh — kernel class (a): the Block1D mixer half, fully fused, HBM-bound.
//
//   y = x + gamma * ( dwconv7( RMSNorm(x) * norm_w ) + conv_bias )          (reference :927-932 / :667-672)
//   a = bf16( RMSNorm(y) * ffn_norm_w )                                     (reference :936 / :676, feeds linear1)
//
// RMSNorm follows ConvRMSNorm/RMSNorm._norm (:65-66, :81-91): fp32 statistics over channels,
// x * rsqrt(mean(x^2) + eps), then * weight.  The depthwise conv is the causal k=7 SConv1d of the mixer
// (:640-647): output t sees normalised rows t-6..t; rows before the start are zeros offline (:398-403) or
// the stream's cached context (:327-382).  We cache the last 6 RAW rows and re-normalise them, which is
// output-identical to caching the normalised rows and makes streaming bit-equal to offline.
//
// Layout: channels-last fp32 [batch, rows, C].  A "group" of LPR threads owns one strip of consecutive
// rows of one batch item; each thread owns 4 adjacent channels (one float4, fully coalesced) and keeps a
// 7-row ring of normalised values in registers (the sliding window).  Rows are processed 7 at a time:
// 7 independent loads in flight per thread, 7 reductions batched per exchange.
//   LPR <= 32 : several groups per warp, reductions are width-LPR shuffles, no shared memory   (C <= 128)
//   LPR  > 32 : one group per CTA (blockDim = C/4), reductions go through a small smem exchange  (C >= 256)
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include "epilogue.cuh"
#include "ptx.cuh"

namespace vv {

struct StateRef {          // how a kernel finds a stream's cached context for one layer (null base = offline)
  const int* slot_ids;     // [nbatch] device
  const int* parity;       // [n_slots] device: which bank holds the current state of a slot
  uint8_t* base;           // state arena
  long long bank_bytes;    // distance between the two banks
  long long slot_bytes;    // distance between slots inside a bank
  long long layer_off;     // this layer's offset inside a slot
  int n_slots;             // slot ids outside [0, n_slots) never touch the arena: the row decodes with a zero context,
                           // saves nothing, and the pass's state_flip_kernel reports it (vv_state::err -> VV_ERR_SLOT);
                           // id -1 is the documented "padding row" (same treatment, no report)
  __device__ __forceinline__ void resolve(int b, const uint8_t*& rd, uint8_t*& wr) const {
    rd = nullptr; wr = nullptr;
    if (base == nullptr) return;
    const int slot = slot_ids[b];
    if ((unsigned)slot >= (unsigned)n_slots) return;
    const int par = parity[slot] & 1;
    rd = base + (long long)par * bank_bytes + (long long)slot * slot_bytes + layer_off;
    wr = base + (long long)(par ^ 1) * bank_bytes + (long long)slot * slot_bytes + layer_off;
  }
};

struct MixerArgs {
  const float* x;        // [nbatch, rows, C]
  float* y;              // [nbatch, rows, C]
  __nv_bfloat16* a;      // [nbatch, rows, C]
  const float* norm_w;   // [C]
  const float* dw_w;     // [7][C]  tap-major (tap 6 multiplies the current row)
  const float* dw_b;     // [C]
  const float* gamma;    // [C]
  const float* ffn_norm_w;  // [C]
  int nbatch, rows, C;
  int strip;             // rows per group
  int strips_per_batch;
  float eps;
  StateRef st;
};

template <int LPR>
__device__ __forceinline__ void reduce7(float (&s)[7], float* xchg, int buf) {
  if constexpr (LPR <= 32) {
#pragma unroll
    for (int m = LPR / 2; m >= 1; m >>= 1) {
#pragma unroll
      for (int i = 0; i < 7; ++i) s[i] += __shfl_xor_sync(0xffffffffu, s[i], m);
    }
  } else {
    constexpr int NW = LPR / 32;
#pragma unroll
    for (int m = 16; m >= 1; m >>= 1) {
#pragma unroll
      for (int i = 0; i < 7; ++i) s[i] += __shfl_xor_sync(0xffffffffu, s[i], m);
    }
    float* slot = xchg + buf * (NW * 8);
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
      for (int i = 0; i < 7; ++i) slot[(threadIdx.x >> 5) * 8 + i] = s[i];
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 7; ++i) {
      float t = 0.f;
#pragma unroll
      for (int w = 0; w < NW; ++w) t += slot[w * 8 + i];
      s[i] = t;
    }
  }
}

// Measured on B200 (tools/kbench.py --kind mixer, 5760 frames): keeping the next 7-row batch in flight while the
// current one is computed, at 3 resident CTAs per SM (the extra 28 registers do not fit 4), lifts stage 4 from
// 4.84 to 5.92 TB/s moved and stage 3 from 3.89 to 5.30 TB/s.
#ifndef VV_MIXER_MINBLOCKS
#define VV_MIXER_MINBLOCKS 3
#endif
#ifndef VV_MIXER_PREFETCH
#define VV_MIXER_PREFETCH 1
#endif
#ifndef VV_MIXER_PREFETCH_MAXLPR
#define VV_MIXER_PREFETCH_MAXLPR 128   // groups wider than this do not keep a second 7-row batch in registers (measured on
                                       // B200, 5760 frames: C = 2048 75 -> 57 us, C = 1024 with two resident CTAs 157 -> 138 us)
#endif
#ifndef VV_MIXER_MINBLOCKS256
#define VV_MIXER_MINBLOCKS256 2        // resident CTAs per SM the compiler must allow for the 256-thread group (C = 1024)
#endif
#ifndef VV_MIXER_MINBLOCKS128
#define VV_MIXER_MINBLOCKS128 3        // ... for the 128-thread group (C = 512): 326 -> 273 us per 5760 frames
#endif
template <int LPR>
__global__ void __launch_bounds__((LPR <= 32) ? 128 : LPR, (LPR <= 32) ? VV_MIXER_MINBLOCKS : (LPR == 256 ? VV_MIXER_MINBLOCKS256 : (LPR == 128 ? VV_MIXER_MINBLOCKS128 : 1)))
mixer_kernel(const MixerArgs p) {
  pdl_launch();
  pdl_wait();
  __shared__ float xchg[(LPR <= 32) ? 1 : 4 * (LPR / 32) * 8];
  int group, lane_in_group;
  if constexpr (LPR <= 32) {
    const int gtid = blockIdx.x * blockDim.x + threadIdx.x;
    group = gtid / LPR;
    lane_in_group = gtid % LPR;
  } else {
    group = blockIdx.x;
    lane_in_group = threadIdx.x;
  }
  const int total_groups = p.nbatch * p.strips_per_batch;
  const bool live = group < total_groups;          // dead groups still take part in shuffles
  const int gsafe = live ? group : 0;
  const int b = gsafe / p.strips_per_batch;
  const int t_begin = (gsafe % p.strips_per_batch) * p.strip;
  const int t_end = live ? min(p.rows, t_begin + p.strip) : t_begin;
  const int C = p.C;
  const int c0 = lane_in_group * 4;
  const float inv_c = 1.0f / (float)C;

  const uint8_t* st_rd8;
  uint8_t* st_wr8;
  p.st.resolve(b, st_rd8, st_wr8);
  const float* st_rd = reinterpret_cast<const float*>(st_rd8);   // [6][C] raw rows t = -6..-1
  float* st_wr = reinterpret_cast<float*>(st_wr8);

  const float4 nw = *reinterpret_cast<const float4*>(p.norm_w + c0);
  const float4 cb = *reinterpret_cast<const float4*>(p.dw_b + c0);
  const float4 gm = *reinterpret_cast<const float4*>(p.gamma + c0);
  const float4 fw = *reinterpret_cast<const float4*>(p.ffn_norm_w + c0);
  float4 w[7];
#pragma unroll
  for (int j = 0; j < 7; ++j) w[j] = *reinterpret_cast<const float4*>(p.dw_w + j * C + c0);

  const float* xb = p.x + (long long)b * p.rows * C + c0;
  float* yb = p.y + (long long)b * p.rows * C + c0;
  __nv_bfloat16* ab = p.a + (long long)b * p.rows * C + c0;

  float4 win[7];
#pragma unroll
  for (int j = 0; j < 7; ++j) win[j] = make_float4(0.f, 0.f, 0.f, 0.f);

  const int ngroups7 = (p.strip + 6 + 6) / 7;      // uniform trip count (strip + 6 warm-up rows)
  // 7 raw rows starting at row tb (rows before 0 come from the stream state or are zero)
  auto load7 = [&](int tb, float4 (&dst)[7]) {
#pragma unroll
    for (int i = 0; i < 7; ++i) {
      const int t = tb + i;
      dst[i] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (t < t_end) {
        if (t >= 0) dst[i] = *reinterpret_cast<const float4*>(xb + (long long)t * C);
        else if (st_rd != nullptr) dst[i] = *reinterpret_cast<const float4*>(st_rd + (6 + t) * C + c0);
      }
    }
  };
  constexpr bool PREFETCH = VV_MIXER_PREFETCH && (LPR <= VV_MIXER_PREFETCH_MAXLPR);
  float4 xn[7];                                     // the NEXT batch is always in flight while this one is computed
  if constexpr (PREFETCH) load7(t_begin - 6, xn);
  for (int g7 = 0; g7 < ngroups7; ++g7) {
    const int tbase = t_begin - 6 + g7 * 7;
    float4 xr[7];
    if constexpr (PREFETCH) {
#pragma unroll
      for (int i = 0; i < 7; ++i) xr[i] = xn[i];
      if (g7 + 1 < ngroups7) load7(tbase + 7, xn);
    } else {
      load7(tbase, xr);
    }
    // streaming: the new context is the last 6 rows of [old context, x]
    if (st_wr != nullptr) {
#pragma unroll
      for (int i = 0; i < 7; ++i) {
        const int t = tbase + i;
        const int dst = t - (p.rows - 6);
        if (t < t_end && dst >= 0 && (t >= t_begin || (t < 0 && t_begin == 0)))
          *reinterpret_cast<float4*>(st_wr + dst * C + c0) = xr[i];
      }
    }
    float ss[7];
#pragma unroll
    for (int i = 0; i < 7; ++i)
      ss[i] = xr[i].x * xr[i].x + xr[i].y * xr[i].y + xr[i].z * xr[i].z + xr[i].w * xr[i].w;
    reduce7<LPR>(ss, xchg, (g7 & 1) * 2);

    float4 yv[7];
#pragma unroll
    for (int i = 0; i < 7; ++i) {
      const float rstd = rsqrtf(ss[i] * inv_c + p.eps);
      win[i] = make_float4(xr[i].x * rstd * nw.x, xr[i].y * rstd * nw.y, xr[i].z * rstd * nw.z,
                           xr[i].w * rstd * nw.w);
      float4 acc = cb;
#pragma unroll
      for (int j = 0; j < 7; ++j) {
        const float4 nv = win[(i + 1 + j) % 7];    // normalised row t-6+j
        acc.x = fmaf(w[j].x, nv.x, acc.x);
        acc.y = fmaf(w[j].y, nv.y, acc.y);
        acc.z = fmaf(w[j].z, nv.z, acc.z);
        acc.w = fmaf(w[j].w, nv.w, acc.w);
      }
      yv[i] = make_float4(fmaf(gm.x, acc.x, xr[i].x), fmaf(gm.y, acc.y, xr[i].y), fmaf(gm.z, acc.z, xr[i].z),
                          fmaf(gm.w, acc.w, xr[i].w));
      ss[i] = yv[i].x * yv[i].x + yv[i].y * yv[i].y + yv[i].z * yv[i].z + yv[i].w * yv[i].w;
    }
    reduce7<LPR>(ss, xchg, (g7 & 1) * 2 + 1);
#pragma unroll
    for (int i = 0; i < 7; ++i) {
      const int t = tbase + i;
      if (t >= t_begin && t < t_end) {
        const float rstd = rsqrtf(ss[i] * inv_c + p.eps);
        *reinterpret_cast<float4*>(yb + (long long)t * C) = yv[i];
        uint2 pk;
        pk.x = pack_bf16x2(yv[i].x * rstd * fw.x, yv[i].y * rstd * fw.y);
        pk.y = pack_bf16x2(yv[i].z * rstd * fw.z, yv[i].w * rstd * fw.w);
        *reinterpret_cast<uint2*>(ab + (long long)t * C) = pk;
      }
    }
  }
}

}  // namespace vv
