// pcm16.cuh — SURVEY §8f N3: the wire / on-disk format step right after decode().
//
// Replaces convert_to_16_bit_wav (reference main.py:2956-2979, the same function at demo/gradio_demo.py:1205), which the
// Gradio path runs on every streamed chunk (main.py:937-951) after a per-sample .item() + .cpu() copy in the streamer
// (vibevoice/modular/streamer.py:50-55):
//
//     peak = max|x| over the chunk;   if peak > 1:  x = x / peak;   pcm = (x * 32767).astype(int16)
//
// numpy does this in fp32 (the caller casts bf16 -> fp32 first, main.py:939-942): an IEEE division, an IEEE multiply and
// a C cast (truncation toward zero).  The kernels below use the same three operations (__fdiv_rn, __fmul_rn,
// __float2int_rz; dividing by 1.0f is exact, so the "if" becomes a select of the divisor), which makes the result
// BIT-EXACT, not merely close.  fmaxf drops NaNs where np.max would propagate them; the reference's result for a chunk
// with a NaN is undefined (NaN -> int16 cast), so that case is not part of the contract.
//
// HBM-bound byte work: 4 (or 2) B in, 2 B out per sample.  Rows up to PCM_ROW_MAX samples (a streamed chunk is 3200) take
// ONE launch with the row held in registers between the peak reduction and the conversion; longer rows (offline
// utterances) take a peak pass (atomicMax on the fp32 bit pattern, valid for non-negative floats) and a convert pass
// whose re-read of a batch of rows mostly hits the 126 MB L2.
#pragma once
#include <cuda_bf16.h>
#include <stdint.h>

namespace vv {

constexpr int PCM_THREADS = 256;
constexpr int PCM_VEC = 8;                                   // samples per thread per step: 32 B fp32 in, 16 B int16 out
constexpr int PCM_ROW_STEPS = 4;
constexpr int PCM_ROW_MAX = PCM_THREADS * PCM_VEC * PCM_ROW_STEPS;   // 8192 samples: the one-launch form
constexpr int PCM_CHUNK = PCM_THREADS * PCM_VEC * 8;         // samples per CTA in the two-pass form (16384)

struct Pcm16Args {
  const void* wave;          // [B, L] fp32 or bf16, rows `row_stride` elements apart
  int wave_is_bf16;
  int B;
  long long L, row_stride;
  int16_t* pcm;              // [B, L] contiguous
  float* peak;               // [B]  max|x| per row (output of the peak pass, input of the convert pass)
  int vec_ok;                // rows and base pointers are 16 B-aligned for 8-sample vectors
};

__device__ __forceinline__ void pcm_load8(const Pcm16Args& p, long long off, float (&v)[PCM_VEC]) {
  if (p.wave_is_bf16) {
    const uint4 q = __ldg(reinterpret_cast<const uint4*>(reinterpret_cast<const __nv_bfloat16*>(p.wave) + off));
    const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {                            // bf16 -> fp32 is a 16-bit shift (exact)
      v[2 * i] = __uint_as_float(w[i] << 16);
      v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
    }
  } else {
    const float4* s = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p.wave) + off);
    const float4 a = __ldg(s), b = __ldg(s + 1);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  }
}
__device__ __forceinline__ float pcm_load1(const Pcm16Args& p, long long off) {
  return p.wave_is_bf16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p.wave)[off])
                        : reinterpret_cast<const float*>(p.wave)[off];
}
// (x / div) * 32767 -> int16, the reference's three fp32 operations in order (no contraction, round-to-nearest, truncate)
__device__ __forceinline__ int pcm_quant(float x, float div) {
  return __float2int_rz(__fmul_rn(__fdiv_rn(x, div), 32767.0f));
}
__device__ __forceinline__ void pcm_store8(int16_t* dst, const float (&v)[PCM_VEC], float div) {
  uint32_t w[4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
    w[i] = ((uint32_t)pcm_quant(v[2 * i], div) & 0xffffu) | ((uint32_t)pcm_quant(v[2 * i + 1], div) << 16);
  *reinterpret_cast<uint4*>(dst) = make_uint4(w[0], w[1], w[2], w[3]);
}
// max over the CTA; every thread gets the result
__device__ __forceinline__ float pcm_block_max(float m, float* red) {
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
  __syncthreads();
  float r = red[0];
#pragma unroll
  for (int w = 1; w < PCM_THREADS / 32; ++w) r = fmaxf(r, red[w]);
  return r;
}

// ---- short rows: one CTA per row, one launch, one read of the row ------------------------------------------------------
__global__ void __launch_bounds__(PCM_THREADS) pcm16_row_kernel(const Pcm16Args p) {
  __shared__ float red[PCM_THREADS / 32];
  const int b = blockIdx.x;
  const long long src0 = (long long)b * p.row_stride;
  int16_t* dst = p.pcm + (long long)b * p.L;
  const int L = (int)p.L;
  float v[PCM_ROW_STEPS][PCM_VEC];
  float m = 0.f;
  if (p.vec_ok) {
#pragma unroll
    for (int s = 0; s < PCM_ROW_STEPS; ++s) {
      const int i0 = (s * PCM_THREADS + threadIdx.x) * PCM_VEC;
      if (i0 + PCM_VEC <= L) {
        pcm_load8(p, src0 + i0, v[s]);
#pragma unroll
        for (int i = 0; i < PCM_VEC; ++i) m = fmaxf(m, fabsf(v[s][i]));
      } else {
#pragma unroll
        for (int i = 0; i < PCM_VEC; ++i) {
          v[s][i] = (i0 + i < L) ? pcm_load1(p, src0 + i0 + i) : 0.f;
          m = fmaxf(m, fabsf(v[s][i]));
        }
      }
    }
  } else {
#pragma unroll
    for (int s = 0; s < PCM_ROW_STEPS; ++s) {
      const int i0 = (s * PCM_THREADS + threadIdx.x) * PCM_VEC;
#pragma unroll
      for (int i = 0; i < PCM_VEC; ++i) {
        v[s][i] = (i0 + i < L) ? pcm_load1(p, src0 + i0 + i) : 0.f;
        m = fmaxf(m, fabsf(v[s][i]));
      }
    }
  }
  const float peak = pcm_block_max(m, red);
  if (threadIdx.x == 0 && p.peak != nullptr) p.peak[b] = peak;
  const float div = (peak > 1.0f) ? peak : 1.0f;
#pragma unroll
  for (int s = 0; s < PCM_ROW_STEPS; ++s) {
    const int i0 = (s * PCM_THREADS + threadIdx.x) * PCM_VEC;
    if (p.vec_ok && i0 + PCM_VEC <= L) {
      pcm_store8(dst + i0, v[s], div);
    } else {
#pragma unroll
      for (int i = 0; i < PCM_VEC; ++i)
        if (i0 + i < L) dst[i0 + i] = (int16_t)pcm_quant(v[s][i], div);
    }
  }
}

// ---- long rows, pass 1: peak[b] = max|x| (peak zeroed by the host; grid = (chunks, B)) ---------------------------------
__global__ void __launch_bounds__(PCM_THREADS) pcm16_peak_kernel(const Pcm16Args p) {
  __shared__ float red[PCM_THREADS / 32];
  const int b = blockIdx.y;
  const long long src0 = (long long)b * p.row_stride;
  const long long c0 = (long long)blockIdx.x * PCM_CHUNK;
  const long long c1 = (c0 + PCM_CHUNK < p.L) ? c0 + PCM_CHUNK : p.L;
  float m = 0.f;
  if (p.vec_ok) {
    for (long long i0 = c0 + (long long)threadIdx.x * PCM_VEC; i0 < c1; i0 += PCM_THREADS * PCM_VEC) {
      if (i0 + PCM_VEC <= c1) {
        float v[PCM_VEC];
        pcm_load8(p, src0 + i0, v);
#pragma unroll
        for (int i = 0; i < PCM_VEC; ++i) m = fmaxf(m, fabsf(v[i]));
      } else {
        for (long long i = i0; i < c1; ++i) m = fmaxf(m, fabsf(pcm_load1(p, src0 + i)));
      }
    }
  } else {
    for (long long i = c0 + threadIdx.x; i < c1; i += PCM_THREADS) m = fmaxf(m, fabsf(pcm_load1(p, src0 + i)));
  }
  const float r = pcm_block_max(m, red);
  // non-negative fp32 values order like their bit patterns
  if (threadIdx.x == 0) atomicMax(reinterpret_cast<unsigned int*>(p.peak) + b, __float_as_uint(r));
}

// ---- long rows, pass 2: convert with the row's divisor -----------------------------------------------------------------
__global__ void __launch_bounds__(PCM_THREADS) pcm16_convert_kernel(const Pcm16Args p) {
  const int b = blockIdx.y;
  const long long src0 = (long long)b * p.row_stride;
  int16_t* dst = p.pcm + (long long)b * p.L;
  const long long c0 = (long long)blockIdx.x * PCM_CHUNK;
  const long long c1 = (c0 + PCM_CHUNK < p.L) ? c0 + PCM_CHUNK : p.L;
  const float peak = p.peak[b];
  const float div = (peak > 1.0f) ? peak : 1.0f;
  if (p.vec_ok) {
    for (long long i0 = c0 + (long long)threadIdx.x * PCM_VEC; i0 < c1; i0 += PCM_THREADS * PCM_VEC) {
      if (i0 + PCM_VEC <= c1) {
        float v[PCM_VEC];
        pcm_load8(p, src0 + i0, v);
        pcm_store8(dst + i0, v, div);
      } else {
        for (long long i = i0; i < c1; ++i) dst[i] = (int16_t)pcm_quant(pcm_load1(p, src0 + i), div);
      }
    }
  } else {
    for (long long i = c0 + threadIdx.x; i < c1; i += PCM_THREADS)
      dst[i] = (int16_t)pcm_quant(pcm_load1(p, src0 + i), div);
  }
}

}  // namespace vv
