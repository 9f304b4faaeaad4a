// sampler.cuh — SURVEY.md 8f N4: the row / element kernels around the GEMMs of the diffusion-head sampler
// (reference: sample_speech_tokens, modeling_vibevoice_inference.py:695-708; VibeVoiceDiffusionHead,
// modular_vibevoice_diffusion_head.py:254-280; DPMSolverMultistepScheduler.step, vibevoice/schedule/dpm_solver.py:935-1022).
//
// One sampling step for n streams runs the head on 2n rows (n conditioned + n unconditioned, classifier-free guidance):
//   sc  = bf16( silu( cond_proj(condition) + t_emb[step] ) )                  samp_silu_add_kernel        (:155, :180, :275-276)
//   mod = sc  x [adaLN_0 | ... | adaLN_{L-1} | adaLN_final]^T                 one tcgen05 GEMM, N = (3L + 2) H
//   x   = bf16(latent, both halves) x noisy_images_proj^T                     samp_dup_latent_kernel + GEMM (K = 64)
//   per layer:  [x += gate * f of the previous layer;]  a = bf16( RMSNorm(x) w (1 + scale) + shift )   samp_norm_mod_kernel
//               h = bf16( silu(g) * u ),  [g | u] = a x [gate_proj ; up_proj]^T   one GEMM (N = 2F, weights interleaved in
//                   128-unit blocks) with the SwiGLU in its epilogue (EPI_SWIGLU_BF16, gemm_tc2.cuh)
//               f  = h x down_proj^T               (GEMM, K = F, as K slices: partial-sum planes added by the next norm)
//   final:      x += gate * f;  a = bf16( RMSNorm(x) (1 + scale_f) + shift_f );  v = a x linear^T (GEMM, N = 64)
//   solver:     eps = v_u + s (v_c - v_u);  m0 = alpha_s x - sigma_s eps;  x' = A x + B m0 + D (m0 - m1)   samp_solver_kernel
// Everything between GEMMs is fp32; GEMM operands are bf16 (like the tokenizer path).  All kernels are HBM-trivial
// (a few MB per step); the step is bound by the GEMMs' weight reads (245 MB of bf16 per step at H = 1536).
#pragma once
#include <cuda_bf16.h>
#include <stdint.h>
#include "ptx.cuh"

namespace vv {

__device__ __forceinline__ float samp_silu(float x) { return x / (1.0f + __expf(-x)); }

// sc[r][h] = bf16(silu(cproj[r][h] + temb[h])),  r < rows
__global__ void samp_silu_add_kernel(const float* __restrict__ cproj, const float* __restrict__ temb,
                                     __nv_bfloat16* __restrict__ sc, int rows, int H) {
  pdl_launch();
  pdl_wait();
  const long long total = (long long)rows * H / 4;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int h = (int)((i * 4) % H);
    const float4 c = reinterpret_cast<const float4*>(cproj)[i];
    const float4 t = *reinterpret_cast<const float4*>(temb + h);
    __nv_bfloat162 lo = __floats2bfloat162_rn(samp_silu(c.x + t.x), samp_silu(c.y + t.y));
    __nv_bfloat162 hi = __floats2bfloat162_rn(samp_silu(c.z + t.z), samp_silu(c.w + t.w));
    uint2 pk;
    pk.x = *reinterpret_cast<uint32_t*>(&lo);
    pk.y = *reinterpret_cast<uint32_t*>(&hi);
    reinterpret_cast<uint2*>(sc)[i] = pk;
  }
}

// xin[r][j] = bf16(x[r mod n][j]),  r < rows.  Sampling: rows = 2n, both guidance halves start every head pass from the same
// latent (:701-702).  Single head pass (test hook): rows = n, every row its own input.
__global__ void samp_dup_latent_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ xin, int n, int rows, int Ld) {
  pdl_launch();
  pdl_wait();
  const int total = rows * Ld;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int r = i / Ld, j = i - r * Ld;
    xin[i] = __float2bfloat16(x[(r % n) * Ld + j]);
  }
}

// One CTA per row.  Optional residual update with the previous layer's FFN output, then adaLN-modulated RMSNorm:
//   x[r] += gate[r] * f[r]                     (HeadLayer.forward :160; gate = mod[r][gate_off ..])
//   a[r]  = bf16( x[r] * rsqrt(mean(x^2) + eps) * w * (1 + scale[r]) + shift[r] )      (RMSNorm :20-41, modulate :43-45)
// w == nullptr: no affine weight (FinalLayer.norm_final, elementwise_affine=False :178).
struct SampNormArgs {
  float* x;                  // [rows, H] fp32, updated in place when f != nullptr
  const float* f;            // [f_planes][rows, H] previous layer's down_proj output as partial sums over K slices (plane 0
                             // carries the bias; added in plane order), or nullptr
  int f_planes;
  long long f_plane_stride;  // elements between planes
  const float* mod;          // [rows, NM] all adaLN outputs of this step
  int NM;
  int gate_off;              // column of the PREVIOUS layer's gate (used with f)
  int shift_off, scale_off;  // columns of THIS norm's shift / scale
  const float* w;            // [H] or nullptr
  __nv_bfloat16* a;          // [rows, H]
  int H;
  float eps;
};

constexpr int SAMP_NORM_THREADS = 256;
constexpr int SAMP_NORM_MAXV = 4;          // float4 per thread: H <= 256 * 4 * 4 = 4096

__global__ void __launch_bounds__(SAMP_NORM_THREADS) samp_norm_mod_kernel(const SampNormArgs p) {
  pdl_launch();
  pdl_wait();
  __shared__ float red[SAMP_NORM_THREADS / 32];
  const int r = blockIdx.x;
  const int nv = p.H / 4;
  float* xr = p.x + (long long)r * p.H;
  const float* mr = p.mod + (long long)r * p.NM;
  float4 v[SAMP_NORM_MAXV];
  float ss = 0.f;
#pragma unroll
  for (int k = 0; k < SAMP_NORM_MAXV; ++k) {
    const int i = threadIdx.x + k * SAMP_NORM_THREADS;
    v[k] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (i < nv) {
      v[k] = reinterpret_cast<const float4*>(xr)[i];
      if (p.f != nullptr) {
        float4 fv = reinterpret_cast<const float4*>(p.f + (long long)r * p.H)[i];
        for (int pl = 1; pl < p.f_planes; ++pl) {
          const float4 pv = reinterpret_cast<const float4*>(p.f + pl * p.f_plane_stride + (long long)r * p.H)[i];
          fv.x += pv.x; fv.y += pv.y; fv.z += pv.z; fv.w += pv.w;
        }
        const float4 gv = *reinterpret_cast<const float4*>(mr + p.gate_off + i * 4);
        v[k].x = fmaf(gv.x, fv.x, v[k].x);
        v[k].y = fmaf(gv.y, fv.y, v[k].y);
        v[k].z = fmaf(gv.z, fv.z, v[k].z);
        v[k].w = fmaf(gv.w, fv.w, v[k].w);
        reinterpret_cast<float4*>(xr)[i] = v[k];
      }
      ss += v[k].x * v[k].x + v[k].y * v[k].y + v[k].z * v[k].z + v[k].w * v[k].w;
    }
  }
#pragma unroll
  for (int m = 16; m >= 1; m >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, m);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
  __syncthreads();
  float tot = 0.f;
#pragma unroll
  for (int wdx = 0; wdx < SAMP_NORM_THREADS / 32; ++wdx) tot += red[wdx];
  const float rstd = rsqrtf(tot / (float)p.H + p.eps);
#pragma unroll
  for (int k = 0; k < SAMP_NORM_MAXV; ++k) {
    const int i = threadIdx.x + k * SAMP_NORM_THREADS;
    if (i < nv) {
      float4 wv = make_float4(1.f, 1.f, 1.f, 1.f);
      if (p.w != nullptr) wv = *reinterpret_cast<const float4*>(p.w + i * 4);
      const float4 sh = *reinterpret_cast<const float4*>(mr + p.shift_off + i * 4);
      const float4 sc = *reinterpret_cast<const float4*>(mr + p.scale_off + i * 4);
      __nv_bfloat162 lo = __floats2bfloat162_rn(fmaf(v[k].x * rstd * wv.x, 1.0f + sc.x, sh.x), fmaf(v[k].y * rstd * wv.y, 1.0f + sc.y, sh.y));
      __nv_bfloat162 hi = __floats2bfloat162_rn(fmaf(v[k].z * rstd * wv.z, 1.0f + sc.z, sh.z), fmaf(v[k].w * rstd * wv.w, 1.0f + sc.w, sh.w));
      uint2 pk;
      pk.x = *reinterpret_cast<uint32_t*>(&lo);
      pk.y = *reinterpret_cast<uint32_t*>(&hi);
      reinterpret_cast<uint2*>(p.a + (long long)r * p.H)[i] = pk;
    }
  }
}

// Classifier-free guidance + one DPM-Solver++ update per latent element (coefficients: host table, oracle/sampler_oracle.py
// step_coefficients).  v [2n, Ld]: rows < n conditioned, rows >= n unconditioned (:703-705).
struct SampSolverArgs {
  const float* v;
  float* x;            // [n, Ld] current latent, updated in place
  float* m_prev;       // [n, Ld] previous data prediction (written every step)
  int n, Ld;
  float cfg_scale, alpha_s, sigma_s, A, B, D;
};

__global__ void samp_solver_kernel(const SampSolverArgs p) {
  pdl_launch();
  pdl_wait();
  const int total = p.n * p.Ld;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const float vc = p.v[i], vu = p.v[total + i];
    const float eps = vu + p.cfg_scale * (vc - vu);
    const float x = p.x[i];
    const float m0 = p.alpha_s * x - p.sigma_s * eps;                       // convert_model_output, v_prediction (:578-581)
    float xn = p.A * x + p.B * m0;
    if (p.D != 0.f) xn += p.D * (m0 - p.m_prev[i]);                          // second-order midpoint term (:756-761)
    p.m_prev[i] = m0;
    p.x[i] = xn;
  }
}

}  // namespace vv
