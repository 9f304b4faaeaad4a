// gelu_bench.cu — throughput of GELU-pair variants on one SM-filling grid (registers only, no memory in the loop).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gelu_bench gelu_bench.cu ; run on the GPU box.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk(float a, float b) { f32x2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void upk(f32x2 v, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) { f32x2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) { f32x2 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) { f32x2 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ float tanhf_a(float x) { float y; asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ uint32_t bf2(float a, float b) { __nv_bfloat162 h = __floats2bfloat162_rn(a, b); return *reinterpret_cast<uint32_t*>(&h); }

// V0: current (bias add, deg-4 Q, 0.5x)
__device__ __forceinline__ uint32_t g0(float a, float b, float ba, float bb) {
  const f32x2 x = add2(pk(a, b), pk(ba, bb));
  const f32x2 s = mul2(x, x);
  f32x2 q = fma2(pk(1.4250002314e-06f, 1.4250002314e-06f), s, pk(-3.6693157976e-05f, -3.6693157976e-05f));
  q = fma2(q, s, pk(-8.7828905299e-05f, -8.7828905299e-05f));
  q = fma2(q, s, pk(3.6392291184e-02f, 3.6392291184e-02f));
  q = fma2(q, s, pk(7.9786971043e-01f, 7.9786971043e-01f));
  float u0, u1; upk(mul2(x, q), u0, u1);
  const f32x2 t = pk(tanhf_a(u0), tanhf_a(u1));
  const f32x2 hx = mul2(x, pk(0.5f, 0.5f));
  float h0, h1; upk(fma2(hx, t, hx), h0, h1);
  return bf2(h0, h1);
}
// V1: prescaled z = x/2 (bias add kept), deg-1 Q'
__device__ __forceinline__ uint32_t g1(float a, float b, float ba, float bb) {
  const f32x2 z = add2(pk(a, b), pk(ba, bb));
  const f32x2 s = mul2(z, z);
  const f32x2 q = fma2(pk(0.2776f, 0.2776f), s, pk(1.6003f, 1.6003f));
  float u0, u1; upk(mul2(z, q), u0, u1);
  const f32x2 t = pk(tanhf_a(u0), tanhf_a(u1));
  float h0, h1; upk(fma2(z, t, z), h0, h1);
  return bf2(h0, h1);
}
// V2: V1 without bias add
__device__ __forceinline__ uint32_t g2(float a, float b, float ba, float bb) {
  const f32x2 z = pk(a, b);
  const f32x2 s = mul2(z, z);
  const f32x2 q = fma2(pk(0.2776f, 0.2776f), s, pk(1.6003f, 1.6003f));
  float u0, u1; upk(mul2(z, q), u0, u1);
  const f32x2 t = pk(tanhf_a(u0), tanhf_a(u1));
  float h0, h1; upk(fma2(z, t, z), h0, h1);
  return bf2(h0, h1);
}
// V3: V1 with tanh.approx.bf16x2 (one cvt to bf16x2, 2 MUFU.BF16, result combined in bf16x2 HFMA2)
__device__ __forceinline__ uint32_t g3(float a, float b, float ba, float bb) {
  const f32x2 z = add2(pk(a, b), pk(ba, bb));
  const f32x2 s = mul2(z, z);
  const f32x2 q = fma2(pk(0.2776f, 0.2776f), s, pk(1.6003f, 1.6003f));
  float u0, u1; upk(mul2(z, q), u0, u1);
  uint32_t uh; asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(uh) : "f"(u1), "f"(u0));
  uint32_t th; asm("tanh.approx.f16x2 %0, %1;" : "=r"(th) : "r"(uh));
  float t0, t1;
  asm("{.reg .f16 l, h; mov.b32 {l, h}, %2; cvt.f32.f16 %0, l; cvt.f32.f16 %1, h;}" : "=f"(t0), "=f"(t1) : "r"(th));
  float h0, h1; upk(fma2(z, pk(t0, t1), z), h0, h1);
  return bf2(h0, h1);
}
// V4: MUFU-free: Phi(x) ~ 0.5 + x*P(x^2) clamped; here a degree-5 P as a cost model (coefficients are placeholders)
__device__ __forceinline__ uint32_t g4(float a, float b, float ba, float bb) {
  const f32x2 z = add2(pk(a, b), pk(ba, bb));
  float z0, z1; upk(z, z0, z1);
  const float c0 = fminf(fmaxf(z0, -2.f), 2.f), c1 = fminf(fmaxf(z1, -2.f), 2.f);
  const f32x2 c = pk(c0, c1);
  const f32x2 s = mul2(c, c);
  f32x2 p = fma2(pk(1.1e-3f, 1.1e-3f), s, pk(-1.2e-2f, -1.2e-2f));
  p = fma2(p, s, pk(6.1e-2f, 6.1e-2f));
  p = fma2(p, s, pk(-1.9e-1f, -1.9e-1f));
  p = fma2(p, s, pk(3.9e-1f, 3.9e-1f));
  p = fma2(p, s, pk(7.9e-1f, 7.9e-1f));
  const f32x2 phi = fma2(c, p, pk(1.0f, 1.0f));     // 1 + erf-like, in [0, 2]
  float h0, h1; upk(mul2(z, phi), h0, h1);
  return bf2(h0, h1);
}

template <int V>
__global__ void __launch_bounds__(512, 1) bench(const float* in, uint32_t* out, int iters) {
  float v[16], bias[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) { v[i] = in[(threadIdx.x * 16 + i) & 1023]; bias[i] = in[(i * 7 + threadIdx.x) & 1023]; }
  uint32_t acc = 0;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      uint32_t r;
      if (V == 0) r = g0(v[2 * i], v[2 * i + 1], bias[2 * i], bias[2 * i + 1]);
      else if (V == 1) r = g1(v[2 * i], v[2 * i + 1], bias[2 * i], bias[2 * i + 1]);
      else if (V == 2) r = g2(v[2 * i], v[2 * i + 1], bias[2 * i], bias[2 * i + 1]);
      else if (V == 3) r = g3(v[2 * i], v[2 * i + 1], bias[2 * i], bias[2 * i + 1]);
      else if (V == 4) r = g4(v[2 * i], v[2 * i + 1], bias[2 * i], bias[2 * i + 1]);
      else r = (i & 3) == 3 ? g4(v[2 * i], v[2 * i + 1], bias[2 * i], bias[2 * i + 1])
                            : g1(v[2 * i], v[2 * i + 1], bias[2 * i], bias[2 * i + 1]);
      acc ^= r;
      v[2 * i] += __uint_as_float((r & 0x7f) | 0x33000000);        // tiny perturbation keeps the loop live
      v[2 * i + 1] -= __uint_as_float(((r >> 16) & 0x7f) | 0x33000000);
    }
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int V> void run(const char* name, const float* in, uint32_t* out) {
  const int iters = 4000, grid = 148, block = 512;
  bench<V><<<grid, block>>>(in, out, 10);
  cudaDeviceSynchronize();
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  bench<V><<<grid, block>>>(in, out, iters);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  const double pairs = (double)grid * block * iters * 8;
  printf("%-28s %8.3f ms  %7.2f Gpair/s  (%.2f SM-cycles per warp-pair-op at 1.7 GHz, per SMSP)\n", name, ms,
         pairs / ms * 1e-6, ms * 1e-3 * 1.7e9 / (pairs / 32 / (148 * 4)));
}
int main() {
  float* in; uint32_t* out;
  cudaMalloc(&in, 4096); cudaMalloc(&out, 148 * 512 * 4);
  float h[1024]; for (int i = 0; i < 1024; ++i) h[i] = (float)((i * 37) % 200 - 100) * 0.02f;
  cudaMemcpy(in, h, 4096, cudaMemcpyHostToDevice);
  run<0>("V0 current deg4", in, out);
  run<1>("V1 prescale deg1", in, out);
  run<2>("V2 V1 no bias", in, out);
  run<3>("V3 V1 tanh.f16x2", in, out);
  run<4>("V4 poly MUFU-free (deg5)", in, out);
  run<5>("V5 mix 3:1 V1:V4", in, out);
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
