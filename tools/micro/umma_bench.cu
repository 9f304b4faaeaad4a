// umma_bench.cu — tcgen05.mma issue/throughput model on B200: cycles per MMA (M=128, K=16, bf16) as a function of N,
// A from shared memory (".ss") or tensor memory (".ts"), one CTA per SM, one issuing thread, operands resident.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I../../vibevoice-rocm_b200/csrc -o umma_bench umma_bench.cu
#include <cstdio>
#include "ptx.cuh"
using namespace vv;

template <int N, bool TS, int DEPS>   // DEPS: number of distinct accumulators cycled through (1 = fully dependent chain)
__global__ void __launch_bounds__(128, 1) k(long long* out, int iters) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  for (int i = threadIdx.x; i < (16384 + N * 128) / 4; i += 128) ((uint32_t*)smem)[i] = 0x3c003c00u + i;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
  if (threadIdx.x < 32) { tmem_alloc(&slot, 512); tmem_relinquish(); }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tb = slot;
  if (threadIdx.x == 0) {
    const uint64_t da = umma_smem_desc<128>(smem_u32(smem));
    const uint64_t db = umma_smem_desc<128>(smem_u32(smem + 16384));
    constexpr uint32_t idesc = umma_idesc_bf16(128, N);
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
        const uint32_t d = tb + (uint32_t)(((it * 4 + kk) % DEPS) * N) % 256;
        if (TS) umma_bf16_ts(d, tb + 256 + kk * 8, db + kk * 2, idesc, 1);
        else umma_bf16(d, da + kk * 2, db + kk * 2, idesc, 1);
      }
    }
    umma_commit(&bar);
    mbar_wait(&bar, 0);
    long long t1 = clock64();
    if (blockIdx.x == 0) out[0] = t1 - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) tmem_dealloc(tb, 512);
}

template <int N, bool TS, int DEPS> void run(long long* d) {
  const int iters = 4000;
  const int smem = 1024 + 16384 + N * 128;
  cudaFuncSetAttribute(k<N, TS, DEPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  k<N, TS, DEPS><<<148, 128, smem>>>(d, 10);
  k<N, TS, DEPS><<<148, 128, smem>>>(d, iters);
  long long h = 0;
  cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
  const double per = (double)h / (iters * 4);
  printf("N=%3d %s accs=%d : %7.1f clk/MMA  (ideal %5.1f)  -> %.0f%% of tensor peak   %s\n", N, TS ? "A=tmem" : "A=smem", DEPS, per,
         N / 2.0, 100.0 * (N / 2.0) / per, cudaGetErrorString(cudaGetLastError()));
}
int main() {
  long long* d; cudaMalloc(&d, 8);
  run<32, false, 1>(d); run<64, false, 1>(d); run<128, false, 1>(d); run<256, false, 1>(d);
  run<64, false, 2>(d); run<128, false, 2>(d);
  run<32, true, 1>(d); run<64, true, 1>(d); run<128, true, 1>(d); run<256, true, 1>(d);
  return 0;
}
