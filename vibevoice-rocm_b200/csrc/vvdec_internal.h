// vvdec_internal.h — host-side structures behind the C ABI in include/vvdec.h.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <map>
#include <mutex>
#include <set>
#include <string>
#include <vector>
#include "../../include/vvdec.h"

namespace vv {

void set_error(const char* fmt, ...);
const char* get_error();

#define VV_CUDA(expr)                                                                       \
  do {                                                                                      \
    cudaError_t _e = (expr);                                                                \
    if (_e != cudaSuccess) {                                                                \
      vv::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return VV_ERR_CUDA;                                                                   \
    }                                                                                       \
  } while (0)

struct DenseW {                 // one packed GEMM weight: bf16 [N, K] K-major + its TMA map
  __nv_bfloat16* w = nullptr;
  float* bias = nullptr;        // [N]
  int N = 0, K = 0, Ktap = 0, n_taps = 1;
  int BN = 0, BK = 0;
  CUtensorMap tm;
  bool has_tm256 = false;       // N % 256 == 0 and BK == 64: second map with a [64 x 256] box (gemm_tc2 BN = 256)
  CUtensorMap tm256;
  bool has_tm64 = false;        // BN == 128 and BK == 64: third map with a [64 x 64] box (gemm_tc2 BN = 64 for launches with few tiles)
  CUtensorMap tm64;
};

struct BlockW {
  float *norm_w = nullptr, *dw_w = nullptr, *dw_b = nullptr, *gamma = nullptr, *ffn_norm_w = nullptr,
        *ffn_gamma = nullptr;
  DenseW lin1, lin2;            // lin1 is stored pre-scaled by 0.5 (weights and bias), see epilogue.cuh
  CUtensorMap tm_w1s;           // lin1 through a [64 x 64] box (ffn_stream_kernel), C = 128 only
  CUtensorMap tm_w1fs;          // lin1f through a [64 x 64] box (block_ws_kernel), C = 128 only
  // block_tma_kernel (C <= 64) folds the per-channel scales into the GEMM operands: lin1f = 0.5 * W1 * diag(ffn_norm_w),
  // lin2f = diag(ffn_gamma) * W2, gb2 = ffn_gamma * b2  (lin1f.bias = 0.5 * b1 as in lin1)
  DenseW lin1f, lin2f;
  float* gb2 = nullptr;
  std::vector<float> h_b1, h_b2, h_fgamma;   // C == 128 / 256: 0.5 * b1, b2, ffn_gamma for ffn_stream_kernel's constant-bank arguments
  std::vector<float> h_b1c, h_gb2c;   // host copies: block_tma_kernel takes 0.5 * b1 and gb2 as by-value (constant-bank) arguments
  long long state_off = 0;      // fp32 [6][C]
};

struct StageW {
  int C = 0, rate = 1, ratio = 1;   // ratio: upsampling factor INTO this stage (stage 0: 1)
  DenseW up;                        // stage 0: stem; else polyphase transposed conv
  long long up_state_off = 0;       // stage 0: bf16 [6][D]; else bf16 [C_prev]
  std::vector<BlockW> blocks;
};

}  // namespace vv

struct vv_decoder {
  vv_config cfg;
  int hop = 1, halo = 0;
  int num_sms = 148;                // of cfg.device (persistent kernels launch one CTA per SM)
  bool finalized = false;
  std::map<std::string, std::vector<float>> staged;        // name -> fp32 host copy until finalize
  std::map<std::string, std::vector<int64_t>> staged_shape;
  std::vector<vv::StageW> stages;
  float* head_w = nullptr;          // [7][C_last] tap-major
  float head_bias = 0.f;
  long long head_state_off = 0;     // fp32 [6][C_last]
  long long state_slot_bytes = 0;
  std::vector<void*> allocs;
  int64_t param_count = 0;
  int64_t launches = 0;
  // run-time options (vv_decoder_set_option); defaults come from the environment once, at create
  int opt_bn256 = 1;                // gemm_tc2 tile width: 0 never 256, 1 auto (>= opt_bn256_waves waves of wide tiles), 2 whenever N % 256 == 0
  int opt_bn256_waves = 2;
  int opt_bn64 = 1;                 // N = 64 layers on the persistent kernel's 64-wide tile (0: one-tile-per-CTA kernel)
  int opt_bn64_small = 1;           // residual-epilogue GEMMs whose 128-wide tiles would occupy fewer than half the SMs use the 64-wide tile
  int opt_ffn_stream256 = 1;        // FFN half at C = 256: 1 = ffn_stream_kernel<256> (hidden on chip), 0 = two GEMMs through HBM
  int opt_block_ws = 1;             // whole-Block1D kernel at C = 128: 1 = block_ws_kernel, 0 = mixer_kernel + ffn_stream_kernel
  int opt_block_kernel = 1;         // whole-Block1D kernel at C <= 64: 1 = block_tma_kernel (x tile staged by TMA), 0 = block_fused_kernel
  // optional per-launch timing (vv_decoder_set_profiling): CUDA events recorded on the decode stream
  bool prof_on = false;
  struct ProfMark { cudaEvent_t ev; int kind; int stage; };
  std::vector<ProfMark> prof_marks;             // marks of the call in flight
  std::vector<cudaEvent_t> prof_pool;           // recycled events
  std::map<std::pair<int, int>, std::pair<double, int64_t>> prof_acc;   // (kind, stage) -> (ms, launches)
  void mark(int kind, int stage, cudaStream_t s);
};

struct vv_state {
  const vv_decoder* dec = nullptr;
  int n_slots = 0;
  long long slot_bytes = 0, bank_bytes = 0;
  uint8_t* base = nullptr;          // 2 banks
  int* parity = nullptr;            // [n_slots]
  int device = 0;                   // the decoder's / encoder's device: every vv_state_* entry point selects it
  // out-of-range slot ids: kernels never index the arena with them (StateRef::resolve) and state_flip / state_zero
  // write the offending id here.  Host-mapped pinned word, so the NEXT call on this state can report VV_ERR_SLOT
  // without synchronising the stream.
  int* err_host = nullptr;
  int* err_dev = nullptr;
};
