// vvdec.cu — C ABI of the B200-native VibeVoice acoustic-tokenizer decoder (include/vvdec.h):
// weight packing, workspace planning, the launch schedule of one decode pass, streaming state.
// Replaces TokenizerDecoder.forward / forward_features (modular_vibevoice_tokenizer.py:914-951) behind
// VibeVoiceAcousticTokenizerModel.decode (:1099-1108).  No CPU fallback: every path below launches CUDA.
#include <stdarg.h>
#include <stdio.h>
#include <string.h>
#include <stdlib.h>
#include <algorithm>
#include <utility>
#include <cmath>
#include "vvdec_internal.h"
#include "rowops.cuh"
#include "ffn_fused.cuh"
#include "ffn_stream.cuh"
#include "block_fused.cuh"
#include "block_tma.cuh"
#include "block_ws.cuh"
#include "gemm_tc2.cuh"
#include "encoder.cuh"
#include "pcm16.cuh"
#include "sampler.cuh"

namespace vv {

static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
const char* get_error() { return g_err; }

// ---- kernel launch -------------------------------------------------------------------------------------
// Every kernel of a decode / encode pass goes through launch_k.  With VV_PDL=1 the launch carries the programmatic
// stream-serialization attribute: the kernel's CTAs may be scheduled while the previous kernel of the stream drains,
// run their prologue (mbarrier init, TMEM allocation, descriptor prefetch) and block in griddepcontrol.wait until the
// predecessor has completed (ptx.cuh: pdl_launch / pdl_wait; every kernel launched here executes both).
// VV_PDL unset: on for small streaming calls only (a state is given and the call holds <= 2048 frames: the K4 step
// drops 2.95 -> 2.83 ms; every launch of such a call is 5-140 us long), off for offline passes (K2 measured 5 % SLOWER
// with it).  VV_PDL=0: never; VV_PDL=1: every launch (the whole GPU suite passes under it).
static thread_local bool g_pdl_small_stream = false;
struct PdlScope {                        // marks the calling thread's current vv_decode / vv_encode call
  explicit PdlScope(bool small_stream) { g_pdl_small_stream = small_stream; }
  ~PdlScope() { g_pdl_small_stream = false; }
};
static bool pdl_enabled() {
  static int mode = -2;
  if (mode == -2) { const char* ev = getenv("VV_PDL"); mode = ev ? (atoi(ev) != 0 ? 1 : 0) : -1; }
  return mode == 1 || (mode == -1 && g_pdl_small_stream);
}
static thread_local bool g_launch_failed = false;
// after a launch_k: the launch's own return code first (cudaGetLastError can be cleared by an unrelated call), then the
// sticky / asynchronous error state
#define VV_LAUNCH_OK()                                                     \
  do {                                                                     \
    if (vv::g_launch_failed) { vv::g_launch_failed = false; (void)cudaGetLastError(); return VV_ERR_CUDA; } \
    VV_CUDA(cudaGetLastError());                                           \
  } while (0)
template <typename... KArgs, typename... Args>
static inline void launch_k(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args&&... args) {
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  const cudaError_t e = cudaLaunchKernelEx(&cfg, kern, std::forward<Args>(args)...);
  if (e != cudaSuccess) {            // callers test g_launch_failed through VV_LAUNCH_OK() right after every launch
    g_launch_failed = true;
    set_error("kernel launch failed: %s", cudaGetErrorString(e));
  }
}

// ---- TMA tensor maps (driver entry point fetched through the runtime: no link against libcuda) ---------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// bf16 (or fp32) tensor [d2][d1][d0] (d0 contiguous), box (b0, b1, b2); swizzle span = b0 * elem bytes (128 or 64).
static int encode_map(CUtensorMap* tm, const void* ptr, int rank, const uint64_t* dims, const uint64_t* strides_b,
                      const uint32_t* box, bool is_f32 = false) {
  EncodeTiledFn fn = get_encode_fn();
  if (fn == nullptr) {
    set_error("cuTensorMapEncodeTiled entry point unavailable");
    return VV_ERR_CUDA;
  }
  cuuint64_t gd[3], gs[2];
  cuuint32_t bx[3], es[3] = {1, 1, 1};
  for (int i = 0; i < rank; ++i) { gd[i] = dims[i]; bx[i] = box[i]; }
  for (int i = 0; i + 1 < rank; ++i) gs[i] = strides_b[i];
  const uint32_t esz = is_f32 ? 4 : 2;
  const CUtensorMapSwizzle sw = (box[0] * esz == 128) ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B;
  CUresult r = fn(tm, is_f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(ptr), gd, gs, bx, es,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (%d) rank=%d dims=%llu,%llu box=%u,%u", (int)r, rank,
              (unsigned long long)dims[0], (unsigned long long)dims[1], box[0], box[1]);
    return VV_ERR_CUDA;
  }
  return VV_OK;
}

static inline uint16_t f32_to_bf16_rne(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);
  u += 0x7fffu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}

static inline long long align_up(long long v, long long a) { return (v + a - 1) / a * a; }

}  // namespace vv

using namespace vv;

void vv_decoder::mark(int kind, int stage, cudaStream_t s) {
  if (!prof_on) return;
  cudaEvent_t ev;
  if (!prof_pool.empty()) { ev = prof_pool.back(); prof_pool.pop_back(); }
  else if (cudaEventCreate(&ev) != cudaSuccess) return;
  cudaEventRecord(ev, s);
  prof_marks.push_back({ev, kind, stage});
}

static void prof_collect(vv_decoder* d, cudaStream_t s) {
  if (!d->prof_on || d->prof_marks.empty()) return;
  cudaStreamSynchronize(s);
  for (size_t i = 1; i < d->prof_marks.size(); ++i) {
    if (d->prof_marks[i].kind < 0) continue;                 // a start mark
    float ms = 0.f;
    cudaEventElapsedTime(&ms, d->prof_marks[i - 1].ev, d->prof_marks[i].ev);
    auto& acc = d->prof_acc[{d->prof_marks[i].kind, d->prof_marks[i].stage}];
    acc.first += ms;
    acc.second += 1;
  }
  for (auto& m : d->prof_marks) d->prof_pool.push_back(m.ev);
  d->prof_marks.clear();
}

extern "C" int32_t vv_decoder_set_profiling(vv_decoder* d, int32_t enabled) {
  if (!d) { set_error("null decoder"); return VV_ERR_INVALID_ARG; }
  d->prof_on = enabled != 0;
  return VV_OK;
}
extern "C" int32_t vv_decoder_read_profile(vv_decoder* d, vv_profile_record* out, int32_t cap) {
  if (!d) return 0;
  int n = 0;
  for (auto& kv : d->prof_acc) {
    if (out && n < cap) { out[n].kind = kv.first.first; out[n].stage = kv.first.second;
                          out[n].launches = kv.second.second; out[n].ms = kv.second.first; }
    ++n;
  }
  if (out) d->prof_acc.clear();
  return n;
}

// =====================================================================================================
// library
// =====================================================================================================
extern "C" int32_t vv_abi_version(void) { return VVDEC_ABI_VERSION; }
extern "C" const char* vv_last_error(void) { return get_error(); }
extern "C" const char* vv_status_string(int32_t s) {
  switch (s) {
    case VV_OK: return "ok";
    case VV_ERR_INVALID_ARG: return "invalid argument";
    case VV_ERR_BAD_SHAPE: return "bad shape";
    case VV_ERR_BAD_DTYPE: return "bad dtype";
    case VV_ERR_UNSUPPORTED: return "unsupported configuration";
    case VV_ERR_NOT_FINALIZED: return "decoder not finalized / weight missing";
    case VV_ERR_WORKSPACE: return "workspace too small";
    case VV_ERR_SLOT: return "stream slot out of range";
    case VV_ERR_CUDA: return "CUDA error";
    case VV_ERR_NO_DEVICE: return "no sm_100 CUDA device";
    case VV_ERR_UNKNOWN_TENSOR: return "unknown tensor name";
    default: return "unknown status";
  }
}

// =====================================================================================================
// decoder life cycle
// =====================================================================================================
static int stage_channels(const vv_config& c, int i) { return c.n_filters << (c.n_stages - 1 - i); }

// geometry / device checks shared by the decoder and the encoder (which mirrors the same config, :1019-1028)
static int32_t init_core(const vv_config* cfg, vv_decoder* d) {
  if (cfg->n_stages < 2 || cfg->n_stages > VV_MAX_STAGES || cfg->kernel_size != 7) {
    set_error("unsupported geometry: n_stages=%d kernel_size=%d (path implements k=7, 2..%d stages)",
              cfg->n_stages, cfg->kernel_size, VV_MAX_STAGES);
    return VV_ERR_UNSUPPORTED;
  }
  const int c_last = stage_channels(*cfg, cfg->n_stages - 1);
  if (c_last != 32 || cfg->vae_dim % 32 != 0 || cfg->vae_dim > 256) {
    set_error("unsupported geometry: n_filters=%d vae_dim=%d (kernels are specialised for n_filters=32, "
              "vae_dim a multiple of 32)", cfg->n_filters, cfg->vae_dim);
    return VV_ERR_UNSUPPORTED;
  }
  if (stage_channels(*cfg, 0) > 2048) { set_error("stage-0 channels > 2048 unsupported"); return VV_ERR_UNSUPPORTED; }
  for (int i = 0; i + 1 < cfg->n_stages; ++i)
    if (cfg->ratios[i] < 1 || cfg->ratios[i] > 16) { set_error("ratio %d out of range", cfg->ratios[i]); return VV_ERR_UNSUPPORTED; }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || cfg->device < 0 || cfg->device >= ndev) {
    set_error("CUDA device %d not available (%d devices)", cfg->device, ndev);
    return VV_ERR_NO_DEVICE;
  }
  cudaDeviceProp prop;
  VV_CUDA(cudaGetDeviceProperties(&prop, cfg->device));
  d->num_sms = prop.multiProcessorCount;
  if (prop.major != 10) {
    set_error("device %d is sm_%d%d; this library only contains sm_100a code", cfg->device, prop.major, prop.minor);
    return VV_ERR_NO_DEVICE;
  }
  d->cfg = *cfg;
  if (const char* ev = getenv("VV_GEMM_BN256")) d->opt_bn256 = atoi(ev);            // defaults; vv_decoder_set_option overrides
  if (const char* ev = getenv("VV_GEMM_BN256_WAVES")) d->opt_bn256_waves = atoi(ev);
  if (const char* ev = getenv("VV_GEMM_BN64")) d->opt_bn64 = atoi(ev);
  if (const char* ev = getenv("VV_BLOCK_KERNEL")) d->opt_block_kernel = atoi(ev) != 0;
  if (const char* ev = getenv("VV_BLOCK_WS")) d->opt_block_ws = atoi(ev) != 0;
  if (const char* ev = getenv("VV_FFN_STREAM256")) d->opt_ffn_stream256 = atoi(ev) != 0;
  d->hop = 1;
  for (int i = 0; i + 1 < cfg->n_stages; ++i) d->hop *= cfg->ratios[i];
  // causal receptive field in latent frames: sum of left taps / layer input rate
  double ctx = 6.0;
  int rate = 1;
  for (int i = 0; i < cfg->n_stages; ++i) {
    if (i > 0) { ctx += 1.0 / rate; rate *= cfg->ratios[i - 1]; }
    ctx += cfg->depths[i] * 6.0 / rate;
  }
  ctx += 6.0 / d->hop;
  d->halo = (int)std::ceil(ctx - 1e-9);
  return VV_OK;
}

extern "C" int32_t vv_decoder_create(const vv_config* cfg, vv_decoder** out) {
  if (cfg == nullptr || out == nullptr) { set_error("null argument"); return VV_ERR_INVALID_ARG; }
  *out = nullptr;
  vv_decoder* d = new vv_decoder();
  const int32_t rc = init_core(cfg, d);
  if (rc != VV_OK) { delete d; return rc; }
  *out = d;
  return VV_OK;
}


extern "C" int32_t vv_decoder_destroy(vv_decoder* d) {
  if (d == nullptr) return VV_OK;
  cudaSetDevice(d->cfg.device);
  for (void* p : d->allocs) cudaFree(p);
  delete d;
  return VV_OK;
}

// run-time options (tests force kernel instantiations the automatic routing would only pick at benchmark sizes)
extern "C" int32_t vv_decoder_set_option(vv_decoder* d, const char* key, int64_t value) {
  if (!d || !key) { set_error("null argument"); return VV_ERR_INVALID_ARG; }
  if (strcmp(key, "gemm_bn256") == 0 && value >= 0 && value <= 2) { d->opt_bn256 = (int)value; return VV_OK; }
  if (strcmp(key, "gemm_bn256_waves") == 0 && value >= 0) { d->opt_bn256_waves = (int)value; return VV_OK; }
  if (strcmp(key, "gemm_bn64") == 0 && value >= 0 && value <= 1) { d->opt_bn64 = (int)value; return VV_OK; }
  if (strcmp(key, "gemm_bn64_small") == 0 && value >= 0 && value <= 1) { d->opt_bn64_small = (int)value; return VV_OK; }
  if (strcmp(key, "block_kernel") == 0 && value >= 0 && value <= 1) { d->opt_block_kernel = (int)value; return VV_OK; }
  if (strcmp(key, "block_ws") == 0 && value >= 0 && value <= 1) { d->opt_block_ws = (int)value; return VV_OK; }
  if (strcmp(key, "ffn_stream256") == 0 && value >= 0 && value <= 1) { d->opt_ffn_stream256 = (int)value; return VV_OK; }
  set_error("unknown option '%s' or value %lld out of range", key, (long long)value);
  return VV_ERR_INVALID_ARG;
}

extern "C" int64_t vv_decoder_param_count(const vv_decoder* d) { return d ? d->param_count : 0; }
extern "C" int32_t vv_decoder_hop(const vv_decoder* d) { return d ? d->hop : 0; }
extern "C" int32_t vv_decoder_halo_frames(const vv_decoder* d) { return d ? d->halo : 0; }
extern "C" int64_t vv_decoder_take_launch_count(vv_decoder* d) {
  if (!d) return 0;
  int64_t n = d->launches;
  d->launches = 0;
  return n;
}

static int32_t stage_tensor(vv_decoder* d, const char* prefix, const char* name, const void* data, int32_t dtype,
                            const int64_t* shape, int32_t ndim, int32_t is_device) {
  if (!d || !name || !data || !shape || ndim < 1 || ndim > 3) { set_error("bad argument"); return VV_ERR_INVALID_ARG; }
  if (dtype != VV_F32 && dtype != VV_BF16) { set_error("dtype %d", dtype); return VV_ERR_BAD_DTYPE; }
  if (strncmp(name, prefix, strlen(prefix)) != 0) { set_error("'%s' is not a '%s*' parameter", name, prefix); return VV_ERR_UNKNOWN_TENSOR; }
  int64_t n = 1;
  for (int i = 0; i < ndim; ++i) n *= shape[i];
  const size_t esz = dtype == VV_F32 ? 4 : 2;
  std::vector<uint8_t> raw((size_t)n * esz);
  if (is_device) {
    cudaSetDevice(d->cfg.device);
    VV_CUDA(cudaMemcpy(raw.data(), data, raw.size(), cudaMemcpyDeviceToHost));
  } else {
    memcpy(raw.data(), data, raw.size());
  }
  std::vector<float>& dst = d->staged[name];
  dst.resize((size_t)n);
  if (dtype == VV_F32) {
    memcpy(dst.data(), raw.data(), raw.size());
  } else {
    const uint16_t* s = reinterpret_cast<const uint16_t*>(raw.data());
    for (int64_t i = 0; i < n; ++i) {
      uint32_t u = (uint32_t)s[i] << 16;
      memcpy(&dst[(size_t)i], &u, 4);
    }
  }
  d->staged_shape[name].assign(shape, shape + ndim);
  d->finalized = false;
  return VV_OK;
}

extern "C" int32_t vv_decoder_load_tensor(vv_decoder* d, const char* name, const void* data, int32_t dtype,
                                          const int64_t* shape, int32_t ndim, int32_t is_device) {
  return stage_tensor(d, "decoder.", name, data, dtype, shape, ndim, is_device);
}

// =====================================================================================================
// finalize: pack the reference state-dict into kernel-native layouts
// =====================================================================================================
namespace {

struct Packer {
  vv_decoder* d;
  int status = VV_OK;

  const std::vector<float>* get(const std::string& name, std::initializer_list<int64_t> shape) {
    auto it = d->staged.find(name);
    if (it == d->staged.end()) {
      set_error("missing tensor '%s'", name.c_str());
      status = VV_ERR_NOT_FINALIZED;
      return nullptr;
    }
    const std::vector<int64_t>& got = d->staged_shape[name];
    if (got.size() != shape.size() || !std::equal(got.begin(), got.end(), shape.begin())) {
      set_error("tensor '%s' has the wrong shape for this geometry", name.c_str());
      status = VV_ERR_BAD_SHAPE;
      return nullptr;
    }
    d->param_count += (int64_t)it->second.size();
    return &it->second;
  }
  template <typename T>
  T* upload(const std::vector<T>& h) {
    void* p = nullptr;
    if (cudaMalloc(&p, std::max<size_t>(h.size() * sizeof(T), 16)) != cudaSuccess ||
        cudaMemcpy(p, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice) != cudaSuccess) {
      set_error("weight upload failed: %s", cudaGetErrorString(cudaGetLastError()));
      status = VV_ERR_CUDA;
      return nullptr;
    }
    d->allocs.push_back(p);
    return reinterpret_cast<T*>(p);
  }
  // bf16 [N, n_taps*Ktap] K-major GEMM weight + bias + TMA map
  void dense(DenseW& w, const std::vector<uint16_t>& wp, const std::vector<float>& bias, int N, int Ktap,
             int n_taps) {
    if (status != VV_OK) return;
    w.N = N; w.Ktap = Ktap; w.n_taps = n_taps; w.K = Ktap * n_taps;
    w.BK = (Ktap % 64 == 0) ? 64 : 32;
    w.BN = (N % 128 == 0) ? 128 : (N % 64 == 0) ? 64 : 32;
    w.w = reinterpret_cast<__nv_bfloat16*>(upload(wp));
    w.bias = upload(bias);
    if (status != VV_OK) return;
    const uint64_t dims[2] = {(uint64_t)w.K, (uint64_t)N};
    const uint64_t strides[1] = {(uint64_t)w.K * 2};
    const uint32_t box[2] = {(uint32_t)w.BK, (uint32_t)w.BN};
    status = encode_map(&w.tm, w.w, 2, dims, strides, box);
    if (status == VV_OK && w.BK == 64 && N % 256 == 0) {
      const uint32_t box256[2] = {64u, 256u};
      status = encode_map(&w.tm256, w.w, 2, dims, strides, box256);
      w.has_tm256 = (status == VV_OK);
    }
    if (status == VV_OK && w.BK == 64 && w.BN == 128) {
      const uint32_t box64[2] = {64u, 64u};
      status = encode_map(&w.tm64, w.w, 2, dims, strides, box64);
      w.has_tm64 = (status == VV_OK);
    }
  }
  // one Block1D (:915-945) under state-dict prefix p: folded mixer vectors + the two FFN GEMM operands
  bool block(const std::string& p, int C, BlockW& bw);
};

std::vector<uint16_t> to_bf16(const std::vector<float>& v) {
  std::vector<uint16_t> o(v.size());
  for (size_t i = 0; i < v.size(); ++i) o[i] = f32_to_bf16_rne(v[i]);
  return o;
}

bool Packer::block(const std::string& p, int C, BlockW& bw) {
  const auto* g = get(p + "gamma", {C});
  const auto* fg = get(p + "ffn_gamma", {C});
  const auto* nw = get(p + "norm.weight", {C});
  const auto* fnw = get(p + "ffn_norm.weight", {C});
  const auto* dw = get(p + "mixer.conv.conv.conv.weight", {C, 1, 7});
  const auto* db = get(p + "mixer.conv.conv.conv.bias", {C});
  const auto* w1 = get(p + "ffn.linear1.weight", {4 * C, C});
  const auto* b1 = get(p + "ffn.linear1.bias", {4 * C});
  const auto* w2 = get(p + "ffn.linear2.weight", {C, 4 * C});
  const auto* b2 = get(p + "ffn.linear2.bias", {C});
  if (status != VV_OK) return false;
  std::vector<float> dwt((size_t)7 * C);                 // tap-major [7][C]
  for (int ch = 0; ch < C; ++ch)
    for (int t = 0; t < 7; ++t) dwt[(size_t)t * C + ch] = (*dw)[(size_t)ch * 7 + t];
  bw.gamma = upload(*g);
  bw.ffn_gamma = upload(*fg);
  bw.norm_w = upload(*nw);
  bw.ffn_norm_w = upload(*fnw);
  bw.dw_w = upload(dwt);
  bw.dw_b = upload(*db);
  {  // linear1 is packed PRE-SCALED by 0.5 (exact): every GELU epilogue works on z = x/2 (epilogue.cuh)
    std::vector<float> w1h(*w1), b1h(*b1);
    for (float& v : w1h) v *= 0.5f;
    for (float& v : b1h) v *= 0.5f;
    dense(bw.lin1, to_bf16(w1h), b1h, 4 * C, C, 1);
  }
  dense(bw.lin2, to_bf16(*w2), *b2, C, 4 * C, 1);
  if (C == 128 || C == 256) {          // host copies for ffn_stream_kernel's by-value constants
    bw.h_b1 = *b1;
    for (float& v : bw.h_b1) v *= 0.5f;
    bw.h_b2 = *b2;
    bw.h_fgamma = *fg;
  }
  if (status == VV_OK && C <= 128) {   // operands of block_tma_kernel / block_ws_kernel with the per-channel scales folded in
    std::vector<float> w1f(*w1), b1h(*b1), w2f(*w2), gb2((size_t)C);
    for (int n = 0; n < 4 * C; ++n)
      for (int k = 0; k < C; ++k) w1f[(size_t)n * C + k] *= 0.5f * (*fnw)[k];
    for (float& v : b1h) v *= 0.5f;
    for (int c = 0; c < C; ++c) {
      for (int k = 0; k < 4 * C; ++k) w2f[(size_t)c * 4 * C + k] *= (*fg)[c];
      gb2[c] = (*fg)[c] * (*b2)[c];
    }
    dense(bw.lin1f, to_bf16(w1f), b1h, 4 * C, C, 1);
    dense(bw.lin2f, to_bf16(w2f), *b2, C, 4 * C, 1);
    bw.gb2 = upload(gb2);
    bw.h_b1c = b1h;
    bw.h_gb2c = gb2;
    if (status == VV_OK && C == 128) {  // block_ws_kernel reads the folded linear1 in chunks of 64 hidden units
      const uint64_t dims[2] = {(uint64_t)C, (uint64_t)(4 * C)};
      const uint64_t strides[1] = {(uint64_t)C * 2};
      const uint32_t box[2] = {64u, 64u};
      status = encode_map(&bw.tm_w1fs, bw.lin1f.w, 2, dims, strides, box);
    }
  }
  if (status == VV_OK && (C == 128 || C == 256)) {   // ffn_stream_kernel reads linear1 in chunks of 64 hidden units
    const uint64_t dims[2] = {(uint64_t)C, (uint64_t)(4 * C)};
    const uint64_t strides[1] = {(uint64_t)C * 2};
    const uint32_t box[2] = {64u, 64u};
    status = encode_map(&bw.tm_w1s, bw.lin1.w, 2, dims, strides, box);
  }
  return status == VV_OK;
}

}  // namespace

extern "C" int32_t vv_decoder_finalize(vv_decoder* d) {
  if (!d) { set_error("null decoder"); return VV_ERR_INVALID_ARG; }
  if (d->finalized) return VV_OK;
  VV_CUDA(cudaSetDevice(d->cfg.device));
  for (void* p : d->allocs) cudaFree(p);
  d->allocs.clear();
  d->stages.clear();
  d->param_count = 0;
  const vv_config& c = d->cfg;
  Packer pk{d};
  const int D = c.vae_dim;
  long long soff = 0;
  int rate = 1;
  for (int i = 0; i < c.n_stages && pk.status == VV_OK; ++i) {
    StageW st;
    st.C = stage_channels(c, i);
    const int C = st.C;
    if (i == 0) {
      // stem SConv1d(D -> C, k7) (:864-867): Wp[co][j*D + ci] = w[co][ci][j]; tap j reads padded row t+j = x[t-6+j]
      const auto* w = pk.get("decoder.upsample_layers.0.0.conv.conv.weight", {C, D, 7});
      const auto* b = pk.get("decoder.upsample_layers.0.0.conv.conv.bias", {C});
      if (!w || !b) break;
      std::vector<uint16_t> wp((size_t)C * 7 * D);
      for (int co = 0; co < C; ++co)
        for (int ci = 0; ci < D; ++ci)
          for (int j = 0; j < 7; ++j)
            wp[(size_t)co * 7 * D + (size_t)j * D + ci] = f32_to_bf16_rne((*w)[((size_t)co * D + ci) * 7 + j]);
      pk.dense(st.up, wp, *b, C, D, 7);
      st.ratio = 1;
      st.up_state_off = soff;
      soff += align_up(6LL * D * 2, 16);
    } else {
      // SConvTranspose1d(Cin -> C, k=2r, stride r) (:875-878) in polyphase form:
      //   out[q*r + rho][co] = b[co] + sum_ci W[ci][co][rho] x[q][ci] + W[ci][co][rho + r] x[q-1][ci]
      // GEMM row n = rho*C + co; K = [tap 0: x[q-1] | tap 1: x[q]]  (tap = +row shift in the haloed buffer)
      const int r = c.ratios[i - 1];
      const int Cin = stage_channels(c, i - 1);
      const std::string p = "decoder.upsample_layers." + std::to_string(i) + ".0.convtr.convtr.";
      const auto* w = pk.get(p + "weight", {Cin, C, 2 * r});
      const auto* b = pk.get(p + "bias", {C});
      if (!w || !b) break;
      const int N = r * C;
      std::vector<uint16_t> wp((size_t)N * 2 * Cin);
      std::vector<float> bias((size_t)N);
      for (int rho = 0; rho < r; ++rho)
        for (int co = 0; co < C; ++co) {
          const size_t n = (size_t)rho * C + co;
          bias[n] = (*b)[co];
          for (int ci = 0; ci < Cin; ++ci) {
            const size_t src = ((size_t)ci * C + co) * (2 * r);
            wp[n * 2 * Cin + ci] = f32_to_bf16_rne((*w)[src + rho + r]);        // tap 0: x[q-1]
            wp[n * 2 * Cin + Cin + ci] = f32_to_bf16_rne((*w)[src + rho]);      // tap 1: x[q]
          }
        }
      pk.dense(st.up, wp, bias, N, Cin, 2);
      st.ratio = r;
      rate *= r;
      st.up_state_off = soff;
      soff += align_up((long long)Cin * 2, 16);
    }
    st.rate = rate;
    for (int j = 0; j < c.depths[i] && pk.status == VV_OK; ++j) {
      const std::string p = "decoder.stages." + std::to_string(i) + "." + std::to_string(j) + ".";
      BlockW bw;
      if (!pk.block(p, C, bw)) break;
      bw.state_off = soff;
      soff += 6LL * C * 4;
      st.blocks.push_back(bw);
    }
    d->stages.push_back(st);
  }
  if (pk.status == VV_OK) {
    const int C = stage_channels(c, c.n_stages - 1);
    const auto* w = pk.get("decoder.head.conv.conv.weight", {1, C, 7});
    const auto* b = pk.get("decoder.head.conv.conv.bias", {1});
    if (w && b) {
      std::vector<float> hw((size_t)7 * C);
      for (int ch = 0; ch < C; ++ch)
        for (int t = 0; t < 7; ++t) hw[(size_t)t * C + ch] = (*w)[(size_t)ch * 7 + t];
      d->head_w = pk.upload(hw);
      d->head_bias = (*b)[0];
      d->head_state_off = soff;
      soff += 6LL * C * 4;
    }
  }
  if (pk.status != VV_OK) return pk.status;
  d->state_slot_bytes = align_up(soff, 256);
  d->staged.clear();
  d->staged_shape.clear();
  d->finalized = true;
  return VV_OK;
}

// =====================================================================================================
// launch helpers
// =====================================================================================================
namespace {

StateRef make_state_ref(vv_state* st, const int32_t* slot_ids, long long layer_off) {
  StateRef r;
  r.slot_ids = slot_ids;
  r.parity = st ? st->parity : nullptr;
  r.base = st ? st->base : nullptr;
  r.bank_bytes = st ? st->bank_bytes : 0;
  r.slot_bytes = st ? st->slot_bytes : 0;
  r.layer_off = layer_off;
  r.n_slots = st ? st->n_slots : 0;
  return r;
}

struct EpiOut {
  int epi = EPI_BIAS_F32;
  const float* gamma = nullptr;
  const float* resid = nullptr;
  float* out_f32 = nullptr;
  __nv_bfloat16* out_bf16 = nullptr;
  int bf16_halo = 0;            // out_bf16 rows per batch item = rows + bf16_halo, first bf16_halo rows skipped
  long long bf16_rpb = 0;       // when > 0: rows per batch item of out_bf16 (overrides rows + bf16_halo)
  int bf16_rows_pb = 0;         // when > 0: the launch is a flat row range starting at bf16_row0 of [items, bf16_rows_pb]
  long long bf16_row0 = 0;
  int k_slices = 1;             // > 1 (EPI_BIAS_F32, one batch item, persistent kernel): K cut into equal slices that run as
                                // separate tiles and write partial sums into k_slices planes of out_f32 ([k_slices, rows, N])
};

// Opt a kernel into its dynamic shared memory size once per (kernel, device): the attribute is per device and
// handles on different devices may share this process; calls may come from different host threads.
template <typename F>
int ensure_smem(F* func, int bytes, int device) {
  static std::mutex mu;
  static std::set<std::pair<const void*, int>> done;
  std::lock_guard<std::mutex> lk(mu);
  const auto key = std::make_pair(reinterpret_cast<const void*>(func), device);
  if (done.count(key)) return VV_OK;
  VV_CUDA(cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  done.insert(key);
  return VV_OK;
}

template <int BN, int BK>
int launch_gemm_inst(const CUtensorMap& tmA, const DenseW& w, const GemmArgs& g, dim3 grid, int device, cudaStream_t s) {
  const size_t smem = gemm_smem_bytes<BN, BK>(g.stages);
  constexpr int per_stage = GEMM_BM * BK * 2 + BN * BK * 2;
  const int rc = ensure_smem(gemm_tc_kernel<BN, BK>,
                             (int)gemm_smem_bytes<BN, BK>(std::min(GEMM_MAX_STAGES, (96 * 1024) / per_stage)), device);
  if (rc != VV_OK) return rc;
  launch_k(gemm_tc_kernel<BN, BK>, dim3(grid), dim3(GEMM_THREADS), smem, s, tmA, w.tm, g);
  return VV_OK;
}

// D = A (*) W with the fused epilogue.  A: bf16 [nbatch, a_rpb, Ktap] (a_rpb >= rows + n_taps - 1).
int launch_gemm_impl(vv_decoder* d, const __nv_bfloat16* A, long long a_rpb, int rows, int nbatch, const DenseW& w,
                     const EpiOut& e, cudaStream_t s);
int launch_gemm(vv_decoder* d, const __nv_bfloat16* A, long long a_rpb, int rows, int nbatch, const DenseW& w,
                const EpiOut& e, cudaStream_t s, int kind = -2, int stage = 0) {
  const int rc = launch_gemm_impl(d, A, a_rpb, rows, nbatch, w, e, s);
  if (rc == VV_OK && kind != -2) d->mark(kind, stage, s);
  return rc;
}
int launch_gemm_impl(vv_decoder* d, const __nv_bfloat16* A, long long a_rpb, int rows, int nbatch, const DenseW& w,
                     const EpiOut& e, cudaStream_t s) {
  GemmArgs g;
  memset(&g, 0, sizeof(g));
  g.rows = rows; g.nbatch = nbatch; g.n_taps = w.n_taps; g.kb_per_tap = w.Ktap / w.BK; g.N = w.N;
  g.epi = e.epi; g.bias = w.bias; g.gamma = e.gamma; g.resid = e.resid;
  g.out_f32 = e.out_f32; g.of_rpb = rows; g.of_off = 0;
  g.out_bf16 = e.out_bf16; g.ob_rpb = e.bf16_rpb > 0 ? e.bf16_rpb : rows + e.bf16_halo; g.ob_off = e.bf16_halo;
  g.ob_rows_pb = e.bf16_rows_pb; g.ob_row0 = e.bf16_row0;
  int nk = g.n_taps * g.kb_per_tap;
  const int KS = e.k_slices > 1 ? e.k_slices : 1;
  if (KS > 1) {
    if (e.epi != EPI_BIAS_F32 || nbatch != 1 || nk % KS != 0 || !(w.BN == G2_BN && w.BK == G2_BK) || d->cfg.gemm_path != 0) {
      set_error("internal: k_slices=%d needs the bias epilogue, one batch item, %d k-blocks divisible and the persistent kernel", KS, nk);
      return VV_ERR_UNSUPPORTED;
    }
    g.k_slices = KS;
    nk /= KS;
  }
  const int per_stage = GEMM_BM * w.BK * 2 + w.BN * w.BK * 2;
  g.stages = std::max(1, std::min(std::min(nk, GEMM_MAX_STAGES), (96 * 1024) / per_stage));
  // tile shape over (rows, batch): 2^k rows x 128/2^k items, whichever needs fewer tiles
  int best_log2 = 7;
  long long best_tiles = -1;
  for (int k = 0; k <= 7; ++k) {
    const int bt = 1 << k, bb = 128 >> k;
    const long long tiles = (long long)((rows + bt - 1) / bt) * ((nbatch + bb - 1) / bb);
    if (best_tiles < 0 || tiles <= best_tiles) { best_tiles = tiles; best_log2 = k; }
  }
  g.bt_log2 = best_log2;
  d->launches++;
  if (d->cfg.gemm_path == 1) {
    GemmRefArgs r;
    r.A = A; r.a_rpb = a_rpb; r.W = w.w; r.Ktap = w.Ktap; r.g = g;
    const long long work = (long long)nbatch * rows * (w.N / 32);
    gemm_ref_kernel<<<(unsigned)((work + 127) / 128), 128, 0, s>>>(r);
    VV_LAUNCH_OK();
    return VV_OK;
  }
  CUtensorMap tmA;
  const uint64_t dims[3] = {(uint64_t)w.Ktap, (uint64_t)a_rpb, (uint64_t)nbatch};
  const uint64_t strides[2] = {(uint64_t)w.Ktap * 2, (uint64_t)a_rpb * w.Ktap * 2};
  const uint32_t box[3] = {(uint32_t)w.BK, 1u << best_log2, 128u >> best_log2};
  int rc = encode_map(&tmA, A, 3, dims, strides, box);
  if (rc != VV_OK) return rc;
  if (w.BN == G2_BN && w.BK == G2_BK && d->cfg.gemm_path != 3) {
    // persistent kernel with TMA-staged epilogue (gemm_tc2.cuh)
    CUtensorMap tmR, tmO;
    const uint32_t bt = 1u << best_log2, bb = 128u >> best_log2;
    const uint64_t odims[3] = {(uint64_t)w.N, (uint64_t)rows, (uint64_t)nbatch * KS};   // KS > 1: planes of partial sums
    if (e.epi == EPI_SWIGLU_BF16) {                     // [rows, N / 2] bf16: a 256-wide tile yields 128 output columns
      if (!w.has_tm256 || KS != 1) { set_error("internal: the SwiGLU epilogue needs N %% 256 == 0"); return VV_ERR_UNSUPPORTED; }
      const uint64_t sdims[3] = {(uint64_t)w.N / 2, (uint64_t)rows, (uint64_t)nbatch};
      const uint64_t ostr[2] = {(uint64_t)w.N, (uint64_t)rows * w.N};
      const uint32_t obox[3] = {64u, bt, bb};
      rc = encode_map(&tmO, e.out_bf16, 3, sdims, ostr, obox, false);
      tmR = tmO;
    } else if (e.epi == EPI_GELU_BF16) {
      const uint64_t ostr[2] = {(uint64_t)w.N * 2, (uint64_t)rows * w.N * 2};
      const uint32_t obox[3] = {64u, bt, bb};
      rc = encode_map(&tmO, e.out_bf16, 3, odims, ostr, obox, false);
      tmR = tmO;
    } else {
      const uint64_t ostr[2] = {(uint64_t)w.N * 4, (uint64_t)rows * w.N * 4};
      const uint32_t obox[3] = {32u, bt, bb};
      rc = encode_map(&tmO, e.out_f32, 3, odims, ostr, obox, true);
      if (rc == VV_OK && e.epi == EPI_RESID) rc = encode_map(&tmR, e.resid, 3, odims, ostr, obox, true);
      else tmR = tmO;
    }
    if (rc != VV_OK) return rc;
    const int num_sms = d->num_sms;
    // 256-wide MMA tiles when there are enough of them to keep the tail short (>= 2 waves: measured on the C = 2048
    // FFN); vv_decoder_set_option("gemm_bn256" / "gemm_bn256_waves") overrides (tests force every instantiation)
    best_tiles *= KS;                                    // K slices run as separate tiles
    const long long ntiles256 = best_tiles * (w.N / 256);
    // ... or when the wide tiles fit ONE wave where the 128-wide ones would need two (a 256-wide tile costs 1.5-1.8x a
    // 128-wide one: the sampler's [gate; up] GEMM at 256 streams is 144 wide tiles against 288)
    const long long ntiles128 = best_tiles * (w.N / w.BN);
    const bool wide = e.epi == EPI_SWIGLU_BF16 ||
                      (w.has_tm256 && d->opt_bn256 > 0 &&
                       (d->opt_bn256 == 2 || ntiles256 >= (long long)d->opt_bn256_waves * num_sms ||
                        (ntiles128 > num_sms && ntiles256 <= num_sms)));
    // few tiles (streaming: a frame step of 256 streams has 256 rows at stage 0, i.e. 2 x 16 tiles of 128 x 128 for linear2,
    // each streaming 2 MB of weights through ONE SM): the 64-wide tile doubles the CTAs pulling weights.  Every output
    // element still accumulates its k-blocks in the same order, so results are bit-identical to the 128 / 256-wide tiles
    // (streaming == offline stays bit-equal whichever way the two calls route).  fp32-out epilogues only (bias, residual).
    const bool small64 = !wide && w.has_tm64 && (e.epi == EPI_RESID || e.epi == EPI_BIAS_F32) && d->opt_bn64_small != 0 &&
                         d->cfg.gemm_path == 0 && best_tiles * (w.N / w.BN) * 2 <= (long long)num_sms;
    const long long ntiles = wide ? ntiles256 : small64 ? best_tiles * (w.N / 64) : best_tiles * (w.N / w.BN);
    const unsigned grid2 = (unsigned)std::min<long long>(ntiles, num_sms);
    const bool deep = nk >= 32;   // K >= 2048: mainloop-dominated
#define VV_G2_LAUNCH(E, D)                                                                                  \
    do {                                                                                                      \
      if (wide) {                                                                                             \
        rc = ensure_smem(gemm_tc2_kernel<E, D, 256>, (int)Gemm2Cfg<E, D, 256>::TOTAL, d->cfg.device);         \
        if (rc != VV_OK) return rc;                                                                           \
        launch_k(gemm_tc2_kernel<E, D, 256>, dim3(grid2), dim3(G2Warps<E>::THREADS), Gemm2Cfg<E, D, 256>::TOTAL, s, tmA, w.tm256, tmR, tmO, g); \
      } else {                                                                                                \
        rc = ensure_smem(gemm_tc2_kernel<E, D, 128>, (int)Gemm2Cfg<E, D, 128>::TOTAL, d->cfg.device);         \
        if (rc != VV_OK) return rc;                                                                           \
        launch_k(gemm_tc2_kernel<E, D, 128>, dim3(grid2), dim3(G2Warps<E>::THREADS), Gemm2Cfg<E, D, 128>::TOTAL, s, tmA, w.tm, tmR, tmO, g);  \
      }                                                                                                       \
    } while (0)
    if (e.epi == EPI_SWIGLU_BF16) {
      if (deep) {
        rc = ensure_smem(gemm_tc2_kernel<EPI_SWIGLU_BF16, true, 256>, (int)Gemm2Cfg<EPI_SWIGLU_BF16, true, 256>::TOTAL, d->cfg.device);
        if (rc != VV_OK) return rc;
        launch_k(gemm_tc2_kernel<EPI_SWIGLU_BF16, true, 256>, dim3(grid2), dim3(G2Warps<EPI_SWIGLU_BF16>::THREADS), Gemm2Cfg<EPI_SWIGLU_BF16, true, 256>::TOTAL, s, tmA, w.tm256, tmR, tmO, g);
      } else {
        rc = ensure_smem(gemm_tc2_kernel<EPI_SWIGLU_BF16, false, 256>, (int)Gemm2Cfg<EPI_SWIGLU_BF16, false, 256>::TOTAL, d->cfg.device);
        if (rc != VV_OK) return rc;
        launch_k(gemm_tc2_kernel<EPI_SWIGLU_BF16, false, 256>, dim3(grid2), dim3(G2Warps<EPI_SWIGLU_BF16>::THREADS), Gemm2Cfg<EPI_SWIGLU_BF16, false, 256>::TOTAL, s, tmA, w.tm256, tmR, tmO, g);
      }
    } else if (small64 && e.epi == EPI_BIAS_F32) {     // fp32-out epilogues on the 64-wide tile
      if (deep) {
        rc = ensure_smem(gemm_tc2_kernel<EPI_BIAS_F32, true, 64>, (int)Gemm2Cfg<EPI_BIAS_F32, true, 64>::TOTAL, d->cfg.device);
        if (rc != VV_OK) return rc;
        launch_k(gemm_tc2_kernel<EPI_BIAS_F32, true, 64>, dim3(grid2), dim3(G2Warps<EPI_BIAS_F32>::THREADS), Gemm2Cfg<EPI_BIAS_F32, true, 64>::TOTAL, s, tmA, w.tm64, tmR, tmO, g);
      } else {
        rc = ensure_smem(gemm_tc2_kernel<EPI_BIAS_F32, false, 64>, (int)Gemm2Cfg<EPI_BIAS_F32, false, 64>::TOTAL, d->cfg.device);
        if (rc != VV_OK) return rc;
        launch_k(gemm_tc2_kernel<EPI_BIAS_F32, false, 64>, dim3(grid2), dim3(G2Warps<EPI_BIAS_F32>::THREADS), Gemm2Cfg<EPI_BIAS_F32, false, 64>::TOTAL, s, tmA, w.tm64, tmR, tmO, g);
      }
    } else if (small64) {
      if (deep) {
        rc = ensure_smem(gemm_tc2_kernel<EPI_RESID, true, 64>, (int)Gemm2Cfg<EPI_RESID, true, 64>::TOTAL, d->cfg.device);
        if (rc != VV_OK) return rc;
        launch_k(gemm_tc2_kernel<EPI_RESID, true, 64>, dim3(grid2), dim3(G2Warps<EPI_RESID>::THREADS), Gemm2Cfg<EPI_RESID, true, 64>::TOTAL, s, tmA, w.tm64, tmR, tmO, g);
      } else {
        rc = ensure_smem(gemm_tc2_kernel<EPI_RESID, false, 64>, (int)Gemm2Cfg<EPI_RESID, false, 64>::TOTAL, d->cfg.device);
        if (rc != VV_OK) return rc;
        launch_k(gemm_tc2_kernel<EPI_RESID, false, 64>, dim3(grid2), dim3(G2Warps<EPI_RESID>::THREADS), Gemm2Cfg<EPI_RESID, false, 64>::TOTAL, s, tmA, w.tm64, tmR, tmO, g);
      }
    } else
    if (e.epi == EPI_BIAS_F32) { if (deep) VV_G2_LAUNCH(EPI_BIAS_F32, true); else VV_G2_LAUNCH(EPI_BIAS_F32, false); }
    else if (e.epi == EPI_GELU_BF16) { if (deep) VV_G2_LAUNCH(EPI_GELU_BF16, true); else VV_G2_LAUNCH(EPI_GELU_BF16, false); }
    else { if (deep) VV_G2_LAUNCH(EPI_RESID, true); else VV_G2_LAUNCH(EPI_RESID, false); }
#undef VV_G2_LAUNCH
    VV_LAUNCH_OK();
    return VV_OK;
  }
  if (w.BN == 64 && w.BK == G2_BK && e.epi == EPI_BIAS_F32 && d->cfg.gemm_path == 0 && d->opt_bn64 != 0) {
    // N = 64 layers (last upsampler, encoder's first strided conv and head) on the persistent kernel's 64-wide tile
    CUtensorMap tmO;
    const uint32_t bt = 1u << best_log2, bb = 128u >> best_log2;
    const uint64_t odims[3] = {(uint64_t)w.N, (uint64_t)rows, (uint64_t)nbatch};
    const uint64_t ostr[2] = {(uint64_t)w.N * 4, (uint64_t)rows * w.N * 4};
    const uint32_t obox[3] = {32u, bt, bb};
    rc = encode_map(&tmO, e.out_f32, 3, odims, ostr, obox, true);
    if (rc != VV_OK) return rc;
    const unsigned grid2 = (unsigned)std::min<long long>(best_tiles * (w.N / 64), d->num_sms);
    if (nk >= 32) {
      rc = ensure_smem(gemm_tc2_kernel<EPI_BIAS_F32, true, 64>, (int)Gemm2Cfg<EPI_BIAS_F32, true, 64>::TOTAL, d->cfg.device);
      if (rc != VV_OK) return rc;
      launch_k(gemm_tc2_kernel<EPI_BIAS_F32, true, 64>, dim3(grid2), dim3(G2Warps<EPI_BIAS_F32>::THREADS), Gemm2Cfg<EPI_BIAS_F32, true, 64>::TOTAL, s, tmA, w.tm, tmO, tmO, g);
    } else {
      rc = ensure_smem(gemm_tc2_kernel<EPI_BIAS_F32, false, 64>, (int)Gemm2Cfg<EPI_BIAS_F32, false, 64>::TOTAL, d->cfg.device);
      if (rc != VV_OK) return rc;
      launch_k(gemm_tc2_kernel<EPI_BIAS_F32, false, 64>, dim3(grid2), dim3(G2Warps<EPI_BIAS_F32>::THREADS), Gemm2Cfg<EPI_BIAS_F32, false, 64>::TOTAL, s, tmA, w.tm, tmO, tmO, g);
    }
    VV_LAUNCH_OK();
    return VV_OK;
  }
  const long long n_cta = best_tiles * (w.N / w.BN);
  if (n_cta > 0x7fffffffLL) { set_error("internal: %lld tiles exceed the launch limit", n_cta); return VV_ERR_WORKSPACE; }
  dim3 grid((unsigned)n_cta);
  if (w.BN == 128 && w.BK == 64) rc = launch_gemm_inst<128, 64>(tmA, w, g, grid, d->cfg.device, s);
  else if (w.BN == 64 && w.BK == 64) rc = launch_gemm_inst<64, 64>(tmA, w, g, grid, d->cfg.device, s);
  else if (w.BN == 32 && w.BK == 64) rc = launch_gemm_inst<32, 64>(tmA, w, g, grid, d->cfg.device, s);
  else if (w.BN == 128 && w.BK == 32) rc = launch_gemm_inst<128, 32>(tmA, w, g, grid, d->cfg.device, s);
  else if (w.BN == 64 && w.BK == 32) rc = launch_gemm_inst<64, 32>(tmA, w, g, grid, d->cfg.device, s);
  else if (w.BN == 32 && w.BK == 32) rc = launch_gemm_inst<32, 32>(tmA, w, g, grid, d->cfg.device, s);
  else { set_error("no GEMM instance for BN=%d BK=%d", w.BN, w.BK); return VV_ERR_UNSUPPORTED; }
  if (rc != VV_OK) return rc;
  VV_LAUNCH_OK();
  return VV_OK;
}

// Whole FFN half in one kernel (C = 32, 64): hidden stays on chip.  a: bf16 [R, C]; resid/out fp32 [R, C].
template <int C>
int launch_ffn_fused_inst(vv_decoder* d, const CUtensorMap& tmA, const CUtensorMap& tmR, const CUtensorMap& tmO,
                          const BlockW& bw, const FfnFusedArgs& fa, cudaStream_t s) {
  const int arc = ensure_smem(ffn_fused_kernel<C>, (int)FfnFusedSmem<C>::TOTAL, d->cfg.device);
  if (arc != VV_OK) return arc;
  const int num_sms = d->num_sms;
  const long long ntiles = (fa.R + 127) / 128;
  const unsigned grid = (unsigned)std::min<long long>(ntiles, num_sms);
  launch_k(ffn_fused_kernel<C>, dim3(grid), dim3(FF_THREADS), FfnFusedSmem<C>::TOTAL, s, tmA, tmR, tmO, bw.lin1.tm, bw.lin2.tm, fa);
  VV_LAUNCH_OK();
  return VV_OK;
}

bool ffn_fused_supported(const vv_decoder* d, int C) {
  return (d->cfg.gemm_path == 0 || d->cfg.gemm_path == 4) &&
         (C == 32 || C == 64 || C == 128 || (C == 256 && d->opt_ffn_stream256 != 0));
}

struct PadOut { long long rpb = 0; int off = 1; };   // padded bf16 copy geometry; rpb 0 = rows_pb + off

int launch_ffn_fused(vv_decoder* d, int C, const BlockW& bw, const __nv_bfloat16* a, const float* resid, float* out,
                     __nv_bfloat16* out_bf16, long long R, int rows_pb, cudaStream_t s, PadOut po = PadOut()) {
  FfnFusedArgs fa;
  fa.ob_rpb = po.rpb > 0 ? po.rpb : rows_pb + po.off; fa.ob_off = po.off;
  fa.R = R; fa.rows_pb = rows_pb; fa.b1 = bw.lin1.bias; fa.b2 = bw.lin2.bias; fa.gamma = bw.ffn_gamma;
  fa.resid = resid; fa.out = out; fa.out_bf16 = out_bf16;
  // C = 256, out == resid and no bf16 copy: the kernel adds its result to the residual in place with TMA reductions
  fa.inplace_add = (C == 256 && out == resid && out_bf16 == nullptr) ? 1 : 0;
  CUtensorMap tmA;
  const uint64_t dims[2] = {(uint64_t)C, (uint64_t)R};
  const uint64_t strides[1] = {(uint64_t)C * 2};
  const uint32_t box[2] = {(uint32_t)std::min(C, 64), 128u};   // C = 128: [128 rows x 64] K blocks (ffn_stream.cuh)
  int rc = encode_map(&tmA, a, 2, dims, strides, box);
  if (rc != VV_OK) return rc;
  CUtensorMap tmR;                                   // residual, fp32 [R, C], boxes of [128 rows, 32 floats]
  const uint64_t rstrides[1] = {(uint64_t)C * 4};
  const uint32_t rbox[2] = {32u, 128u};
  rc = encode_map(&tmR, resid, 2, dims, rstrides, rbox, true);
  if (rc != VV_OK) return rc;
  CUtensorMap tmO;                                   // output, same geometry as the residual
  rc = encode_map(&tmO, out, 2, dims, rstrides, rbox, true);
  if (rc != VV_OK) return rc;
  d->launches++;
  if (C == 32) return launch_ffn_fused_inst<32>(d, tmA, tmR, tmO, bw, fa, s);
  if (C == 64) return launch_ffn_fused_inst<64>(d, tmA, tmR, tmO, bw, fa, s);
  // C = 128 / 256: hidden on chip, weights streamed (ffn_stream.cuh); a is read in [128 rows x 64] K blocks
  if (bw.h_b1.size() != (size_t)4 * C || bw.h_b2.size() != (size_t)C || bw.h_fgamma.size() != (size_t)C) { set_error("ffn_stream: constants missing"); return VV_ERR_NOT_FINALIZED; }
  memcpy(fa.b1c, bw.h_b1.data(), sizeof(float) * 4 * C);
  memcpy(fa.b2c, bw.h_b2.data(), sizeof(float) * C);
  memcpy(fa.gammac, bw.h_fgamma.data(), sizeof(float) * C);
  const int num_sms = d->num_sms;
  const long long ntiles = (R + 127) / 128;
  const unsigned grid = (unsigned)std::min<long long>(ntiles, num_sms);
  if (R + 256LL * num_sms >= (1LL << 31)) { set_error("ffn_stream: %lld rows exceed the 31-bit row index", R); return VV_ERR_BAD_SHAPE; }
  if (C == 128) {
    rc = ensure_smem(ffn_stream_kernel<128>, (int)FfnStreamCfg<128>::TOTAL, d->cfg.device);
    if (rc != VV_OK) return rc;
    launch_k(ffn_stream_kernel<128>, dim3(grid), dim3(FS_THREADS), FfnStreamCfg<128>::TOTAL, s, tmA, tmR, tmO, bw.tm_w1s, bw.lin2.tm, fa);
  } else {
    if (!bw.lin2.has_tm256) { set_error("ffn_stream: no 256-row map for linear2"); return VV_ERR_NOT_FINALIZED; }
    rc = ensure_smem(ffn_stream_kernel<256>, (int)FfnStreamCfg<256>::TOTAL, d->cfg.device);
    if (rc != VV_OK) return rc;
    launch_k(ffn_stream_kernel<256>, dim3(grid), dim3(FS_THREADS), FfnStreamCfg<256>::TOTAL, s, tmA, tmR, tmO, bw.tm_w1s, bw.lin2.tm256, fa);
  }
  VV_LAUNCH_OK();
  return VV_OK;
}

// Whole Block1D (mixer + FFN) in one kernel (C = 32, 64; block_fused.cuh).  x: fp32 [nbatch, rows_pb, C];
// out: fp32, same shape, must NOT alias x (tiles read 6 rows of their left neighbour).
bool block_fused_supported(const vv_decoder* d, int C) {
  return d->cfg.gemm_path == 0 && (C == 32 || C == 64 || (C == 128 && d->opt_block_ws != 0));
}

template <int C, bool STATE>
int launch_block_fused_inst(vv_decoder* d, const CUtensorMap& tmO, const BlockW& bw, const BlockFusedArgs& a,
                            cudaStream_t s) {
  const int arc = ensure_smem(block_fused_kernel<C, STATE>, (int)BlockFusedCfg<C>::TOTAL, d->cfg.device);
  if (arc != VV_OK) return arc;
  const int num_sms = d->num_sms;
  const long long ntiles = (long long)a.nbatch * a.tiles_pb;
  const unsigned grid = (unsigned)std::min<long long>(ntiles, num_sms);
  launch_k(block_fused_kernel<C, STATE>, dim3(grid), dim3(BlockFusedCfg<C>::THREADS), BlockFusedCfg<C>::TOTAL, s, tmO, bw.lin1.tm, bw.lin2.tm, a);
  VV_LAUNCH_OK();
  return VV_OK;
}

template <int C, bool STATE>
int launch_block_tma_inst(vv_decoder* d, const CUtensorMap& tmX, const CUtensorMap& tmO, const BlockW& bw,
                          const BlockTmaArgs& a, cudaStream_t s) {
  const int arc = ensure_smem(block_tma_kernel<C, STATE>, (int)BlockTmaCfg<C>::TOTAL, d->cfg.device);
  if (arc != VV_OK) return arc;
  const long long ntiles = (long long)a.nbatch * a.tiles_pb;
  const unsigned grid = (unsigned)std::min<long long>(ntiles, d->num_sms);
  launch_k(block_tma_kernel<C, STATE>, dim3(grid), dim3(BlockTmaCfg<C>::THREADS), BlockTmaCfg<C>::TOTAL, s, tmX, tmO,
           bw.lin1f.tm, bw.lin2f.tm, a);
  VV_LAUNCH_OK();
  return VV_OK;
}

int launch_block_tma(vv_decoder* d, int C, const BlockW& bw, const float* x, float* out, __nv_bfloat16* out_bf16,
                     int nbatch, int rows_pb, vv_state* st, const int32_t* slots, cudaStream_t s, PadOut po) {
  BlockTmaArgs a;
  a.x = x; a.nbatch = nbatch; a.rows_pb = rows_pb; a.tiles_pb = (rows_pb + 127) / 128; a.eps = d->cfg.eps;
  a.norm_w = bw.norm_w; a.dw_w = bw.dw_w; a.dw_b = bw.dw_b; a.gamma = bw.gamma;
  a.out_bf16 = out_bf16;
  if (bw.h_b1c.size() != (size_t)4 * C || bw.h_gb2c.size() != (size_t)C || 4 * C > 256) { set_error("block_tma: constants missing"); return VV_ERR_NOT_FINALIZED; }
  memset(a.b1c, 0, sizeof(a.b1c));
  memset(a.gb2c, 0, sizeof(a.gb2c));
  memcpy(a.b1c, bw.h_b1c.data(), sizeof(float) * 4 * C);
  memcpy(a.gb2c, bw.h_gb2c.data(), sizeof(float) * C);
  a.st = make_state_ref(st, slots, bw.state_off);
  a.ob_rpb = po.rpb > 0 ? po.rpb : rows_pb + po.off; a.ob_off = po.off;
  if ((long long)nbatch * a.tiles_pb >= (1LL << 31) / 8) { set_error("block_tma: too many tiles"); return VV_ERR_BAD_SHAPE; }
  // x and out: fp32 [nbatch, rows_pb, C]; load boxes [8 halo + 128 rows, 32 floats] (rows before the item start and past
  // its end are zero-filled), store boxes [128 rows, 32 floats] (rows past the end are clipped)
  CUtensorMap tmX, tmO;
  const uint64_t dims[3] = {(uint64_t)C, (uint64_t)rows_pb, (uint64_t)nbatch};
  const uint64_t strides[2] = {(uint64_t)C * 4, (uint64_t)rows_pb * C * 4};
  const uint32_t xbox[3] = {32u, (uint32_t)BlockTmaCfg<32>::XROWS, 1u};
  const uint32_t obox[3] = {32u, 128u, 1u};
  int rc = encode_map(&tmX, x, 3, dims, strides, xbox, true);
  if (rc == VV_OK) rc = encode_map(&tmO, out, 3, dims, strides, obox, true);
  if (rc != VV_OK) return rc;
  d->launches++;
  const bool streaming = a.st.base != nullptr;
  if (C == 32) return streaming ? launch_block_tma_inst<32, true>(d, tmX, tmO, bw, a, s) : launch_block_tma_inst<32, false>(d, tmX, tmO, bw, a, s);
  return streaming ? launch_block_tma_inst<64, true>(d, tmX, tmO, bw, a, s) : launch_block_tma_inst<64, false>(d, tmX, tmO, bw, a, s);
}

// Whole Block1D at C = 128 (block_ws.cuh): x tile staged by TMA, weights streamed, a tile inside the slot.
template <bool STATE>
int launch_block_ws_inst(vv_decoder* d, const CUtensorMap& tmX, const CUtensorMap& tmO, const BlockW& bw,
                         const BlockWsArgs& a, cudaStream_t s) {
  const int arc = ensure_smem(block_ws_kernel<STATE>, (int)BlockWsCfg::TOTAL, d->cfg.device);
  if (arc != VV_OK) return arc;
  const long long ntiles = (long long)a.nbatch * a.tiles_pb;
  const unsigned grid = (unsigned)std::min<long long>(ntiles, d->num_sms);
  launch_k(block_ws_kernel<STATE>, dim3(grid), dim3(BlockWsCfg::THREADS), BlockWsCfg::TOTAL, s, tmX, tmO, bw.tm_w1fs,
           bw.lin2f.tm, a);
  VV_LAUNCH_OK();
  return VV_OK;
}

int launch_block_ws(vv_decoder* d, const BlockW& bw, const float* x, float* out, __nv_bfloat16* out_bf16, int nbatch,
                    int rows_pb, vv_state* st, const int32_t* slots, cudaStream_t s, PadOut po) {
  constexpr int C = BlockWsCfg::C;
  BlockWsArgs a;
  a.nbatch = nbatch; a.rows_pb = rows_pb; a.tiles_pb = (rows_pb + 127) / 128; a.eps = d->cfg.eps;
  a.norm_w = bw.norm_w; a.dw_w = bw.dw_w; a.dw_b = bw.dw_b; a.gamma = bw.gamma;
  a.out_bf16 = out_bf16;
  if (bw.h_b1c.size() != (size_t)4 * C || bw.h_gb2c.size() != (size_t)C) { set_error("block_ws: constants missing"); return VV_ERR_NOT_FINALIZED; }
  memcpy(a.b1c, bw.h_b1c.data(), sizeof(a.b1c));
  memcpy(a.gb2c, bw.h_gb2c.data(), sizeof(a.gb2c));
  a.st = make_state_ref(st, slots, bw.state_off);
  a.ob_rpb = po.rpb > 0 ? po.rpb : rows_pb + po.off; a.ob_off = po.off;
  if ((long long)nbatch * a.tiles_pb >= (1LL << 31) / 8) { set_error("block_ws: too many tiles"); return VV_ERR_BAD_SHAPE; }
  CUtensorMap tmX, tmO;
  const uint64_t dims[3] = {(uint64_t)C, (uint64_t)rows_pb, (uint64_t)nbatch};
  const uint64_t strides[2] = {(uint64_t)C * 4, (uint64_t)rows_pb * C * 4};
  const uint32_t xbox[3] = {32u, (uint32_t)BlockWsCfg::XROWS, 1u};
  const uint32_t obox[3] = {32u, 128u, 1u};
  int rc = encode_map(&tmX, x, 3, dims, strides, xbox, true);
  if (rc == VV_OK) rc = encode_map(&tmO, out, 3, dims, strides, obox, true);
  if (rc != VV_OK) return rc;
  d->launches++;
  return a.st.base != nullptr ? launch_block_ws_inst<true>(d, tmX, tmO, bw, a, s) : launch_block_ws_inst<false>(d, tmX, tmO, bw, a, s);
}

int launch_block_fused(vv_decoder* d, int C, const BlockW& bw, const float* x, float* out, __nv_bfloat16* out_bf16,
                       int nbatch, int rows_pb, vv_state* st, const int32_t* slots, cudaStream_t s,
                       PadOut po = PadOut()) {
  if (x == out) { set_error("block_fused: input and output must not alias"); return VV_ERR_INVALID_ARG; }
  if (C == 128) return launch_block_ws(d, bw, x, out, out_bf16, nbatch, rows_pb, st, slots, s, po);
  if (d->opt_block_kernel == 1) return launch_block_tma(d, C, bw, x, out, out_bf16, nbatch, rows_pb, st, slots, s, po);
  BlockFusedArgs a;
  a.x = x; a.nbatch = nbatch; a.rows_pb = rows_pb; a.tiles_pb = (rows_pb + 127) / 128; a.eps = d->cfg.eps;
  a.norm_w = bw.norm_w; a.dw_w = bw.dw_w; a.dw_b = bw.dw_b; a.gamma = bw.gamma; a.ffn_norm_w = bw.ffn_norm_w;
  a.b1 = bw.lin1.bias; a.b2 = bw.lin2.bias; a.ffn_gamma = bw.ffn_gamma; a.out_bf16 = out_bf16;
  a.st = make_state_ref(st, slots, bw.state_off);
  a.ob_rpb = po.rpb > 0 ? po.rpb : rows_pb + po.off; a.ob_off = po.off;
  if ((long long)nbatch * a.tiles_pb >= (1LL << 31) / 8) { set_error("block_fused: too many tiles"); return VV_ERR_BAD_SHAPE; }
  CUtensorMap tmO;                                   // output fp32 [nbatch, rows_pb, C], boxes of [128 rows, 32 floats]
  const uint64_t dims[3] = {(uint64_t)C, (uint64_t)rows_pb, (uint64_t)nbatch};
  const uint64_t strides[2] = {(uint64_t)C * 4, (uint64_t)rows_pb * C * 4};
  const uint32_t box[3] = {32u, 128u, 1u};
  int rc = encode_map(&tmO, out, 3, dims, strides, box, true);
  if (rc != VV_OK) return rc;
  d->launches++;
  const bool streaming = a.st.base != nullptr;
  if (C == 32) return streaming ? launch_block_fused_inst<32, true>(d, tmO, bw, a, s) : launch_block_fused_inst<32, false>(d, tmO, bw, a, s);
  return streaming ? launch_block_fused_inst<64, true>(d, tmO, bw, a, s) : launch_block_fused_inst<64, false>(d, tmO, bw, a, s);
}

// FFN half as two GEMMs (C >= 256).  At C = 256 the pair is HBM-bound by the bf16 hidden tensor's round trip (26 bytes
// per element), so an L2-resident variant exists: run the pair in row chunks whose hidden slice fits the 126 MB L2 and
// reuse one hidden buffer.  MEASURED SLOWER on B200 (K2 stage 3, 15 blocks: 31.2 ms unchunked, 43.5 / 47.5 / 65.5 ms
// with 64 / 32 / 16 MB chunks — 70+ short persistent launches per block cost more in prologue and tile quantisation
// than the saved traffic), so it is off by default; VV_FFN_CHUNK_MB=<MB> turns it on for experiments.
//   a: bf16 [R, C]; resid (y), out (x): fp32 [R, C]; h: bf16 scratch [R, 4C]; optional padded bf16 copy.
long long ffn_chunk_rows(const vv_decoder* d, int C, long long R) {
  static long long chunk_mb = -1;
  if (chunk_mb < 0) {
    const char* e = getenv("VV_FFN_CHUNK_MB");
    chunk_mb = e ? atoll(e) : 0;
  }
  const bool tc2 = d->cfg.gemm_path == 0 || d->cfg.gemm_path == 4 || d->cfg.gemm_path == 2;
  if (!tc2 || chunk_mb <= 0 || C > 256 || R * C * 8 <= chunk_mb * (1LL << 20) * 2) return R;
  const long long rows = (chunk_mb << 20) / ((long long)C * 8);
  return std::max<long long>(128, rows / 128 * 128);
}

int launch_ffn_pair(vv_decoder* d, int C, const BlockW& bw, const __nv_bfloat16* A, const float* Y, float* X,
                    __nv_bfloat16* H, __nv_bfloat16* XB, PadOut po, int nb, int rows_pb, int stage, cudaStream_t s) {
  const long long R = (long long)nb * rows_pb;
  const long long chunk = ffn_chunk_rows(d, C, R);
  int rc;
  if (chunk >= R) {
    EpiOut e1;
    e1.epi = EPI_GELU_BF16; e1.out_bf16 = H;
    rc = launch_gemm(d, A, R, (int)R, 1, bw.lin1, e1, s, VV_K_FFN1, stage);
    if (rc != VV_OK) return rc;
    EpiOut e2;
    e2.epi = EPI_RESID; e2.gamma = bw.ffn_gamma; e2.resid = Y; e2.out_f32 = X;
    if (XB != nullptr) {
      e2.out_bf16 = XB; e2.bf16_halo = po.off; e2.bf16_rpb = po.rpb > 0 ? po.rpb : rows_pb + po.off;
      return launch_gemm(d, H, rows_pb, rows_pb, nb, bw.lin2, e2, s, VV_K_FFN2, stage);
    }
    return launch_gemm(d, H, R, (int)R, 1, bw.lin2, e2, s, VV_K_FFN2, stage);
  }
  for (long long r0 = 0; r0 < R; r0 += chunk) {
    const int n = (int)std::min(chunk, R - r0);
    EpiOut e1;
    e1.epi = EPI_GELU_BF16; e1.out_bf16 = H;
    rc = launch_gemm(d, A + r0 * C, n, n, 1, bw.lin1, e1, s);
    if (rc != VV_OK) return rc;
    EpiOut e2;
    e2.epi = EPI_RESID; e2.gamma = bw.ffn_gamma; e2.resid = Y + r0 * C; e2.out_f32 = X + r0 * C;
    if (XB != nullptr) {
      e2.out_bf16 = XB; e2.bf16_halo = po.off; e2.bf16_rpb = po.rpb > 0 ? po.rpb : rows_pb + po.off;
      e2.bf16_rows_pb = rows_pb; e2.bf16_row0 = r0;
    }
    rc = launch_gemm(d, H, n, n, 1, bw.lin2, e2, s);
    if (rc != VV_OK) return rc;
  }
  d->mark(VV_K_FFN_PAIR, stage, s);  // one profiling record for the whole chunked pair
  return VV_OK;
}

int pick_strip(long long total_rows, int rows) {
  // longer strips amortise the 6 warm-up rows; shorter ones keep the machine full on small problems
  int strip = 113;
  while (strip > 15 && total_rows / strip < 148LL * 64) strip = (strip - 1) / 2 + 1;
  return std::max(1, std::min(strip, rows));
}

int launch_mixer(vv_decoder* d, const BlockW& bw, int C, const float* x, float* y, __nv_bfloat16* a, int nbatch,
                 int rows, vv_state* st, const int32_t* slots, cudaStream_t s) {
  MixerArgs p;
  p.x = x; p.y = y; p.a = a;
  p.norm_w = bw.norm_w; p.dw_w = bw.dw_w; p.dw_b = bw.dw_b; p.gamma = bw.gamma; p.ffn_norm_w = bw.ffn_norm_w;
  p.nbatch = nbatch; p.rows = rows; p.C = C; p.eps = d->cfg.eps;
  p.strip = pick_strip((long long)nbatch * rows, rows);
  p.strips_per_batch = (rows + p.strip - 1) / p.strip;
  p.st = make_state_ref(st, slots, bw.state_off);
  const long long groups = (long long)nbatch * p.strips_per_batch;
  const int lpr = C / 4;
  d->launches++;
#define VV_MIXER_CASE(L)                                                                         \
  case L:                                                                                        \
    if (L <= 32) launch_k(mixer_kernel<L>, dim3((unsigned)((groups * L + 127) / 128)), dim3(128), 0, s, p);        \
    else launch_k(mixer_kernel<L>, dim3((unsigned)groups), dim3(L), 0, s, p);                                      \
    break;
  switch (lpr) {
    VV_MIXER_CASE(8) VV_MIXER_CASE(16) VV_MIXER_CASE(32) VV_MIXER_CASE(64) VV_MIXER_CASE(128)
    VV_MIXER_CASE(256) VV_MIXER_CASE(512)
    default: set_error("no mixer kernel for C=%d", C); return VV_ERR_UNSUPPORTED;
  }
#undef VV_MIXER_CASE
  VV_LAUNCH_OK();
  return VV_OK;
}

}  // namespace

// =====================================================================================================
// one decode pass over [nb, T] frames, everything resident in the workspace
// =====================================================================================================
namespace {

struct StageBufs { long long x, y, a, h, xb, total; };

// bytes of the per-stage buffers for nb items x T frames (all offsets 1 KB aligned)
StageBufs stage_bufs(const vv_decoder* d, int i, long long nb, long long T) {
  const StageW& st = d->stages[i];
  const long long rows_pb = T * st.rate, R = nb * rows_pb, C = st.C;
  StageBufs b;
  long long off = 0;
  b.x = off; off += align_up(R * C * 4, 1024);
  b.y = off; off += align_up(R * C * 4, 1024);
  b.a = off; off += align_up(R * C * 2, 1024);
  b.h = off; off += align_up(R * C * 4 * 2, 1024);
  b.xb = off; off += align_up(nb * (rows_pb + 1) * C * 2, 1024);
  b.total = off;
  return b;
}

long long pass_bytes(const vv_decoder* d, long long nb, long long T) {
  long long arena = 0;
  for (int i = 0; i < (int)d->stages.size(); ++i) arena = std::max(arena, stage_bufs(d, i, nb, T).total);
  return align_up(nb * (T + 6) * d->cfg.vae_dim * 2, 1024) + 2 * arena + 1024;
}

struct LatView { const void* p; int is_bf16; long long sb, st, sd; };

int run_pass(vv_decoder* d, const LatView& lat, int nb, int T, void* audio, int out_dtype, long long out_bstride,
             uint8_t* ws, vv_state* st, const int32_t* slots, cudaStream_t s) {
  const int D = d->cfg.vae_dim;
  long long arena = 0;
  const int S = (int)d->stages.size();
  for (int i = 0; i < S; ++i) arena = std::max(arena, stage_bufs(d, i, nb, T).total);
  __nv_bfloat16* LB = reinterpret_cast<__nv_bfloat16*>(ws);
  uint8_t* ar[2];
  ar[0] = ws + align_up((long long)nb * (T + 6) * D * 2, 1024);
  ar[1] = ar[0] + arena;

  d->mark(-1, 0, s);
  {  // latents -> bf16 [nb, T+6, D] with the causal context rows
    PrepArgs p;
    p.lat = lat.p; p.lat_is_bf16 = lat.is_bf16; p.sb = lat.sb; p.st = lat.st; p.sd = lat.sd;
    p.out = LB; p.nbatch = nb; p.T = T; p.D = D;
    p.st_ref = make_state_ref(st, slots, d->stages[0].up_state_off);
    const long long total = (long long)nb * (T + 6) * D;
    launch_k(prep_latents_kernel, dim3((unsigned)std::min<long long>((total + 255) / 256, 148 * 16)), dim3(256), 0, s, p);
    d->launches++;
    VV_LAUNCH_OK();
    d->mark(VV_K_PREP, 0, s);
  }
  int rc;
  for (int i = 0; i < S; ++i) {
    const StageW& sw = d->stages[i];
    const StageBufs bo = stage_bufs(d, i, nb, T);
    uint8_t* base = ar[i & 1];
    float* X = reinterpret_cast<float*>(base + bo.x);
    float* Y = reinterpret_cast<float*>(base + bo.y);
    __nv_bfloat16* A = reinterpret_cast<__nv_bfloat16*>(base + bo.a);
    __nv_bfloat16* H = reinterpret_cast<__nv_bfloat16*>(base + bo.h);
    __nv_bfloat16* XB = reinterpret_cast<__nv_bfloat16*>(base + bo.xb);
    const int rows_pb = T * sw.rate;
    const int C = sw.C;
    if (i == 0) {
      EpiOut e;
      e.epi = EPI_BIAS_F32; e.out_f32 = X;
      rc = launch_gemm(d, LB, T + 6, T, nb, sw.up, e, s, VV_K_STEM, 0);
      if (rc != VV_OK) return rc;
    } else {
      // polyphase transposed conv from the previous stage's haloed bf16 copy
      const StageW& pw = d->stages[i - 1];
      const StageBufs pbo = stage_bufs(d, i - 1, nb, T);
      __nv_bfloat16* PXB = reinterpret_cast<__nv_bfloat16*>(ar[(i - 1) & 1] + pbo.xb);
      const int prow = T * pw.rate;
      HaloArgs hp;
      hp.xb = PXB; hp.nbatch = nb; hp.rows = prow; hp.C = pw.C;
      hp.st_ref = make_state_ref(st, slots, sw.up_state_off);
      launch_k(upsample_halo_kernel, dim3((unsigned)std::min((nb * pw.C + 255) / 256, 148 * 8)), dim3(256), 0, s, hp);
      d->launches++;
      VV_LAUNCH_OK();
      d->mark(VV_K_HALO, i, s);
      EpiOut e;
      e.epi = EPI_BIAS_F32; e.out_f32 = X;
      rc = launch_gemm(d, PXB, prow + 1, prow, nb, sw.up, e, s, VV_K_UPSAMPLE, i);
      if (rc != VV_OK) return rc;
    }
    const int nblk = (int)sw.blocks.size();
    for (int j = 0; j < nblk; ++j) {
      const BlockW& bw = sw.blocks[j];
      if (block_fused_supported(d, C)) {
        // whole block in one kernel, ping-pong between the two fp32 buffers of the stage (X holds the result)
        const bool feeds = (j == nblk - 1) && (i + 1 < S);
        rc = launch_block_fused(d, C, bw, X, Y, feeds ? XB : nullptr, nb, rows_pb, st, slots, s);
        if (rc != VV_OK) return rc;
        d->mark(VV_K_BLOCK_FUSED, i, s);
        std::swap(X, Y);
        continue;
      }
      rc = launch_mixer(d, bw, C, X, Y, A, nb, rows_pb, st, slots, s);
      if (rc != VV_OK) return rc;
      d->mark(VV_K_MIXER, i, s);
      const bool feeds_up = (j == nblk - 1) && (i + 1 < S);
      if (ffn_fused_supported(d, C)) {
        if (C == 256 && !feeds_up) {     // in place: y += gamma_f * ffn(a) by TMA reductions; y becomes the block output
          rc = launch_ffn_fused(d, C, bw, A, Y, Y, nullptr, (long long)nb * rows_pb, rows_pb, s);
          std::swap(X, Y);
        } else {
          rc = launch_ffn_fused(d, C, bw, A, Y, X, feeds_up ? XB : nullptr, (long long)nb * rows_pb, rows_pb, s);
        }
        if (rc != VV_OK) return rc;
        d->mark(VV_K_FFN_FUSED, i, s);
        continue;
      }
      rc = launch_ffn_pair(d, C, bw, A, Y, X, H, feeds_up ? XB : nullptr, PadOut(), nb, rows_pb, i, s);
      if (rc != VV_OK) return rc;
    }
    if (nblk == 0 && i + 1 < S) { set_error("stage %d has no blocks", i); return VV_ERR_UNSUPPORTED; }
    if (i == S - 1) {
      HeadArgs hp;
      hp.x = X; hp.w = d->head_w; hp.bias = d->head_bias; hp.out = audio; hp.out_is_bf16 = (out_dtype == VV_BF16);
      hp.out_batch_stride = out_bstride; hp.nbatch = nb; hp.rows = rows_pb;
      hp.strip = pick_strip((long long)nb * rows_pb, rows_pb);
      hp.strips_per_batch = (rows_pb + hp.strip - 1) / hp.strip;
      hp.st = make_state_ref(st, slots, d->head_state_off);
      const long long groups = (long long)nb * hp.strips_per_batch;
      launch_k(head_kernel, dim3((unsigned)((groups * 8 + 127) / 128)), dim3(128), 0, s, hp);
      d->launches++;
      VV_LAUNCH_OK();
      d->mark(VV_K_HEAD, i, s);
    }
  }
  if (st != nullptr) {
    launch_k(state_flip_kernel, dim3((nb + 127) / 128), dim3(128), 0, s, slots, nb, st->parity, st->n_slots, st->err_dev);
    d->launches++;
    VV_LAUNCH_OK();
    d->mark(VV_K_STATE, 0, s);
  }
  return VV_OK;
}

}  // namespace

// =====================================================================================================
// decode
// =====================================================================================================
extern "C" int64_t vv_decoder_workspace_bytes(const vv_decoder* d, int32_t B, int32_t T) {
  if (!d || !d->finalized || B < 1 || T < 1) return 0;
  return pass_bytes(d, B, T);
}

static int32_t state_take_error(vv_state* st);
static int32_t decode_impl(vv_decoder* d, const void* latents, int32_t lat_dtype, int32_t layout, int32_t B,
                             int32_t T, void* audio, int32_t out_dtype, void* workspace, int64_t workspace_bytes,
                             vv_state* state, const int32_t* slot_ids, void* cuda_stream) {
  if (!d || !latents || !audio || !workspace) { set_error("null argument"); return VV_ERR_INVALID_ARG; }
  if (!d->finalized) { set_error("vv_decoder_finalize has not succeeded"); return VV_ERR_NOT_FINALIZED; }
  if (B < 1 || T < 1) { set_error("B=%d T=%d", B, T); return VV_ERR_BAD_SHAPE; }
  if ((lat_dtype != VV_F32 && lat_dtype != VV_BF16) || (out_dtype != VV_F32 && out_dtype != VV_BF16)) {
    set_error("dtype must be f32 or bf16");
    return VV_ERR_BAD_DTYPE;
  }
  if (layout != VV_LAYOUT_BTD && layout != VV_LAYOUT_BDT) { set_error("layout %d", layout); return VV_ERR_INVALID_ARG; }
  if (state != nullptr && slot_ids == nullptr) { set_error("state given without slot_ids"); return VV_ERR_INVALID_ARG; }
  if (state != nullptr && state->dec != d) { set_error("state belongs to another decoder"); return VV_ERR_INVALID_ARG; }
  if (state != nullptr) { const int32_t prev = state_take_error(state); if (prev != VV_OK) return prev; }
  if (((uintptr_t)workspace & 1023) != 0) { set_error("workspace must be 1 KB aligned"); return VV_ERR_INVALID_ARG; }
  VV_CUDA(cudaSetDevice(d->cfg.device));
  cudaStream_t s = reinterpret_cast<cudaStream_t>(cuda_stream);
  const int D = d->cfg.vae_dim;
  const long long hop = d->hop;
  const size_t lsz = lat_dtype == VV_F32 ? 4 : 2, osz = out_dtype == VV_F32 ? 4 : 2;
  LatView lv;
  lv.is_bf16 = (lat_dtype == VV_BF16);
  if (layout == VV_LAYOUT_BTD) { lv.sb = (long long)T * D; lv.st = D; lv.sd = 1; }
  else { lv.sb = (long long)D * T; lv.st = 1; lv.sd = T; }
  uint8_t* ws = reinterpret_cast<uint8_t*>(workspace);
  const long long out_bstride = (long long)T * hop;

  if (pass_bytes(d, B, T) <= workspace_bytes) {          // whole call in one pass
    lv.p = latents;
    return run_pass(d, lv, B, T, audio, out_dtype, out_bstride, ws, state, slot_ids, s);
  }
  if (pass_bytes(d, 1, T) <= workspace_bytes) {          // split by utterance (independent items)
    int nb = 1;
    while (nb < B && pass_bytes(d, nb + 1, T) <= workspace_bytes) ++nb;
    for (int b0 = 0; b0 < B; b0 += nb) {
      const int n = std::min(nb, B - b0);
      lv.p = reinterpret_cast<const uint8_t*>(latents) + (size_t)b0 * lv.sb * lsz;
      int rc = run_pass(d, lv, n, T, reinterpret_cast<uint8_t*>(audio) + (size_t)b0 * out_bstride * osz, out_dtype,
                        out_bstride, ws, state, state ? slot_ids + b0 : nullptr, s);
      if (rc != VV_OK) return rc;
    }
    return VV_OK;
  }
  // split in time, one item at a time, carrying the causal context in a state (the caller's, or a
  // zero-initialised scratch one carved from the workspace: offline == streaming with zero context)
  vv_state scratch;
  long long reserve = 0;
  if (state == nullptr) {
    reserve = align_up(2 * d->state_slot_bytes + 1024, 1024);
    if (reserve >= workspace_bytes) { set_error("workspace too small for time chunking"); return VV_ERR_WORKSPACE; }
    scratch.dec = d; scratch.n_slots = 1; scratch.slot_bytes = d->state_slot_bytes;
    scratch.bank_bytes = d->state_slot_bytes;
    scratch.base = ws + workspace_bytes - reserve;
    scratch.parity = reinterpret_cast<int*>(scratch.base + 2 * d->state_slot_bytes);   // [0]=parity, [1]=slot id 0
  }
  const long long avail = workspace_bytes - reserve;
  int Tc = 0;
  for (int lo = 1, hi = T; lo <= hi;) {                  // largest chunk that fits
    const int mid = lo + (hi - lo) / 2;
    if (pass_bytes(d, 1, mid) <= avail) { Tc = mid; lo = mid + 1; } else { hi = mid - 1; }
  }
  if (Tc < 1) {
    set_error("workspace of %lld bytes cannot hold one frame (%lld needed)", (long long)workspace_bytes,
              pass_bytes(d, 1, 1) + reserve);
    return VV_ERR_WORKSPACE;
  }
  for (int b = 0; b < B; ++b) {
    vv_state* stp = state;
    const int32_t* sl = state ? slot_ids + b : nullptr;
    if (state == nullptr) {
      VV_CUDA(cudaMemsetAsync(scratch.base, 0, (size_t)(2 * d->state_slot_bytes + 16), s));
      stp = &scratch;
      sl = scratch.parity + 1;
    }
    for (int t0 = 0; t0 < T; t0 += Tc) {
      const int n = std::min(Tc, T - t0);
      lv.p = reinterpret_cast<const uint8_t*>(latents) + ((size_t)b * lv.sb + (size_t)t0 * lv.st) * lsz;
      int rc = run_pass(d, lv, 1, n, reinterpret_cast<uint8_t*>(audio) + ((size_t)b * out_bstride + (size_t)t0 * hop) * osz,
                        out_dtype, out_bstride, ws, stp, sl, s);
      if (rc != VV_OK) return rc;
    }
  }
  return VV_OK;
}

extern "C" int32_t vv_decode(vv_decoder* d, const void* latents, int32_t lat_dtype, int32_t layout, int32_t B,
                             int32_t T, void* audio, int32_t out_dtype, void* workspace, int64_t workspace_bytes,
                             vv_state* state, const int32_t* slot_ids, void* cuda_stream) {
  PdlScope pdl(state != nullptr && (long long)B * T <= 2048);
  const int32_t rc = decode_impl(d, latents, lat_dtype, layout, B, T, audio, out_dtype, workspace, workspace_bytes,
                                 state, slot_ids, cuda_stream);
  if (d != nullptr) prof_collect(d, reinterpret_cast<cudaStream_t>(cuda_stream));
  return rc;
}

// =====================================================================================================
// streaming state
// =====================================================================================================
// A slot id outside [0, n_slots) reached a kernel in an earlier call on this state: report it once (VV_ERR_SLOT).
static int32_t state_take_error(vv_state* st) {
  if (st == nullptr || st->err_host == nullptr) return VV_OK;
  const int e = *reinterpret_cast<volatile int*>(st->err_host);
  if (e == 0) return VV_OK;
  *reinterpret_cast<volatile int*>(st->err_host) = 0;
  set_error("an earlier call on this state used slot id %d, outside [0, %d): that row was decoded with a zero context "
            "and its state was not saved (create the cache with more slots)", e < 0 ? e : e - 1, st->n_slots);
  return VV_ERR_SLOT;
}

extern "C" int32_t vv_state_create(const vv_decoder* d, int32_t n_slots, vv_state** out) {
  if (!d || !out || n_slots < 1) { set_error("bad argument"); return VV_ERR_INVALID_ARG; }
  if (!d->finalized) { set_error("decoder not finalized"); return VV_ERR_NOT_FINALIZED; }
  VV_CUDA(cudaSetDevice(d->cfg.device));
  vv_state* st = new vv_state();
  st->dec = d; st->n_slots = n_slots; st->slot_bytes = d->state_slot_bytes; st->device = d->cfg.device;
  st->bank_bytes = st->slot_bytes * n_slots;
  void* p = nullptr;
  cudaError_t e = cudaMalloc(&p, (size_t)(2 * st->bank_bytes));
  if (e == cudaSuccess) { st->base = (uint8_t*)p; e = cudaMalloc(&p, sizeof(int) * n_slots); }
  if (e == cudaSuccess) { st->parity = (int*)p; e = cudaMemset(st->base, 0, (size_t)(2 * st->bank_bytes)); }
  if (e == cudaSuccess) e = cudaMemset(st->parity, 0, sizeof(int) * n_slots);
  if (e == cudaSuccess) e = cudaHostAlloc(&p, sizeof(int), cudaHostAllocMapped);
  if (e == cudaSuccess) {
    st->err_host = (int*)p;
    *st->err_host = 0;
    e = cudaHostGetDevicePointer(&p, st->err_host, 0);
    if (e == cudaSuccess) st->err_dev = (int*)p;
  }
  if (e != cudaSuccess) {
    set_error("state allocation failed: %s", cudaGetErrorString(e));
    cudaFree(st->base); cudaFree(st->parity);
    if (st->err_host) cudaFreeHost(st->err_host);
    delete st;
    return VV_ERR_CUDA;
  }
  *out = st;
  return VV_OK;
}
extern "C" int32_t vv_state_destroy(vv_state* st) {
  if (!st) return VV_OK;
  cudaSetDevice(st->device);
  cudaFree(st->base);
  cudaFree(st->parity);
  if (st->err_host) cudaFreeHost(st->err_host);
  delete st;
  return VV_OK;
}
extern "C" int32_t vv_state_n_slots(const vv_state* st) { return st ? st->n_slots : 0; }
extern "C" int64_t vv_state_bytes_per_slot(const vv_state* st) { return st ? st->slot_bytes : 0; }
extern "C" int32_t vv_state_check(vv_state* st) {
  if (!st) { set_error("null state"); return VV_ERR_INVALID_ARG; }
  return state_take_error(st);
}

extern "C" int32_t vv_state_zero(vv_state* st, const int32_t* slot_ids, int32_t n, void* cuda_stream) {
  if (!st) { set_error("null state"); return VV_ERR_INVALID_ARG; }
  if (n == 0) return VV_OK;
  if (n > 0 && !slot_ids) { set_error("slot_ids is null"); return VV_ERR_INVALID_ARG; }
  const int32_t prev = state_take_error(st);
  if (prev != VV_OK) return prev;
  VV_CUDA(cudaSetDevice(st->device));
  cudaStream_t s = reinterpret_cast<cudaStream_t>(cuda_stream);
  const int ny = n < 0 ? st->n_slots : n;
  dim3 grid(8, (unsigned)ny);
  state_zero_kernel<<<grid, 256, 0, s>>>(st->base, st->bank_bytes, st->slot_bytes, n < 0 ? nullptr : slot_ids, ny,
                                         st->parity, st->n_slots, st->err_dev);
  VV_CUDA(cudaGetLastError());
  return VV_OK;
}

// checkpoint helpers: select the state's device and wait for ITS work (all streams of that device: the caller does not
// pass the stream the decode ran on), then locate the slot's current bank
static int state_slot_ptr(const vv_state* st, int slot, uint8_t** p) {
  if (!st || slot < 0 || slot >= st->n_slots) { set_error("slot %d out of range", slot); return VV_ERR_SLOT; }
  int par = 0;
  VV_CUDA(cudaSetDevice(st->device));
  VV_CUDA(cudaDeviceSynchronize());
  VV_CUDA(cudaMemcpy(&par, st->parity + slot, sizeof(int), cudaMemcpyDeviceToHost));
  *p = st->base + (long long)(par & 1) * st->bank_bytes + (long long)slot * st->slot_bytes;
  return VV_OK;
}
extern "C" int32_t vv_state_export(const vv_state* st, int32_t slot, void* host_dst, int64_t bytes) {
  uint8_t* p;
  int rc = state_slot_ptr(st, slot, &p);
  if (rc != VV_OK) return rc;
  rc = state_take_error(const_cast<vv_state*>(st));
  if (rc != VV_OK) return rc;
  if (!host_dst || bytes != st->slot_bytes) { set_error("export needs %lld bytes", (long long)st->slot_bytes); return VV_ERR_INVALID_ARG; }
  VV_CUDA(cudaMemcpy(host_dst, p, (size_t)bytes, cudaMemcpyDeviceToHost));
  return VV_OK;
}
extern "C" int32_t vv_state_import(vv_state* st, int32_t slot, const void* host_src, int64_t bytes) {
  uint8_t* p;
  int rc = state_slot_ptr(st, slot, &p);
  if (rc != VV_OK) return rc;
  if (!host_src || bytes != st->slot_bytes) { set_error("import needs %lld bytes", (long long)st->slot_bytes); return VV_ERR_INVALID_ARG; }
  VV_CUDA(cudaMemcpy(p, host_src, (size_t)bytes, cudaMemcpyHostToDevice));
  return VV_OK;
}

// =====================================================================================================
// per-kernel entry points (offline form, no state) for the parity tests and bench.py's roofline leg
// =====================================================================================================
static int check_block(vv_decoder* d, int stage, int block) {
  if (!d || !d->finalized) { set_error("decoder not finalized"); return VV_ERR_NOT_FINALIZED; }
  if (stage < 0 || stage >= (int)d->stages.size() || block < 0 || block >= (int)d->stages[stage].blocks.size()) {
    set_error("stage %d block %d out of range", stage, block);
    return VV_ERR_INVALID_ARG;
  }
  return VV_OK;
}

extern "C" int32_t vv_kernel_mixer(vv_decoder* d, int32_t stage, int32_t block, const float* x, float* y, void* a_out,
                                   int32_t B, int32_t rows, void* cuda_stream) {
  int rc = check_block(d, stage, block);
  if (rc != VV_OK) return rc;
  if (!x || !y || !a_out || B < 1 || rows < 1) { set_error("bad argument"); return VV_ERR_INVALID_ARG; }
  VV_CUDA(cudaSetDevice(d->cfg.device));
  const StageW& sw = d->stages[stage];
  return launch_mixer(d, sw.blocks[block], sw.C, x, y, reinterpret_cast<__nv_bfloat16*>(a_out), B, rows, nullptr,
                      nullptr, reinterpret_cast<cudaStream_t>(cuda_stream));
}

extern "C" int32_t vv_kernel_ffn(vv_decoder* d, int32_t stage, int32_t block, const void* a, const float* resid,
                                 float* out, void* hidden, int32_t B, int32_t rows, void* cuda_stream) {
  int rc = check_block(d, stage, block);
  if (rc != VV_OK) return rc;
  if (!a || !resid || !out || !hidden || B < 1 || rows < 1) { set_error("bad argument"); return VV_ERR_INVALID_ARG; }
  VV_CUDA(cudaSetDevice(d->cfg.device));
  cudaStream_t s = reinterpret_cast<cudaStream_t>(cuda_stream);
  const BlockW& bw = d->stages[stage].blocks[block];
  const long long R = (long long)B * rows;
  if (ffn_fused_supported(d, d->stages[stage].C))
    return launch_ffn_fused(d, d->stages[stage].C, bw, reinterpret_cast<const __nv_bfloat16*>(a), resid, out, nullptr, R,
                            rows, s);
  EpiOut e1;
  e1.epi = EPI_GELU_BF16; e1.out_bf16 = reinterpret_cast<__nv_bfloat16*>(hidden);
  rc = launch_gemm(d, reinterpret_cast<const __nv_bfloat16*>(a), R, (int)R, 1, bw.lin1, e1, s);
  if (rc != VV_OK) return rc;
  EpiOut e2;
  e2.epi = EPI_RESID; e2.gamma = bw.ffn_gamma; e2.resid = resid; e2.out_f32 = out;
  return launch_gemm(d, reinterpret_cast<const __nv_bfloat16*>(hidden), R, (int)R, 1, bw.lin2, e2, s);
}

extern "C" int32_t vv_kernel_block(vv_decoder* d, int32_t stage, int32_t block, const float* x, float* out,
                                   int32_t B, int32_t rows, void* cuda_stream) {
  int rc = check_block(d, stage, block);
  if (rc != VV_OK) return rc;
  if (!x || !out || B < 1 || rows < 1) { set_error("bad argument"); return VV_ERR_INVALID_ARG; }
  VV_CUDA(cudaSetDevice(d->cfg.device));
  const StageW& sw = d->stages[stage];
  if (!block_fused_supported(d, sw.C)) { set_error("no fused block kernel for C=%d", sw.C); return VV_ERR_UNSUPPORTED; }
  return launch_block_fused(d, sw.C, sw.blocks[block], x, out, nullptr, B, rows, nullptr, nullptr,
                            reinterpret_cast<cudaStream_t>(cuda_stream));
}

extern "C" int32_t vv_kernel_upsample(vv_decoder* d, int32_t stage, void* xb, float* y, int32_t B, int32_t rows,
                                      void* cuda_stream) {
  if (!d || !d->finalized) { set_error("decoder not finalized"); return VV_ERR_NOT_FINALIZED; }
  if (stage < 1 || stage >= (int)d->stages.size()) { set_error("stage %d has no transposed conv", stage); return VV_ERR_INVALID_ARG; }
  if (!xb || !y || B < 1 || rows < 1) { set_error("bad argument"); return VV_ERR_INVALID_ARG; }
  VV_CUDA(cudaSetDevice(d->cfg.device));
  cudaStream_t s = reinterpret_cast<cudaStream_t>(cuda_stream);
  const StageW& sw = d->stages[stage];
  HaloArgs hp;
  hp.xb = reinterpret_cast<__nv_bfloat16*>(xb); hp.nbatch = B; hp.rows = rows; hp.C = sw.up.Ktap;
  hp.st_ref = make_state_ref(nullptr, nullptr, 0);
  launch_k(upsample_halo_kernel, dim3((unsigned)std::min((B * hp.C + 255) / 256, 148 * 8)), dim3(256), 0, s, hp);
  d->launches++;
  VV_LAUNCH_OK();
  EpiOut e;
  e.epi = EPI_BIAS_F32; e.out_f32 = y;
  return launch_gemm(d, reinterpret_cast<const __nv_bfloat16*>(xb), rows + 1, rows, B, sw.up, e, s);
}

extern "C" int32_t vv_kernel_head(vv_decoder* d, const float* x, void* out, int32_t out_dtype, int32_t B, int32_t rows,
                                  void* cuda_stream) {
  if (!d || !d->finalized) { set_error("decoder not finalized"); return VV_ERR_NOT_FINALIZED; }
  if (!x || !out || B < 1 || rows < 1) { set_error("bad argument"); return VV_ERR_INVALID_ARG; }
  VV_CUDA(cudaSetDevice(d->cfg.device));
  HeadArgs hp;
  hp.x = x; hp.w = d->head_w; hp.bias = d->head_bias; hp.out = out; hp.out_is_bf16 = (out_dtype == VV_BF16);
  hp.out_batch_stride = rows; hp.nbatch = B; hp.rows = rows;
  hp.strip = pick_strip((long long)B * rows, rows);
  hp.strips_per_batch = (rows + hp.strip - 1) / hp.strip;
  hp.st = make_state_ref(nullptr, nullptr, 0);
  const long long groups = (long long)B * hp.strips_per_batch;
  launch_k(head_kernel, dim3((unsigned)((groups * 8 + 127) / 128)), dim3(128), 0, reinterpret_cast<cudaStream_t>(cuda_stream), hp);
  d->launches++;
  VV_LAUNCH_OK();
  return VV_OK;
}

// GELU as the FFN epilogue evaluates it, exposed for the accuracy test (device pointers, n floats).
namespace { __global__ void gelu_test_kernel(const float* x, float* y, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) y[i] = vv::gelu_erf(x[i]);
} }
extern "C" int32_t vv_test_gelu(const float* x, float* y, int32_t n, void* cuda_stream) {
  if (!x || !y || n < 1) { set_error("bad argument"); return VV_ERR_INVALID_ARG; }
  gelu_test_kernel<<<(n + 255) / 256, 256, 0, reinterpret_cast<cudaStream_t>(cuda_stream)>>>(x, y, n);
  VV_LAUNCH_OK();
  return VV_OK;
}

// PCM16 output step (reference convert_to_16_bit_wav, main.py:2956-2979): see pcm16.cuh.
extern "C" int32_t vv_pcm16(const void* wave, int32_t wave_dtype, int32_t B, int64_t L, int64_t row_stride, int16_t* pcm,
                            float* peak, void* cuda_stream) {
  if (!wave || !pcm || B < 1 || L < 1 || row_stride < L) { set_error("bad argument"); return VV_ERR_INVALID_ARG; }
  if (wave_dtype != VV_F32 && wave_dtype != VV_BF16) { set_error("wave must be f32 or bf16"); return VV_ERR_BAD_DTYPE; }
  if (B > 65535) { set_error("at most 65535 rows per call"); return VV_ERR_BAD_SHAPE; }
  cudaStream_t s = reinterpret_cast<cudaStream_t>(cuda_stream);
  Pcm16Args p;
  p.wave = wave; p.wave_is_bf16 = (wave_dtype == VV_BF16); p.B = B; p.L = L; p.row_stride = row_stride;
  p.pcm = pcm; p.peak = peak;
  p.vec_ok = ((reinterpret_cast<uintptr_t>(wave) | reinterpret_cast<uintptr_t>(pcm)) % 16 == 0) && row_stride % 8 == 0 &&
             (B == 1 || L % 8 == 0);
  if (L <= PCM_ROW_MAX) {
    pcm16_row_kernel<<<(unsigned)B, PCM_THREADS, 0, s>>>(p);
  } else {
    if (!peak) { set_error("rows longer than %d samples need the peak buffer", PCM_ROW_MAX); return VV_ERR_INVALID_ARG; }
    const long long chunks = (L + PCM_CHUNK - 1) / PCM_CHUNK;
    if (chunks > 0x7fffffffLL) { set_error("row too long"); return VV_ERR_BAD_SHAPE; }
    VV_CUDA(cudaMemsetAsync(peak, 0, sizeof(float) * (size_t)B, s));
    pcm16_peak_kernel<<<dim3((unsigned)chunks, (unsigned)B), PCM_THREADS, 0, s>>>(p);
    pcm16_convert_kernel<<<dim3((unsigned)chunks, (unsigned)B), PCM_THREADS, 0, s>>>(p);
  }
  VV_LAUNCH_OK();
  return VV_OK;
}

// ---- vv_pcm16_static: vv_pcm16 for callers that cannot allocate --------------------------------------------------------
// No counterpart in the reference.  Same arithmetic and same kernels as
// vv_pcm16; the only difference is where the per-row peak scratch lives: in this file-scope device array instead of a
// caller-provided buffer, so the call needs no allocation of any kind (no cudaMalloc / cudaFree, no allocating library
// routine): usable under CUDA graph capture, in serving loops and in harnesses that forbid allocation.
//   * every GPU that loads this library holds the array (VV_PCM16_STATIC_MAX_ROWS floats = 16 KB) until the process exits;
//   * ALL calls on a device share it and it keeps the last call's contents: the two-pass form clears the rows it uses
//     (the peak pass is an atomicMax and must start at zero) with cudaMemsetAsync on the caller's stream;
//   * therefore two calls that overlap on the GPU (different streams, no ordering between them) RACE on it.
// In host code `g_pcm16_static_peak` is a host-side stand-in for the symbol, NOT a device address: the address a kernel
// can use comes from cudaGetSymbolAddress, resolved once per device with that device current (below) and cached.
static __device__ __align__(256) float g_pcm16_static_peak[VV_PCM16_STATIC_MAX_ROWS];

namespace {
constexpr int kMaxDevices = 64;
float* g_pcm16_static_ptr[kMaxDevices] = {};
std::mutex g_pcm16_static_mu;

// device address of the scratch array on the CURRENT device (set-up: the first call on a device resolves it)
int pcm16_static_scratch(float** out) {
  int dev = -1;
  VV_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) { set_error("device ordinal %d outside [0, %d)", dev, kMaxDevices); return VV_ERR_NO_DEVICE; }
  std::lock_guard<std::mutex> lk(g_pcm16_static_mu);
  if (g_pcm16_static_ptr[dev] == nullptr) {
    void* ptr = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&ptr, g_pcm16_static_peak);
    if (e != cudaSuccess || ptr == nullptr) {
      set_error("cudaGetSymbolAddress(pcm16 scratch) failed on device %d: %s", dev, cudaGetErrorString(e));
      (void)cudaGetLastError();
      return VV_ERR_CUDA;
    }
    g_pcm16_static_ptr[dev] = reinterpret_cast<float*>(ptr);
  }
  *out = g_pcm16_static_ptr[dev];
  return VV_OK;
}
}  // namespace

extern "C" int32_t vv_pcm16_static(const void* wave, int32_t wave_dtype, int32_t B, int64_t L, int64_t row_stride,
                                   int16_t* pcm, void* cuda_stream) {
  if (!wave || !pcm || B < 1 || L < 1 || row_stride < L) { set_error("bad argument"); return VV_ERR_INVALID_ARG; }
  if (wave_dtype != VV_F32 && wave_dtype != VV_BF16) { set_error("wave must be f32 or bf16"); return VV_ERR_BAD_DTYPE; }
  if (B > VV_PCM16_STATIC_MAX_ROWS) {
    set_error("vv_pcm16_static: %d rows exceed the static scratch of %d rows per call (split the batch, or use vv_pcm16 "
              "with a peak buffer of your own)", B, VV_PCM16_STATIC_MAX_ROWS);
    return VV_ERR_BAD_SHAPE;
  }
  float* peak = nullptr;
  const int rc = pcm16_static_scratch(&peak);
  if (rc != VV_OK) return rc;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(cuda_stream);
  Pcm16Args p;
  p.wave = wave; p.wave_is_bf16 = (wave_dtype == VV_BF16); p.B = B; p.L = L; p.row_stride = row_stride;
  p.pcm = pcm; p.peak = peak;
  p.vec_ok = ((reinterpret_cast<uintptr_t>(wave) | reinterpret_cast<uintptr_t>(pcm)) % 16 == 0) && row_stride % 8 == 0 &&
             (B == 1 || L % 8 == 0);
  if (L <= PCM_ROW_MAX) {
    pcm16_row_kernel<<<(unsigned)B, PCM_THREADS, 0, s>>>(p);            // writes peak[b], reads nothing from the scratch
  } else {
    const long long chunks = (L + PCM_CHUNK - 1) / PCM_CHUNK;
    if (chunks > 0x7fffffffLL) { set_error("row too long"); return VV_ERR_BAD_SHAPE; }
    VV_CUDA(cudaMemsetAsync(peak, 0, sizeof(float) * (size_t)B, s));    // the previous call's peaks are still in there
    pcm16_peak_kernel<<<dim3((unsigned)chunks, (unsigned)B), PCM_THREADS, 0, s>>>(p);
    pcm16_convert_kernel<<<dim3((unsigned)chunks, (unsigned)B), PCM_THREADS, 0, s>>>(p);
  }
  VV_CUDA(cudaGetLastError());
  return VV_OK;
}

#ifdef VV_FS_TRACE
extern "C" int32_t vv_debug_read_trace(long long* host, int32_t n) {
  return (int32_t)cudaMemcpyFromSymbol(host, vv::vv_fs_trace, sizeof(long long) * (size_t)n);
}
#endif
#ifdef VV_BW_TRACE
extern "C" int32_t vv_debug_read_bw_trace(long long* host, int32_t n) {
  return (int32_t)cudaMemcpyFromSymbol(host, vv::vv_bw_trace, sizeof(long long) * (size_t)n);
}
#endif

// =====================================================================================================
// encoder path (TokenizerEncoder, :687-813)
// =====================================================================================================
#include "vvenc.inl"

// =====================================================================================================
// diffusion-head sampler
// =====================================================================================================
#include "vvsamp.inl"
