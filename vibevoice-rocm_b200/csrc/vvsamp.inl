// vvsamp.inl — host side of the diffusion-head SAMPLER behind the C ABI (included at the end of vvdec.cu:
// it reuses that file's weight packer and GEMM launcher).  Replaces, for one frame step of n streams,
//   VibeVoiceForConditionalGenerationInference.sample_speech_tokens   modeling_vibevoice_inference.py:695-708
// = ddpm_num_inference_steps x [ VibeVoiceDiffusionHead.forward on 2n rows (modular_vibevoice_diffusion_head.py:254-280),
//   classifier-free guidance (:703-705), DPMSolverMultistepScheduler.step (vibevoice/schedule/dpm_solver.py:935-1022) ].
// Step-invariant work is hoisted: cond_proj(condition) once per call; the timestep embeddings (TimestepEmbedder :48-93) and
// the solver's scalar coefficients once per step COUNT, on the host in fp32 (they do not depend on data).

struct vv_sampler {
  vv_decoder core;                       // device, SM count, options, launch / allocation bookkeeping (no decoder geometry)
  vv_sampler_config scfg;
  int NM = 0;                            // columns of the concatenated adaLN output: (3 L + 2) H
  vv::DenseW w_noisy, w_cond, w_ada, w_final;
  std::vector<vv::DenseW> w_gu, w_down;  // per layer: [gate_proj ; up_proj] (N = 2F) and down_proj (K = F)
  std::vector<float*> norm_w;            // per layer [H]
  std::vector<float> h_t0, h_t2;         // host fp32 copies of t_embedder.mlp.{0,2}.weight
  struct StepPlan {
    std::vector<float> coef;             // [steps][5] = alpha_s, sigma_s, A, B, D
    std::vector<long long> timesteps;
    float* temb = nullptr;               // device [steps, H]
  };
  std::map<int, StepPlan> plans;
  float* temb_probe = nullptr;           // device [H]: vv_sampler_head's embedding of its one timestep
};

namespace {

// ---- scheduler tables (dpm_solver.py:28-85, :252-264, :321-424, :483-487, :627-695, :745-768) ------------------------
void samp_sigmas(int train_steps, int n, std::vector<long long>& ts, std::vector<float>& sig) {
  auto abar = [](double t) { const double c = std::cos((t + 0.008) / 1.008 * M_PI / 2); return c * c; };
  std::vector<float> all((size_t)train_steps);
  float ac = 1.0f;                                          // float32 running product, like torch.cumprod on float32
  for (int i = 0; i < train_steps; ++i) {
    const float beta = (float)std::min(1.0 - abar((i + 1.0) / train_steps) / abar((double)i / train_steps), 0.999);
    ac *= (1.0f - beta);
    all[(size_t)i] = std::sqrt((1.0f - ac) / ac);
  }
  ts.resize((size_t)n);
  sig.resize((size_t)n + 1);
  for (int k = 0; k < n; ++k) {                             // np.linspace(0, T - 1, n + 1).round()[::-1][:-1]
    const double v = (double)(train_steps - 1) * (double)(n - k) / (double)n;
    ts[(size_t)k] = (long long)std::nearbyint(v);           // round-half-even, like numpy
    sig[(size_t)k] = all[(size_t)ts[(size_t)k]];            // np.interp at integer positions = the table entry
  }
  sig[(size_t)n] = 0.0f;                                    // final_sigmas_type == "zero"
}

void samp_coefficients(const std::vector<float>& sig, std::vector<float>& coef) {
  const int n = (int)sig.size() - 1;
  coef.assign((size_t)n * 5, 0.f);
  auto as = [](float s, float& a, float& st) { a = 1.0f / std::sqrt(s * s + 1.0f); st = s * a; };
  for (int i = 0; i < n; ++i) {
    float a_s, s_s, a_t, s_t;
    as(sig[(size_t)i], a_s, s_s);
    as(sig[(size_t)i + 1], a_t, s_t);
    const float lam_t = std::log(a_t) - std::log(s_t);      // +inf at the last step
    const float lam_s = std::log(a_s) - std::log(s_s);
    const float h = lam_t - lam_s;
    const float em = std::exp(-h) - 1.0f;
    float D = 0.f;
    if (!(i == 0 || i == n - 1)) {                          // second order (midpoint) between the two first-order ends
      float a_p, s_p;
      as(sig[(size_t)i - 1], a_p, s_p);
      const float r0 = (lam_s - (std::log(a_p) - std::log(s_p))) / h;
      D = -0.5f * (a_t * em) / r0;
    }
    float* c = &coef[(size_t)i * 5];
    c[0] = a_s; c[1] = s_s; c[2] = s_t / s_s; c[3] = -(a_t * em); c[4] = D;
  }
}

// TimestepEmbedder.forward (:90-93) for one timestep, host fp32: Linear(256 -> H) -> SiLU -> Linear(H -> H), no biases
void samp_t_embed(const vv_sampler* sp, double t, float* out) {
  const int H = sp->scfg.hidden, Fq = 256, half = Fq / 2;
  std::vector<float> e((size_t)Fq), mid((size_t)H);
  for (int k = 0; k < half; ++k) {
    const float f = std::exp(-std::log(10000.0f) * (float)k / (float)half);
    const float arg = (float)t * f;
    e[(size_t)k] = std::cos(arg);
    e[(size_t)half + k] = std::sin(arg);
  }
  for (int h = 0; h < H; ++h) {
    double acc = 0.0;
    const float* w = &sp->h_t0[(size_t)h * Fq];
    for (int k = 0; k < Fq; ++k) acc += (double)w[k] * e[(size_t)k];
    const float a = (float)acc;
    mid[(size_t)h] = a / (1.0f + std::exp(-a));
  }
  for (int h = 0; h < H; ++h) {
    double acc = 0.0;
    const float* w = &sp->h_t2[(size_t)h * H];
    for (int k = 0; k < H; ++k) acc += (double)w[k] * mid[(size_t)k];
    out[h] = (float)acc;
  }
}

struct SampBufs { long long cond, cproj, sc, mod, xin, x, a, gu, hsw, f, v, xlat, mprev, total; };

// K slices of the down_proj GEMM (N = hidden, K = ffn: few output tiles, long K): as many as keep every slice-tile on its
// own SM, at most 4, dividing the k-block count.  A function of the row count only, so a captured graph and an eager
// call of the same shape take the same summation order.
constexpr int SAMP_MAX_KS = 4;
int samp_down_slices(const vv_sampler* sp, long long R) {
  const long long tiles = ((R + 127) / 128) * (sp->scfg.hidden / 128);
  const int nk = sp->scfg.ffn / 64;
  int ks = 1;
  for (int k = 2; k <= SAMP_MAX_KS; ++k)
    if (nk % k == 0 && tiles * k <= sp->core.num_sms && sp->scfg.hidden % 128 == 0) ks = k;
  return ks;
}

SampBufs samp_bufs(const vv_sampler* sp, long long R) {
  const long long H = sp->scfg.hidden, F = sp->scfg.ffn, Ld = sp->scfg.latent, NM = sp->NM;
  SampBufs b;
  long long off = 0;
  auto take = [&](long long bytes) { const long long o = off; off += align_up(bytes, 1024); return o; };
  b.cond = take(R * H * 2); b.cproj = take(R * H * 4); b.sc = take(R * H * 2); b.mod = take(R * NM * 4);
  b.xin = take(R * Ld * 2); b.x = take(R * H * 4); b.a = take(R * H * 2); b.gu = take(1024);   // (the [gate | up] tensor no longer exists: SwiGLU runs in the GEMM epilogue)
  b.hsw = take(R * F * 2); b.f = take(SAMP_MAX_KS * R * H * 4); b.v = take(R * Ld * 4); b.xlat = take(R * Ld * 4); b.mprev = take(R * Ld * 4);
  b.total = off + 1024;
  return b;
}

__global__ void samp_pack_cond_kernel(const void* cond, const void* neg, int is_bf16, __nv_bfloat16* out, int n, int H) {
  const long long half = (long long)n * H, total = 2 * half;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const void* src = i < half ? cond : neg;
    const long long j = i < half ? i : i - half;
    out[i] = is_bf16 ? reinterpret_cast<const __nv_bfloat16*>(src)[j] : __float2bfloat16(reinterpret_cast<const float*>(src)[j]);
  }
}

unsigned samp_grid(long long work, int block) { return (unsigned)std::max<long long>(1, std::min<long long>((work + block - 1) / block, 148LL * 8)); }

// One head forward on R rows whose bf16 inputs (xin [R, latent]) and modulation input (sc [R, H]) are in the workspace.
int samp_head_pass(vv_sampler* sp, uint8_t* ws, const SampBufs& b, int R, cudaStream_t s) {
  vv_decoder* d = &sp->core;
  const int H = sp->scfg.hidden, L = sp->scfg.layers;
  __nv_bfloat16* SC = reinterpret_cast<__nv_bfloat16*>(ws + b.sc);
  float* MOD = reinterpret_cast<float*>(ws + b.mod);
  __nv_bfloat16* XIN = reinterpret_cast<__nv_bfloat16*>(ws + b.xin);
  float* X = reinterpret_cast<float*>(ws + b.x);
  __nv_bfloat16* A = reinterpret_cast<__nv_bfloat16*>(ws + b.a);
  __nv_bfloat16* HS = reinterpret_cast<__nv_bfloat16*>(ws + b.hsw);
  float* Fo = reinterpret_cast<float*>(ws + b.f);
  float* V = reinterpret_cast<float*>(ws + b.v);
  const int ks = samp_down_slices(sp, R);
  int rc;
  EpiOut e;
  e.epi = EPI_BIAS_F32;
  e.out_f32 = MOD;                                           // every adaLN output of the step in one GEMM
  if ((rc = launch_gemm(d, SC, R, R, 1, sp->w_ada, e, s)) != VV_OK) return rc;
  e.out_f32 = X;                                             // noisy_images_proj (:272)
  if ((rc = launch_gemm(d, XIN, R, R, 1, sp->w_noisy, e, s)) != VV_OK) return rc;
  for (int l = 0; l <= L; ++l) {
    SampNormArgs na;
    na.x = X; na.f = l > 0 ? Fo : nullptr; na.mod = MOD; na.NM = sp->NM; na.H = H; na.eps = sp->scfg.eps; na.a = A;
    na.gate_off = l > 0 ? (l - 1) * 3 * H + 2 * H : 0;       // chunk(3): shift | scale | gate (:159)
    na.shift_off = l * 3 * H; na.scale_off = l * 3 * H + H;  // final layer (l == L): chunk(2) at 3 L H (:187)
    na.w = l < L ? sp->norm_w[(size_t)l] : nullptr;
    na.f_planes = ks; na.f_plane_stride = (long long)R * H;
    launch_k(samp_norm_mod_kernel, dim3((unsigned)R), dim3(SAMP_NORM_THREADS), 0, s, na);
    d->launches++;
    VV_LAUNCH_OK();
    if (l == L) break;
    EpiOut eg;                                                 // [gate; up] GEMM with the SwiGLU in its epilogue: h = bf16
    eg.epi = EPI_SWIGLU_BF16; eg.out_bf16 = HS;
    if ((rc = launch_gemm(d, A, R, R, 1, sp->w_gu[(size_t)l], eg, s)) != VV_OK) return rc;
    e.out_f32 = Fo;                                            // ks planes of partial sums; the next norm kernel adds them
    e.k_slices = ks;
    if ((rc = launch_gemm(d, HS, R, R, 1, sp->w_down[(size_t)l], e, s)) != VV_OK) return rc;
    e.k_slices = 1;
  }
  e.out_f32 = V;                                             // FinalLayer.linear (:189)
  return launch_gemm(d, A, R, R, 1, sp->w_final, e, s);
}

}  // namespace

// =====================================================================================================
// C ABI
// =====================================================================================================
extern "C" int32_t vv_sampler_create(const vv_sampler_config* cfg, vv_sampler** out) {
  if (cfg == nullptr || out == nullptr) { set_error("null argument"); return VV_ERR_INVALID_ARG; }
  *out = nullptr;
  if (cfg->hidden < 128 || cfg->hidden % 128 != 0 || cfg->hidden > 4096 || cfg->ffn < 128 || cfg->ffn % 128 != 0 ||
      cfg->latent != 64 || cfg->layers < 1 || cfg->layers > 16 || cfg->train_steps < 2) {
    set_error("unsupported diffusion head: hidden=%d ffn=%d latent=%d layers=%d (kernels take hidden %% 128 == 0 <= 4096, "
              "ffn %% 128 == 0, latent == 64)", cfg->hidden, cfg->ffn, cfg->latent, cfg->layers);
    return VV_ERR_UNSUPPORTED;
  }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || cfg->device < 0 || cfg->device >= ndev) {
    set_error("CUDA device %d not available (%d devices)", cfg->device, ndev);
    return VV_ERR_NO_DEVICE;
  }
  cudaDeviceProp prop;
  VV_CUDA(cudaGetDeviceProperties(&prop, cfg->device));
  if (prop.major != 10) {
    set_error("device %d is sm_%d%d; this library only contains sm_100a code", cfg->device, prop.major, prop.minor);
    return VV_ERR_NO_DEVICE;
  }
  vv_sampler* sp = new vv_sampler();
  sp->scfg = *cfg;
  memset(&sp->core.cfg, 0, sizeof(sp->core.cfg));
  sp->core.cfg.device = cfg->device;
  sp->core.cfg.gemm_path = 0;
  sp->core.num_sms = prop.multiProcessorCount;
  sp->NM = (3 * cfg->layers + 2) * cfg->hidden;
  *out = sp;
  return VV_OK;
}

extern "C" int32_t vv_sampler_destroy(vv_sampler* sp) {
  if (sp == nullptr) return VV_OK;
  cudaSetDevice(sp->scfg.device);
  for (void* p : sp->core.allocs) cudaFree(p);
  for (auto& kv : sp->plans) cudaFree(kv.second.temb);
  cudaFree(sp->temb_probe);
  delete sp;
  return VV_OK;
}

extern "C" int32_t vv_sampler_load_tensor(vv_sampler* sp, const char* name, const void* data, int32_t dtype,
                                          const int64_t* shape, int32_t ndim, int32_t is_device) {
  return stage_tensor(sp ? &sp->core : nullptr, "", name, data, dtype, shape, ndim, is_device);
}

extern "C" int32_t vv_sampler_finalize(vv_sampler* sp) {
  if (!sp) { set_error("null sampler"); return VV_ERR_INVALID_ARG; }
  vv_decoder* d = &sp->core;
  if (d->finalized) return VV_OK;
  VV_CUDA(cudaSetDevice(sp->scfg.device));
  for (void* p : d->allocs) cudaFree(p);
  d->allocs.clear();
  d->param_count = 0;
  sp->w_gu.clear(); sp->w_down.clear(); sp->norm_w.clear();
  const int H = sp->scfg.hidden, F = sp->scfg.ffn, Ld = sp->scfg.latent, L = sp->scfg.layers;
  Packer pk{d};
  auto zeros = [](int n) { return std::vector<float>((size_t)n, 0.f); };      // every Linear of the head is bias=False
  if (const auto* w = pk.get("noisy_images_proj.weight", {H, Ld})) pk.dense(sp->w_noisy, to_bf16(*w), zeros(H), H, Ld, 1);
  if (const auto* w = pk.get("cond_proj.weight", {H, H})) pk.dense(sp->w_cond, to_bf16(*w), zeros(H), H, H, 1);
  if (const auto* w = pk.get("t_embedder.mlp.0.weight", {H, 256})) sp->h_t0 = *w;
  if (const auto* w = pk.get("t_embedder.mlp.2.weight", {H, H})) sp->h_t2 = *w;
  std::vector<float> ada((size_t)sp->NM * H);                                  // [adaLN_0 | ... | adaLN_{L-1} | adaLN_final] rows
  for (int l = 0; l < L && pk.status == VV_OK; ++l) {
    const std::string p = "layers." + std::to_string(l) + ".";
    const auto* g = pk.get(p + "ffn.gate_proj.weight", {F, H});
    const auto* u = pk.get(p + "ffn.up_proj.weight", {F, H});
    const auto* dn = pk.get(p + "ffn.down_proj.weight", {H, F});
    const auto* nw = pk.get(p + "norm.weight", {H});
    const auto* ad = pk.get(p + "adaLN_modulation.1.weight", {3 * H, H});
    if (pk.status != VV_OK) break;
    // rows interleaved in blocks of 128 hidden units: [gate 0..127 | up 0..127 | gate 128..255 | up 128..255 | ...], so a
    // 256-wide GEMM tile holds gate and up of the same units and its epilogue applies SwiGLU (EPI_SWIGLU_BF16)
    std::vector<float> gu((size_t)2 * F * H);
    for (int t = 0; t < F / 128; ++t) {
      memcpy(gu.data() + (size_t)(2 * t) * 128 * H, g->data() + (size_t)t * 128 * H, sizeof(float) * (size_t)128 * H);
      memcpy(gu.data() + (size_t)(2 * t + 1) * 128 * H, u->data() + (size_t)t * 128 * H, sizeof(float) * (size_t)128 * H);
    }
    vv::DenseW wgu, wdn;
    pk.dense(wgu, to_bf16(gu), zeros(2 * F), 2 * F, H, 1);
    pk.dense(wdn, to_bf16(*dn), zeros(H), H, F, 1);
    sp->w_gu.push_back(wgu);
    sp->w_down.push_back(wdn);
    sp->norm_w.push_back(pk.upload(*nw));
    memcpy(ada.data() + (size_t)l * 3 * H * H, ad->data(), sizeof(float) * (size_t)3 * H * H);
  }
  if (pk.status == VV_OK) {
    const auto* ad = pk.get("final_layer.adaLN_modulation.1.weight", {2 * H, H});
    const auto* fl = pk.get("final_layer.linear.weight", {Ld, H});
    if (ad && fl) {
      memcpy(ada.data() + (size_t)L * 3 * H * H, ad->data(), sizeof(float) * (size_t)2 * H * H);
      pk.dense(sp->w_ada, to_bf16(ada), zeros(sp->NM), sp->NM, H, 1);
      pk.dense(sp->w_final, to_bf16(*fl), zeros(Ld), Ld, H, 1);
    }
  }
  if (pk.status == VV_OK) {
    void* p = nullptr;
    if (cudaMalloc(&p, sizeof(float) * (size_t)H) != cudaSuccess) { set_error("allocation failed"); return VV_ERR_CUDA; }
    cudaFree(sp->temb_probe);
    sp->temb_probe = (float*)p;
  }
  if (pk.status != VV_OK) return pk.status;
  d->staged.clear();
  d->staged_shape.clear();
  d->finalized = true;
  return VV_OK;
}

extern "C" int64_t vv_sampler_param_count(const vv_sampler* sp) { return sp ? sp->core.param_count : 0; }
extern "C" int64_t vv_sampler_take_launch_count(vv_sampler* sp) { return sp ? vv_decoder_take_launch_count(&sp->core) : 0; }
extern "C" int64_t vv_sampler_workspace_bytes(const vv_sampler* sp, int32_t n) {
  if (!sp || !sp->core.finalized || n < 1) return 0;
  return samp_bufs(sp, 2LL * n).total;
}

// Tables for a step count: solver coefficients + the timestep embeddings of its timesteps (host fp32 -> device).  Allocates
// once per distinct `steps`; vv_sampler_sample calls it lazily (call it yourself before capturing a CUDA graph).
extern "C" int32_t vv_sampler_prepare(vv_sampler* sp, int32_t steps) {
  if (!sp || !sp->core.finalized) { set_error("sampler not finalized"); return VV_ERR_NOT_FINALIZED; }
  if (steps < 1 || steps > sp->scfg.train_steps) { set_error("steps=%d", steps); return VV_ERR_INVALID_ARG; }
  if (sp->plans.count(steps)) return VV_OK;
  VV_CUDA(cudaSetDevice(sp->scfg.device));
  vv_sampler::StepPlan plan;
  std::vector<float> sig;
  samp_sigmas(sp->scfg.train_steps, steps, plan.timesteps, sig);
  samp_coefficients(sig, plan.coef);
  const int H = sp->scfg.hidden;
  std::vector<float> temb((size_t)steps * H);
  for (int i = 0; i < steps; ++i) samp_t_embed(sp, (double)plan.timesteps[(size_t)i], &temb[(size_t)i * H]);
  void* p = nullptr;
  VV_CUDA(cudaMalloc(&p, temb.size() * sizeof(float)));
  VV_CUDA(cudaMemcpy(p, temb.data(), temb.size() * sizeof(float), cudaMemcpyHostToDevice));
  plan.temb = (float*)p;
  sp->plans[steps] = std::move(plan);
  return VV_OK;
}

extern "C" int32_t vv_sampler_sample(vv_sampler* sp, const void* condition, const void* neg_condition, int32_t cond_dtype,
                                     int32_t n, const float* noise, float cfg_scale, int32_t steps, float* out,
                                     void* workspace, int64_t workspace_bytes, void* cuda_stream) {
  if (!sp || !condition || !neg_condition || !noise || !out || !workspace) { set_error("null argument"); return VV_ERR_INVALID_ARG; }
  if (!sp->core.finalized) { set_error("vv_sampler_finalize has not succeeded"); return VV_ERR_NOT_FINALIZED; }
  if (n < 1 || 2LL * n >= (1 << 24)) { set_error("n=%d", n); return VV_ERR_BAD_SHAPE; }
  if (cond_dtype != VV_F32 && cond_dtype != VV_BF16) { set_error("condition dtype must be f32 or bf16"); return VV_ERR_BAD_DTYPE; }
  if (((uintptr_t)workspace & 1023) != 0) { set_error("workspace must be 1 KB aligned"); return VV_ERR_INVALID_ARG; }
  const int R = 2 * n;
  const SampBufs b = samp_bufs(sp, R);
  if (b.total > workspace_bytes) { set_error("workspace of %lld bytes, %lld needed", (long long)workspace_bytes, b.total); return VV_ERR_WORKSPACE; }
  int rc = vv_sampler_prepare(sp, steps);
  if (rc != VV_OK) return rc;
  const vv_sampler::StepPlan& plan = sp->plans[steps];
  VV_CUDA(cudaSetDevice(sp->scfg.device));
  cudaStream_t s = reinterpret_cast<cudaStream_t>(cuda_stream);
  vv_decoder* d = &sp->core;
  uint8_t* ws = reinterpret_cast<uint8_t*>(workspace);
  const int H = sp->scfg.hidden, Ld = sp->scfg.latent;
  __nv_bfloat16* COND = reinterpret_cast<__nv_bfloat16*>(ws + b.cond);
  float* CPROJ = reinterpret_cast<float*>(ws + b.cproj);
  float* XLAT = reinterpret_cast<float*>(ws + b.xlat);
  float* MPREV = reinterpret_cast<float*>(ws + b.mprev);
  PdlScope pdl(true);                                        // every launch here is short: programmatic dependent launch on
  launch_k(samp_pack_cond_kernel, dim3(samp_grid((long long)R * H, 256)), dim3(256), 0, s, condition, neg_condition,
           (int)(cond_dtype == VV_BF16), COND, (int)n, H);
  d->launches++;
  VV_LAUNCH_OK();
  EpiOut e;
  e.epi = EPI_BIAS_F32; e.out_f32 = CPROJ;                   // cond_proj (:274): step-invariant, once per call
  if ((rc = launch_gemm(d, COND, R, R, 1, sp->w_cond, e, s)) != VV_OK) return rc;
  VV_CUDA(cudaMemcpyAsync(XLAT, noise, sizeof(float) * (size_t)n * Ld, cudaMemcpyDeviceToDevice, s));   // speech = randn (:698)
  for (int i = 0; i < steps; ++i) {
    launch_k(samp_silu_add_kernel, dim3(samp_grid((long long)R * H / 4, 256)), dim3(256), 0, s, (const float*)CPROJ,
             (const float*)(plan.temb + (size_t)i * H), reinterpret_cast<__nv_bfloat16*>(ws + b.sc), R, H);
    launch_k(samp_dup_latent_kernel, dim3(samp_grid((long long)R * Ld, 256)), dim3(256), 0, s, (const float*)XLAT,
             reinterpret_cast<__nv_bfloat16*>(ws + b.xin), (int)n, R, Ld);
    d->launches += 2;
    VV_LAUNCH_OK();
    if ((rc = samp_head_pass(sp, ws, b, R, s)) != VV_OK) return rc;
    SampSolverArgs sa;
    sa.v = reinterpret_cast<const float*>(ws + b.v); sa.x = XLAT; sa.m_prev = MPREV; sa.n = n; sa.Ld = Ld; sa.cfg_scale = cfg_scale;
    const float* c = &plan.coef[(size_t)i * 5];
    sa.alpha_s = c[0]; sa.sigma_s = c[1]; sa.A = c[2]; sa.B = c[3]; sa.D = c[4];
    launch_k(samp_solver_kernel, dim3(samp_grid((long long)n * Ld, 256)), dim3(256), 0, s, sa);
    d->launches++;
    VV_LAUNCH_OK();
  }
  VV_CUDA(cudaMemcpyAsync(out, XLAT, sizeof(float) * (size_t)n * Ld, cudaMemcpyDeviceToDevice, s));
  return VV_OK;
}

// Per-kernel hook (tests): ONE head forward, VibeVoiceDiffusionHead.forward(noisy, t, condition) (:254-280) on R rows.
extern "C" int32_t vv_sampler_head(vv_sampler* sp, const float* noisy, float timestep, const void* condition,
                                   int32_t cond_dtype, int32_t R, float* v_out, void* workspace, int64_t workspace_bytes,
                                   void* cuda_stream) {
  if (!sp || !noisy || !condition || !v_out || !workspace) { set_error("null argument"); return VV_ERR_INVALID_ARG; }
  if (!sp->core.finalized) { set_error("vv_sampler_finalize has not succeeded"); return VV_ERR_NOT_FINALIZED; }
  if (R < 2 || R % 2 != 0) { set_error("R=%d (an even row count: the workspace is laid out for 2n rows)", R); return VV_ERR_BAD_SHAPE; }
  if (cond_dtype != VV_F32 && cond_dtype != VV_BF16) { set_error("condition dtype must be f32 or bf16"); return VV_ERR_BAD_DTYPE; }
  const SampBufs b = samp_bufs(sp, R);
  if (b.total > workspace_bytes || ((uintptr_t)workspace & 1023) != 0) { set_error("workspace too small or unaligned"); return VV_ERR_WORKSPACE; }
  VV_CUDA(cudaSetDevice(sp->scfg.device));
  cudaStream_t s = reinterpret_cast<cudaStream_t>(cuda_stream);
  vv_decoder* d = &sp->core;
  uint8_t* ws = reinterpret_cast<uint8_t*>(workspace);
  const int H = sp->scfg.hidden, Ld = sp->scfg.latent;
  std::vector<float> temb((size_t)H);
  samp_t_embed(sp, (double)timestep, temb.data());
  VV_CUDA(cudaStreamSynchronize(s));                          // test hook: the pageable copy below must not race earlier work
  VV_CUDA(cudaMemcpy(sp->temb_probe, temb.data(), sizeof(float) * (size_t)H, cudaMemcpyHostToDevice));
  __nv_bfloat16* COND = reinterpret_cast<__nv_bfloat16*>(ws + b.cond);
  const size_t esz = cond_dtype == VV_BF16 ? 2 : 4;
  launch_k(samp_pack_cond_kernel, dim3(samp_grid((long long)R * H, 256)), dim3(256), 0, s, condition,
           (const void*)(reinterpret_cast<const uint8_t*>(condition) + (size_t)(R / 2) * H * esz), (int)(cond_dtype == VV_BF16),
           COND, (int)(R / 2), H);
  d->launches++;
  VV_LAUNCH_OK();
  EpiOut e;
  e.epi = EPI_BIAS_F32; e.out_f32 = reinterpret_cast<float*>(ws + b.cproj);
  int rc = launch_gemm(d, COND, R, R, 1, sp->w_cond, e, s);
  if (rc != VV_OK) return rc;
  launch_k(samp_silu_add_kernel, dim3(samp_grid((long long)R * H / 4, 256)), dim3(256), 0, s,
           (const float*)reinterpret_cast<float*>(ws + b.cproj), (const float*)sp->temb_probe,
           reinterpret_cast<__nv_bfloat16*>(ws + b.sc), (int)R, H);
  // all R rows of `noisy` are distinct here: stage them through xlat (sized for R rows) with n = rows = R
  float* XL = reinterpret_cast<float*>(ws + b.xlat);
  VV_CUDA(cudaMemcpyAsync(XL, noisy, sizeof(float) * (size_t)R * Ld, cudaMemcpyDeviceToDevice, s));
  launch_k(samp_dup_latent_kernel, dim3(samp_grid((long long)R * Ld, 256)), dim3(256), 0, s, (const float*)XL,
           reinterpret_cast<__nv_bfloat16*>(ws + b.xin), (int)R, (int)R, Ld);
  d->launches += 2;
  VV_LAUNCH_OK();
  if ((rc = samp_head_pass(sp, ws, b, R, s)) != VV_OK) return rc;
  VV_CUDA(cudaMemcpyAsync(v_out, ws + b.v, sizeof(float) * (size_t)R * Ld, cudaMemcpyDeviceToDevice, s));
  return VV_OK;
}
