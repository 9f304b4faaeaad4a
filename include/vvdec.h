/*
 * vvdec.h — C ABI of the B200-native VibeVoice acoustic-tokenizer decoder (libvvdec_b200.so).
 *
 * Drop-in boundary for ONE path of beecave-homelab/VibeVoice-ROCm:
 *   VibeVoiceAcousticTokenizerModel.decode(latents, cache, sample_indices, use_cache, debug)
 *     reference: vibevoice/modular/modular_vibevoice_tokenizer.py:1099-1108
 *   called by the generation loop at vibevoice/modular/modeling_vibevoice_inference.py:635-641.
 *
 * The reference is pure Python/PyTorch and has no FFI of its own; this header is the interface a
 * maintainer binds with ctypes (see INTEGRATION.md, and vibevoice-rocm_b200/vvdec/_capi.py which is
 * exactly that binding).  Plain C types only: device pointers, sizes, a cudaStream_t passed as
 * void*.  No C++ exceptions cross the boundary; every entry point returns a vv_status.
 *
 * All device pointers must live on the device the decoder was created on.  Calls enqueue work on
 * the given stream and never synchronise it (except vv_decoder_finalize and vv_state_export/import,
 * which are set-up / checkpoint operations).
 */
#ifndef VVDEC_H_
#define VVDEC_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VVDEC_ABI_VERSION 2
#define VV_MAX_STAGES 8

typedef enum vv_status {
  VV_OK = 0,
  VV_ERR_INVALID_ARG = -1,    /* null pointer, negative size, bad enum                         */
  VV_ERR_BAD_SHAPE = -2,      /* tensor shape does not match the geometry                       */
  VV_ERR_BAD_DTYPE = -3,
  VV_ERR_UNSUPPORTED = -4,    /* geometry / config variant outside the path (see vv_config)     */
  VV_ERR_NOT_FINALIZED = -5,  /* decode before vv_decoder_finalize, or a weight is missing      */
  VV_ERR_WORKSPACE = -6,      /* workspace too small                                            */
  VV_ERR_SLOT = -7,           /* stream slot out of range                                       */
  VV_ERR_CUDA = -8,           /* a CUDA call failed; see vv_last_error()                        */
  VV_ERR_NO_DEVICE = -9,      /* no CUDA device / not an sm_100 device                          */
  VV_ERR_UNKNOWN_TENSOR = -10 /* vv_decoder_load_tensor: name is not a decoder parameter        */
} vv_status;

typedef enum vv_dtype { VV_F32 = 0, VV_BF16 = 1 } vv_dtype;

/* latents layout, mirroring the reference's shape sniffing at :1102-1105 */
typedef enum vv_layout {
  VV_LAYOUT_BTD = 0,          /* [B, T, vae_dim]  (what generate() passes, :634-641)            */
  VV_LAYOUT_BDT = 1           /* [B, vae_dim, T]                                                */
} vv_layout;

/*
 * Geometry = the fields of VibeVoiceAcousticTokenizerConfig that fix shapes on the decode path
 * (vibevoice/modular/configuration_vibevoice.py:16-44; parsed at modular_vibevoice_tokenizer.py:
 * 1011-1058).  Variants the path does not implement are rejected with VV_ERR_UNSUPPORTED by the
 * host wrapper before this struct is built: non-causal, LayerNorm, mixer_layer != depthwise_conv,
 * weight/spectral norm, non-constant padding.
 */
typedef struct vv_config {
  int32_t vae_dim;                    /* 64                                                     */
  int32_t n_filters;                  /* decoder_n_filters, 32                                  */
  int32_t n_stages;                   /* len(depths) = len(ratios) + 1, 7                       */
  int32_t ratios[VV_MAX_STAGES];      /* decoder_ratios, [8,5,5,4,2,2]                          */
  int32_t depths[VV_MAX_STAGES];      /* decoder order (already reversed, :1028): [8,3,3,3,3,3,3] */
  int32_t kernel_size;                /* 7 (depthwise mixer, stem, head)                        */
  float   eps;                        /* layernorm_eps 1e-5                                     */
  int32_t device;                     /* CUDA device ordinal                                    */
  int32_t gemm_path;                  /* 0 = product path (whole-Block1D kernel at C<=64, fused FFN with
                                         streamed weights at C=128, tcgen05 GEMMs elsewhere);
                                         1 = CUDA-core validation GEMM everywhere;
                                         2 = unfused tcgen05 GEMMs everywhere (A/B against 0);
                                         3 = as 2 but only the non-persistent GEMM kernel;
                                         4 = as 0 but mixer and FFN halves as separate kernels  */
} vv_config;

typedef struct vv_decoder vv_decoder; /* opaque: packed weights + launch plan                    */
typedef struct vv_state vv_state;     /* opaque: device-resident per-stream conv state            */

/* ---- library ------------------------------------------------------------------------------- */
int32_t     vv_abi_version(void);
const char* vv_status_string(int32_t status);
/* Last CUDA / validation message recorded on this thread ("" if none). */
const char* vv_last_error(void);

/* ---- decoder life cycle (replaces TokenizerDecoder.__init__, :823-912, + from_pretrained) -- */
int32_t vv_decoder_create(const vv_config* cfg, vv_decoder** out);
int32_t vv_decoder_destroy(vv_decoder* dec);

/*
 * Hand one reference state-dict tensor to the decoder.  `name` is the key under the tokenizer model
 * ("decoder.upsample_layers.0.0.conv.conv.weight", "decoder.stages.3.1.ffn.linear1.bias", ... —
 * SURVEY.md Appendix B).  `data` may be a host or a device pointer (is_device says which); it is
 * copied, the caller keeps ownership.  shape is the PyTorch shape.
 */
int32_t vv_decoder_load_tensor(vv_decoder* dec, const char* name, const void* data, int32_t dtype,
                               const int64_t* shape, int32_t ndim, int32_t is_device);
/* Pack weights into kernel-native layouts (bf16 K-major GEMM operands, polyphase-split transposed
 * convs).  Fails with VV_ERR_NOT_FINALIZED and names the first missing tensor in vv_last_error(). */
int32_t vv_decoder_finalize(vv_decoder* dec);
int64_t vv_decoder_param_count(const vv_decoder* dec);
/*
 * Run-time options of one decoder (tests and A/B runs; the defaults are the product routing).  Keys:
 *   "gemm_bn256"       0 = never use the 128x256 GEMM tile, 1 = automatic (default: when the launch has at least
 *                      gemm_bn256_waves waves of wide tiles), 2 = whenever N % 256 == 0
 *   "gemm_bn256_waves" threshold of the automatic mode, in waves of one tile per SM (default 2)
 *   "gemm_bn64"        1 (default) = N = 64 layers on the persistent kernel's 64-wide tile, 0 = one-tile-per-CTA kernel
 *   "gemm_bn64_small"  1 (default) = residual-epilogue GEMMs whose 128-wide tiles would occupy fewer than half the SMs
 *                      (streaming frame steps) use the 64-wide tile; 0 = never.  Results are bit-identical either way
 *   "block_kernel"     1 (default) = block_tma_kernel for whole Block1Ds at C <= 64, 0 = block_fused_kernel
 *   "block_ws"         1 (default) = block_ws_kernel (whole Block1D in one kernel) at C = 128, 0 = mixer_kernel +
 *                      ffn_stream_kernel
 *   "ffn_stream256"    1 (default) = FFN half at C = 256 in one kernel with the hidden activation on chip
 *                      (ffn_stream_kernel<256>), 0 = two GEMMs with the hidden tensor through HBM
 * Unknown keys / values out of range: VV_ERR_INVALID_ARG.  The environment variables VV_GEMM_BN256,
 * VV_GEMM_BN256_WAVES, VV_GEMM_BN64, VV_BLOCK_KERNEL, VV_BLOCK_WS, VV_FFN_STREAM256 set the defaults a decoder is created with.
 */
int32_t vv_decoder_set_option(vv_decoder* dec, const char* key, int64_t value);
/* samples produced per latent frame (prod(ratios) = 3200) and causal receptive field in frames (58) */
int32_t vv_decoder_hop(const vv_decoder* dec);
int32_t vv_decoder_halo_frames(const vv_decoder* dec);

/* ---- decode (replaces VibeVoiceAcousticTokenizerModel.decode, :1099-1108) ------------------- */
/* Scratch bytes vv_decode needs for a [B, T] call (activations ping-pong + GEMM operand staging). */
int64_t vv_decoder_workspace_bytes(const vv_decoder* dec, int32_t B, int32_t T);

/*
 * latents : device, [B,T,vae_dim] or [B,vae_dim,T] per `layout`, dtype lat_dtype.
 * audio   : device, [B, 1, T*hop] contiguous, dtype out_dtype.
 * state   : NULL for offline decode (use_cache=False / cache=None, :317,:469).  Otherwise the
 *           per-stream conv state; slot_ids (device int32 [B], one distinct slot per row — the
 *           reference's sample_indices, :322-323) selects each row's stream.  State is read as the
 *           causal left context and updated in place (SConv1d._forward_streaming :327-382,
 *           SConvTranspose1d._forward_streaming :478-549); a zeroed slot reproduces the reference's
 *           "first chunk" zeros context (:337-346, :488-492).
 * Any T >= 1 is accepted in both modes.
 */
int32_t vv_decode(vv_decoder* dec, const void* latents, int32_t lat_dtype, int32_t layout,
                  int32_t B, int32_t T, void* audio, int32_t out_dtype,
                  void* workspace, int64_t workspace_bytes,
                  vv_state* state, const int32_t* slot_ids, void* cuda_stream);

/* Number of kernels vv_decode enqueued since the counter was last read (bench.py's gpu_launches). */
int64_t vv_decoder_take_launch_count(vv_decoder* dec);

/*
 * Per-launch timing for bench.py's roofline leg.  While enabled, vv_decode records a CUDA event on the
 * decode stream after every kernel launch and synchronises the stream before returning; the elapsed
 * time of each launch is accumulated per (kind, stage).  vv_decoder_read_profile copies up to `cap`
 * records, clears the accumulator and returns the number of records available.
 */
typedef enum vv_kernel_kind {
  VV_K_PREP = 0, VV_K_STEM = 1, VV_K_MIXER = 2, VV_K_FFN1 = 3, VV_K_FFN2 = 4, VV_K_HALO = 5,
  VV_K_UPSAMPLE = 6, VV_K_HEAD = 7, VV_K_STATE = 8, VV_K_FFN_FUSED = 9, VV_K_BLOCK_FUSED = 10,
  VV_K_ENC_STEM = 11, VV_K_DOWNSAMPLE = 12, VV_K_PAD = 13, VV_K_ENC_HEAD = 14,  /* encoder path */
  VV_K_FFN_PAIR = 15   /* both FFN GEMMs of a block run in L2-resident row chunks: one record for the pair */
} vv_kernel_kind;
typedef struct vv_profile_record {
  int32_t kind;        /* vv_kernel_kind */
  int32_t stage;
  int64_t launches;
  double  ms;          /* summed device time of those launches */
} vv_profile_record;
int32_t vv_decoder_set_profiling(vv_decoder* dec, int32_t enabled);
int32_t vv_decoder_read_profile(vv_decoder* dec, vv_profile_record* out, int32_t cap);

/* ---- streaming state (replaces VibeVoiceTokenizerStreamingCache, :193-256) ------------------ */
int32_t vv_state_create(const vv_decoder* dec, int32_t n_slots, vv_state** out);   /* zero-filled */
int32_t vv_state_destroy(vv_state* st);
int32_t vv_state_n_slots(const vv_state* st);
int64_t vv_state_bytes_per_slot(const vv_state* st);
/*
 * Slot ids are consumed on the device (no host sync, unlike the reference's sample_indices.tolist(), :204,230,238), so a
 * bad id cannot fail the call that carries it.  Kernels never index the state with an id outside [0, n_slots): such a row
 * is decoded with a zero context and saves nothing, the id is recorded, and the NEXT vv_decode / vv_encode /
 * vv_state_zero / vv_state_export on the state -- or vv_state_check, which does nothing else -- returns VV_ERR_SLOT once
 * (the record is host-mapped memory: no synchronisation; to observe an error of a call still in flight, synchronise the
 * stream first).  Slot id -1 is a PADDING ROW: decoded with a zero context, nothing saved, no error -- it lets a caller
 * pad a step to a fixed row count (one CUDA graph per bucket of stream counts, vvdec.FrameStep).
 */
int32_t vv_state_check(vv_state* st);
/* set_to_zero(sample_indices) (:234-241): slot_ids is a device int32 [n] array; n<0 zeroes all. */
int32_t vv_state_zero(vv_state* st, const int32_t* slot_ids, int32_t n, void* cuda_stream);
/* checkpoint one slot to / from host memory (vv_state_bytes_per_slot bytes); synchronises. */
int32_t vv_state_export(const vv_state* st, int32_t slot, void* host_dst, int64_t bytes);
int32_t vv_state_import(vv_state* st, int32_t slot, const void* host_src, int64_t bytes);

/* ---- encoder (replaces TokenizerEncoder.forward :810-813 under VibeVoiceAcousticTokenizerModel.encode :1081-1085) --
 * The encoder mirrors the decoder's vv_config (same struct, decoder-order fields): stage i has n_filters * 2^i
 * channels, depth depths[n_stages-1-i] and is entered through a strided conv of ratio ratios[n_stages-1-i]
 * (:701, :1019-1028).  Tensor names are the reference's "encoder.*" keys. */
typedef struct vv_encoder vv_encoder;
int32_t vv_encoder_create(const vv_config* cfg, vv_encoder** out);
int32_t vv_encoder_destroy(vv_encoder* enc);
int32_t vv_encoder_load_tensor(vv_encoder* enc, const char* name, const void* data, int32_t dtype,
                               const int64_t* shape, int32_t ndim, int32_t is_device);
int32_t vv_encoder_finalize(vv_encoder* enc);
int32_t vv_encoder_set_option(vv_encoder* enc, const char* key, int64_t value);   /* same keys as vv_decoder_set_option */
int64_t vv_encoder_param_count(const vv_encoder* enc);
int32_t vv_encoder_hop(const vv_encoder* enc);
/* left context, in latent frames, that makes a time chunk starting on a hop boundary reproduce the monolithic encode
 * (time-chunk sharding across GPUs, SURVEY.md 8e last row): 58 for the shipped geometry, like the decoder's */
int32_t vv_encoder_halo_frames(const vv_encoder* enc);
/* latent frames for L samples: ceil(L / hop) (each strided conv pads its input up to a whole stride, :127-133) */
int64_t vv_encoder_frames(const vv_encoder* enc, int64_t L);
int64_t vv_encoder_workspace_bytes(const vv_encoder* enc, int32_t B, int64_t L);
/*
 * audio : device, [B, L] (= the reference's [B, 1, L]) contiguous, f32 or bf16.
 * mean  : device, fp32 [B, frames, vae_dim] contiguous — VibeVoiceTokenizerEncoderOutput.mean (:1084-1085; its
 *         std is the config's fix_std, a host-side constant).
 * state : NULL for offline encode (cache=None / use_cache=False).  Otherwise a state made by vv_encoder_state_create
 *         (all vv_state_* calls apply); slot_ids (device int32 [B]) = sample_indices.  Each layer's left context comes
 *         from the slot and is updated in place (SConv1d._forward_streaming :327-382); a zeroed slot = first chunk.
 *         Streaming chunks must be whole hops (L % hop == 0, VV_ERR_BAD_SHAPE otherwise): then no strided conv pads
 *         on the right and the stream reproduces the offline result.
 * A workspace smaller than vv_encoder_workspace_bytes(B, L) is honoured by encoding the utterances in groups;
 * one utterance must fit (VV_ERR_WORKSPACE otherwise).
 */
int32_t vv_encoder_state_create(const vv_encoder* enc, int32_t n_slots, vv_state** out);   /* zero-filled */
int32_t vv_encode(vv_encoder* enc, const void* audio, int32_t audio_dtype, int32_t B, int64_t L, float* mean,
                  void* workspace, int64_t workspace_bytes, vv_state* state, const int32_t* slot_ids,
                  void* cuda_stream);
int64_t vv_encoder_take_launch_count(vv_encoder* enc);
int32_t vv_encoder_set_profiling(vv_encoder* enc, int32_t enabled);
int32_t vv_encoder_read_profile(vv_encoder* enc, vv_profile_record* out, int32_t cap);

/* Encoder per-kernel hooks (offline form; tests):
 *  stem SConv1d(1 -> 32, k7) (:732-734): audio [B, L] -> x fp32 [B, L, 32];
 *  strided SConv1d(C/2 -> C, k = 2r, stride r) INTO encoder stage `stage` (1..n_stages-1) (:737-743): xb is the padded bf16
 *    operand copy [B, (ceil(rows / r) + 1) * r, C/2] whose rows [r, r + rows) hold the input (the call zeroes the r
 *    context rows and the ragged tail); y fp32 [B, ceil(rows / r), C];
 *  head SConv1d(C_last -> vae_dim, k7) (:771): xb bf16 [B, rows + 6, C_last], rows [6, 6 + rows) = input; mean fp32
 *    [B, rows, vae_dim]. */
int32_t vv_kernel_enc_stem(vv_encoder* enc, const void* audio, int32_t audio_dtype, float* x, int32_t B, int64_t L,
                           void* cuda_stream);
int32_t vv_kernel_downsample(vv_encoder* enc, int32_t stage, void* xb, float* y, int32_t B, int32_t rows,
                             void* cuda_stream);
int32_t vv_kernel_enc_head(vv_encoder* enc, void* xb, float* mean, int32_t B, int32_t rows, void* cuda_stream);

/* ---- per-kernel entry points (offline form; tests and bench.py's roofline leg time these alone) -- */
/* Fused Block1D mixer half on a channels-last fp32 activation x [B, rows, C]:
 *   y = x + gamma * (dwconv7(RMSNorm(x) * norm_w) + bias);   a = bf16( RMSNorm(y) * ffn_norm_w )
 * (modular_vibevoice_tokenizer.py:927-932, :936).  stage/block pick the packed weights. */
int32_t vv_kernel_mixer(vv_decoder* dec, int32_t stage, int32_t block, const float* x, float* y,
                        void* a_out, int32_t B, int32_t rows, void* cuda_stream);
/* Fused Block1D FFN half: out = resid + ffn_gamma * (W2 gelu(W1 a + b1) + b2)  (:937-942).
 * a: bf16 [B*rows, C]; hidden: bf16 scratch [B*rows, 4C]; resid/out: fp32 [B*rows, C] (may alias). */
int32_t vv_kernel_ffn(vv_decoder* dec, int32_t stage, int32_t block, const void* a, const float* resid,
                      float* out, void* hidden, int32_t B, int32_t rows, void* cuda_stream);
/* Whole Block1D (mixer half + FFN half, :927-942) in ONE kernel, stages with C <= 64 only (VV_ERR_UNSUPPORTED
 * otherwise): x, out fp32 [B, rows, C]; out must not alias x. */
int32_t vv_kernel_block(vv_decoder* dec, int32_t stage, int32_t block, const float* x, float* out,
                        int32_t B, int32_t rows, void* cuda_stream);
/* Polyphase transposed conv INTO stage `stage` (1..n_stages-1), (:875-878 + trim :563-571):
 * xb: bf16 [B, rows+1, Cin], rows 1..rows = input, row 0 = causal context (zeroed by this call);
 * y: fp32 [B, rows*ratio, Cout]. */
int32_t vv_kernel_upsample(vv_decoder* dec, int32_t stage, void* xb, float* y, int32_t B, int32_t rows,
                           void* cuda_stream);
/* Head SConv1d(C_last -> 1, k7) (:912, :950): x fp32 [B, rows, C_last] -> out [B, rows] (f32 | bf16). */
int32_t vv_kernel_head(vv_decoder* dec, const float* x, void* out, int32_t out_dtype, int32_t B,
                       int32_t rows, void* cuda_stream);

/* ---- PCM16 output step ------------------------------------------------------------------------------
 * Replaces convert_to_16_bit_wav (reference main.py:2956-2979; same function at demo/gradio_demo.py:1205), run by the
 * Gradio path on every streamed chunk (main.py:937-951) and on the final waveform:
 *     per row of wave [B, L]:  peak = max|x|;  if peak > 1: x = x / peak;  pcm = (int16) trunc(x * 32767)
 * in fp32 with IEEE division / multiplication and truncation, i.e. bit-exact with the numpy original.  A row is one
 * chunk / utterance (the reference converts each sample's chunk separately).
 *   wave: device [B, L] f32 | bf16, rows row_stride elements apart (decode() output [B, 1, L] has row_stride = L);
 *   pcm:  device int16 [B, L] contiguous;   peak: device fp32 [B], receives max|x| per row -- may be NULL when
 *   L <= 8192 (one launch, row kept in registers); required for longer rows (peak pass + convert pass).
 * Stateless; runs on the calling thread's current device. */
int32_t vv_pcm16(const void* wave, int32_t wave_dtype, int32_t B, int64_t L, int64_t row_stride, int16_t* pcm,
                 float* peak, void* cuda_stream);

/* ---- PCM16 without any allocation -----------------------------------------------------------------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot allocate (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  Same result as vv_pcm16, bit for bit; the
 * per-row peak scratch the long-row form needs lives in a file-scope `static __device__` array inside the library instead
 * of a caller buffer, and the call performs no cudaMalloc / cudaFree and calls no allocating routine.
 *   limit    : B <= VV_PCM16_STATIC_MAX_ROWS rows per call (VV_ERR_BAD_SHAPE beyond it; L is unlimited);
 *   cost     : every GPU that loads the library holds the array (4 B x VV_PCM16_STATIC_MAX_ROWS = 16 KB) until exit;
 *   sharing  : all calls on a device share the array, which keeps the last call's contents; each call clears what it
 *              needs on `cuda_stream`, so calls ordered on one stream are safe, while calls that OVERLAP on the GPU
 *              (different streams without ordering) RACE on it -- use vv_pcm16 with separate peak buffers for those;
 *   set-up   : the first call on a device resolves the array's device address (cudaGetSymbolAddress, once per device,
 *              with that device current) -- make that first call outside stream capture.
 * Runs on the calling thread's current device, like vv_pcm16. */
#define VV_PCM16_STATIC_MAX_ROWS 4096
int32_t vv_pcm16_static(const void* wave, int32_t wave_dtype, int32_t B, int64_t L, int64_t row_stride, int16_t* pcm,
                        void* cuda_stream);

/* ---- diffusion-head sampler (the step right before the decoder) -------------------------------------
 * Replaces VibeVoiceForConditionalGenerationInference.sample_speech_tokens (modeling_vibevoice_inference.py:695-708): for n
 * streams, ddpm_num_inference_steps x [ VibeVoiceDiffusionHead.forward on 2n rows (modular_vibevoice_diffusion_head.py:
 * 254-280), classifier-free guidance (:703-705), DPMSolverMultistepScheduler.step (vibevoice/schedule/dpm_solver.py:
 * 935-1022) ] with the scheduler as the model builds it (modeling_vibevoice.py:138-142: cosine betas, v-prediction; every
 * other argument at its default: order 2, dpmsolver++, midpoint, lower_order_final, final sigma 0, linspace spacing).
 * Tensor names are the head's own state-dict keys ("noisy_images_proj.weight", "layers.0.ffn.gate_proj.weight", ...;
 * under "model.prediction_head." in the full model).  The initial noise (torch.randn, :698) is an INPUT. */
typedef struct vv_sampler_config {
  int32_t hidden;        /* VibeVoiceDiffusionHeadConfig.hidden_size (1536 / 3584); multiple of 128, <= 4096      */
  int32_t layers;        /* head_layers (4)                                                                      */
  int32_t ffn;           /* int(hidden_size * head_ffn_ratio) (4608 / 10752); multiple of 128                    */
  int32_t latent;        /* latent_size (64; the kernels take exactly 64)                                        */
  float   eps;           /* rms_norm_eps (1e-5)                                                                  */
  int32_t train_steps;   /* ddpm_num_steps (1000)                                                                */
  int32_t device;
} vv_sampler_config;
typedef struct vv_sampler vv_sampler;
int32_t vv_sampler_create(const vv_sampler_config* cfg, vv_sampler** out);
int32_t vv_sampler_destroy(vv_sampler* sp);
int32_t vv_sampler_load_tensor(vv_sampler* sp, const char* name, const void* data, int32_t dtype, const int64_t* shape,
                               int32_t ndim, int32_t is_device);
int32_t vv_sampler_finalize(vv_sampler* sp);
int64_t vv_sampler_param_count(const vv_sampler* sp);
/* Solver coefficients and timestep embeddings for a step count (set_timesteps, :321-424; TimestepEmbedder :48-93), computed on
 * the host in fp32 and uploaded; allocates once per distinct `steps`.  vv_sampler_sample calls it lazily: call it yourself
 * before capturing a CUDA graph. */
int32_t vv_sampler_prepare(vv_sampler* sp, int32_t steps);
int64_t vv_sampler_workspace_bytes(const vv_sampler* sp, int32_t n);
/* condition, neg_condition: device [n, hidden] (f32 | bf16);  noise: device fp32 [n, latent] = the initial latent;
 * out: device fp32 [n, latent] = sample_speech_tokens' return value.  Enqueues on the stream, never synchronises. */
int32_t vv_sampler_sample(vv_sampler* sp, const void* condition, const void* neg_condition, int32_t cond_dtype, int32_t n,
                          const float* noise, float cfg_scale, int32_t steps, float* out, void* workspace,
                          int64_t workspace_bytes, void* cuda_stream);
/* Per-kernel hook (tests): ONE head forward (:254-280) on R rows (R even): noisy fp32 [R, latent], one timestep for all rows,
 * condition [R, hidden] -> v fp32 [R, latent].  Synchronises the stream (it uploads the timestep embedding). */
int32_t vv_sampler_head(vv_sampler* sp, const float* noisy, float timestep, const void* condition, int32_t cond_dtype,
                        int32_t R, float* v_out, void* workspace, int64_t workspace_bytes, void* cuda_stream);
int64_t vv_sampler_take_launch_count(vv_sampler* sp);

/* The GELU of the FFN epilogue (:589, :594), alone: y[i] = gelu(x[i]); device pointers.  Accuracy test hook. */
int32_t vv_test_gelu(const float* x, float* y, int32_t n, void* cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* VVDEC_H_ */
