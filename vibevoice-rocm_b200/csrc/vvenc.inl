// vvenc.inl — host side of the ENCODER path behind the C ABI (included at the end of vvdec.cu: it reuses that file's
// weight packer and launch helpers).  Replaces TokenizerEncoder.forward (modular_vibevoice_tokenizer.py:810-813) as
// called by VibeVoiceAcousticTokenizerModel.encode (:1081-1085), offline and streaming form (whole-hop chunks).
//
// The encoder is the decoder's stage list walked backwards: stage i has C = n_filters * 2^i channels, the decoder's
// depth of stage n-1-i (:1019-1028) and is entered through a strided SConv1d with the decoder's ratio n-1-i (:701).

struct vv_encoder {
  vv_decoder core;              // cfg (decoder-order fields), stages in ENCODER order, launch / profiling bookkeeping
  float* stem_w = nullptr;      // [7][32] tap-major
  float* stem_b = nullptr;      // [32]
  vv::DenseW head;              // SConv1d(C_last -> vae_dim, k7) as a 7-tap GEMM
  // stream-state layout inside a slot (core.state_slot_bytes): stem fp32 [6]; per stage i >= 1 (StageW::up_state_off)
  // bf16 [r][C/2] = the strided conv's context; per block fp32 [6][C] (BlockW::state_off); head bf16 [6][C_last]
  long long stem_state_off = 0, head_state_off = 0;
};

namespace {

int enc_channels(const vv_config& c, int i) { return c.n_filters << i; }
int enc_depth(const vv_config& c, int i) { return c.depths[c.n_stages - 1 - i]; }
int enc_ratio(const vv_config& c, int i) { return c.ratios[c.n_stages - 1 - i]; }   // into stage i >= 1

// rows of every stage for L input samples: each strided conv yields ceil(rows / r) (:127-133)
void enc_rows(const vv_config& c, long long L, long long* rows) {
  rows[0] = L;
  for (int i = 1; i < c.n_stages; ++i) {
    const int r = enc_ratio(c, i);
    rows[i] = (rows[i - 1] + r - 1) / r;
  }
}

struct EncBufs { long long x, y, a, h, xb, total; long long xb_rpb; int xb_off; };

// per-stage buffers for nb items (1 KB aligned).  xb = padded bf16 copy of the stage output, the A operand of the
// next strided conv ([nb, (rows_next + 1) * r, C]) or of the head ([nb, rows + 6, C]).
EncBufs enc_bufs(const vv_encoder* e, int i, long long nb, const long long* rows) {
  const vv_config& c = e->core.cfg;
  const long long C = enc_channels(c, i), R = nb * rows[i];
  const bool whole_block = block_fused_supported(&e->core, (int)C);
  const bool fused_ffn = whole_block || ffn_fused_supported(&e->core, (int)C);
  EncBufs b;
  long long off = 0;
  b.x = off; off += align_up(R * C * 4, 1024);
  b.y = off; off += align_up(R * C * 4, 1024);
  b.a = off; if (!whole_block) off += align_up(R * C * 2, 1024);
  b.h = off; if (!fused_ffn) off += align_up(R * C * 4 * 2, 1024);
  if (i + 1 < c.n_stages) {
    const int r = enc_ratio(c, i + 1);
    b.xb_rpb = (rows[i + 1] + 1) * r;
    b.xb_off = r;
  } else {
    b.xb_rpb = rows[i] + 6;
    b.xb_off = 6;
  }
  b.xb = off; off += align_up(nb * b.xb_rpb * C * 2, 1024);
  b.total = off;
  return b;
}

long long enc_pass_bytes(const vv_encoder* e, long long nb, long long L) {
  long long rows[VV_MAX_STAGES];
  enc_rows(e->core.cfg, L, rows);
  long long arena = 0;
  for (int i = 0; i < e->core.cfg.n_stages; ++i) arena = std::max(arena, enc_bufs(e, i, nb, rows).total);
  return 2 * arena + 1024;
}

int enc_run_pass(vv_encoder* e, const void* audio, int audio_is_bf16, long long audio_bstride, int nb, long long L,
                 float* mean, uint8_t* ws, vv_state* st, const int32_t* slots, cudaStream_t s) {
  vv_decoder* d = &e->core;
  const vv_config& c = d->cfg;
  const int S = c.n_stages;
  long long rows[VV_MAX_STAGES];
  enc_rows(c, L, rows);
  long long arena = 0;
  for (int i = 0; i < S; ++i) arena = std::max(arena, enc_bufs(e, i, nb, rows).total);
  uint8_t* ar[2] = {ws, ws + arena};
  d->mark(-1, 0, s);
  int rc;
  for (int i = 0; i < S; ++i) {
    const StageW& sw = d->stages[i];
    const EncBufs bo = enc_bufs(e, i, nb, rows);
    uint8_t* base = ar[i & 1];
    float* X = reinterpret_cast<float*>(base + bo.x);
    float* Y = reinterpret_cast<float*>(base + bo.y);
    __nv_bfloat16* A = reinterpret_cast<__nv_bfloat16*>(base + bo.a);
    __nv_bfloat16* H = reinterpret_cast<__nv_bfloat16*>(base + bo.h);
    __nv_bfloat16* XB = reinterpret_cast<__nv_bfloat16*>(base + bo.xb);
    const int C = sw.C;
    if (rows[i] * (long long)nb >= (1LL << 31)) { set_error("encode: %lld rows in one pass", rows[i] * nb); return VV_ERR_BAD_SHAPE; }
    const int rows_pb = (int)rows[i];
    if (i == 0) {
      EncStemArgs p;
      p.audio = audio; p.audio_is_bf16 = audio_is_bf16; p.audio_bstride = audio_bstride;
      p.w = e->stem_w; p.b = e->stem_b; p.x = X; p.nbatch = nb; p.L = rows_pb;
      p.st = make_state_ref(st, slots, e->stem_state_off);
      const long long threads = ((long long)rows_pb + ENC_STEM_ROWS - 1) / ENC_STEM_ROWS * nb * 8;
      launch_k(enc_stem_kernel, dim3((unsigned)std::min<long long>((threads + 255) / 256, 148LL * 64)), dim3(256), 0, s, p);
      d->launches++;
      VV_LAUNCH_OK();
      d->mark(VV_K_ENC_STEM, 0, s);
    } else {
      // strided conv from the previous stage's padded bf16 copy: 2 taps of K_tap = r * C/2 (encoder.cuh)
      const EncBufs pbo = enc_bufs(e, i - 1, nb, rows);
      const __nv_bfloat16* PXB = reinterpret_cast<const __nv_bfloat16*>(ar[(i - 1) & 1] + pbo.xb);
      EpiOut eo;
      eo.epi = EPI_BIAS_F32; eo.out_f32 = X;
      rc = launch_gemm(d, PXB, rows[i] + 1, rows_pb, nb, sw.up, eo, s, VV_K_DOWNSAMPLE, i);
      if (rc != VV_OK) return rc;
    }
    PadOut po;
    po.rpb = bo.xb_rpb; po.off = bo.xb_off;
    const int nblk = (int)sw.blocks.size();
    if (nblk == 0) { set_error("encoder stage %d has no blocks", i); return VV_ERR_UNSUPPORTED; }
    for (int j = 0; j < nblk; ++j) {
      const BlockW& bw = sw.blocks[j];
      const bool last = (j == nblk - 1);
      if (block_fused_supported(d, C)) {
        rc = launch_block_fused(d, C, bw, X, Y, last ? XB : nullptr, nb, rows_pb, st, slots, s, po);
        if (rc != VV_OK) return rc;
        d->mark(VV_K_BLOCK_FUSED, i, s);
        std::swap(X, Y);
        continue;
      }
      rc = launch_mixer(d, bw, C, X, Y, A, nb, rows_pb, st, slots, s);
      if (rc != VV_OK) return rc;
      d->mark(VV_K_MIXER, i, s);
      if (ffn_fused_supported(d, C)) {
        if (C == 256 && !last) {         // in place (see run_pass): y becomes the block output
          rc = launch_ffn_fused(d, C, bw, A, Y, Y, nullptr, (long long)nb * rows_pb, rows_pb, s, po);
          std::swap(X, Y);
        } else {
          rc = launch_ffn_fused(d, C, bw, A, Y, X, last ? XB : nullptr, (long long)nb * rows_pb, rows_pb, s, po);
        }
        if (rc != VV_OK) return rc;
        d->mark(VV_K_FFN_FUSED, i, s);
        continue;
      }
      rc = launch_ffn_pair(d, C, bw, A, Y, X, H, last ? XB : nullptr, po, nb, rows_pb, i, s);
      if (rc != VV_OK) return rc;
    }
    {  // context rows (zeros, or the stream's cached rows) and right padding of this stage's bf16 copy; the data rows
       // were written by the last block's epilogue.  With a state this also saves the stream's new context.
      EncContextArgs cp;
      cp.xb = XB; cp.nbatch = nb; cp.rows = rows_pb; cp.off = bo.xb_off; cp.C = C; cp.rpb = bo.xb_rpb;
      cp.st = make_state_ref(st, slots, i + 1 < S ? d->stages[i + 1].up_state_off : e->head_state_off);
      const long long cells = (bo.xb_rpb - rows_pb) * nb * (C / 8);
      launch_k(enc_context_kernel, dim3((unsigned)std::max<long long>(1, std::min<long long>((cells + 255) / 256, 148LL * 16))), dim3(256), 0, s, cp);
      d->launches++;
      VV_LAUNCH_OK();
      d->mark(VV_K_PAD, i, s);
    }
    if (i == S - 1) {   // self.norm is Identity (disable_last_norm, :767-770); head k7 -> mean [nb, T, vae_dim]
      EpiOut eo;
      eo.epi = EPI_BIAS_F32; eo.out_f32 = mean;
      rc = launch_gemm(d, XB, rows[i] + 6, rows_pb, nb, e->head, eo, s, VV_K_ENC_HEAD, i);
      if (rc != VV_OK) return rc;
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
// C ABI
// =====================================================================================================
extern "C" int32_t vv_encoder_create(const vv_config* cfg, vv_encoder** out) {
  if (cfg == nullptr || out == nullptr) { set_error("null argument"); return VV_ERR_INVALID_ARG; }
  *out = nullptr;
  vv_encoder* e = new vv_encoder();
  const int32_t rc = init_core(cfg, &e->core);
  if (rc != VV_OK) { delete e; return rc; }
  *out = e;
  return VV_OK;
}

extern "C" int32_t vv_encoder_destroy(vv_encoder* e) {
  if (e == nullptr) return VV_OK;
  cudaSetDevice(e->core.cfg.device);
  for (void* p : e->core.allocs) cudaFree(p);
  delete e;
  return VV_OK;
}

extern "C" int32_t vv_encoder_load_tensor(vv_encoder* e, const char* name, const void* data, int32_t dtype,
                                          const int64_t* shape, int32_t ndim, int32_t is_device) {
  return stage_tensor(e ? &e->core : nullptr, "encoder.", name, data, dtype, shape, ndim, is_device);
}

extern "C" int32_t vv_encoder_finalize(vv_encoder* e) {
  if (!e) { set_error("null encoder"); return VV_ERR_INVALID_ARG; }
  vv_decoder* d = &e->core;
  if (d->finalized) return VV_OK;
  VV_CUDA(cudaSetDevice(d->cfg.device));
  for (void* p : d->allocs) cudaFree(p);
  d->allocs.clear();
  d->stages.clear();
  d->param_count = 0;
  const vv_config& c = d->cfg;
  Packer pk{d};
  long long soff = 0;
  e->stem_state_off = soff;
  soff += 32;                                               // fp32 [6], padded to 32 B
  {  // stem SConv1d(1 -> n_filters, k7) (:732-734), fp32 tap-major
    const int C = enc_channels(c, 0);
    const auto* w = pk.get("encoder.downsample_layers.0.0.conv.conv.weight", {C, 1, 7});
    const auto* b = pk.get("encoder.downsample_layers.0.0.conv.conv.bias", {C});
    if (w && b) {
      std::vector<float> wt((size_t)7 * C);
      for (int ch = 0; ch < C; ++ch)
        for (int t = 0; t < 7; ++t) wt[(size_t)t * C + ch] = (*w)[(size_t)ch * 7 + t];
      e->stem_w = pk.upload(wt);
      e->stem_b = pk.upload(*b);
    }
  }
  for (int i = 0; i < c.n_stages && pk.status == VV_OK; ++i) {
    StageW st;
    st.C = enc_channels(c, i);
    const int C = st.C;
    if (i > 0) {
      // SConv1d(C/2 -> C, k = 2r, stride r) (:737-743): Wp[co][j*Cin + ci] = w[co][ci][j]; taps 0..r-1 read padded
      // row-group q, taps r..2r-1 row-group q+1 (encoder.cuh)
      const int r = enc_ratio(c, i), Cin = C / 2;
      const std::string p = "encoder.downsample_layers." + std::to_string(i) + ".0.conv.conv.";
      const auto* w = pk.get(p + "weight", {C, Cin, 2 * r});
      const auto* b = pk.get(p + "bias", {C});
      if (!w || !b) break;
      const size_t K = (size_t)2 * r * Cin;
      std::vector<uint16_t> wp((size_t)C * K);
      for (int co = 0; co < C; ++co)
        for (int ci = 0; ci < Cin; ++ci)
          for (int j = 0; j < 2 * r; ++j)
            wp[(size_t)co * K + (size_t)j * Cin + ci] = f32_to_bf16_rne((*w)[((size_t)co * Cin + ci) * (2 * r) + j]);
      pk.dense(st.up, wp, *b, C, r * Cin, 2);
      st.ratio = r;
      st.up_state_off = soff;                               // bf16 [r][Cin]: context_size = k - 1 - (r - 1) = r rows (:282)
      soff += align_up((long long)r * Cin * 2, 16);
    }
    for (int j = 0; j < enc_depth(c, i) && pk.status == VV_OK; ++j) {
      BlockW bw;
      if (!pk.block("encoder.stages." + std::to_string(i) + "." + std::to_string(j) + ".", C, bw)) break;
      bw.state_off = soff;
      soff += 6LL * C * 4;
      st.blocks.push_back(bw);
    }
    d->stages.push_back(st);
  }
  if (pk.status == VV_OK) {   // head SConv1d(C_last -> vae_dim, k7) (:771): tap j reads padded row t + j = x[t-6+j]
    const int C = enc_channels(c, c.n_stages - 1), D = c.vae_dim;
    const auto* w = pk.get("encoder.head.conv.conv.weight", {D, C, 7});
    const auto* b = pk.get("encoder.head.conv.conv.bias", {D});
    if (w && b) {
      std::vector<uint16_t> wp((size_t)D * 7 * C);
      for (int co = 0; co < D; ++co)
        for (int ci = 0; ci < C; ++ci)
          for (int j = 0; j < 7; ++j)
            wp[(size_t)co * 7 * C + (size_t)j * C + ci] = f32_to_bf16_rne((*w)[((size_t)co * C + ci) * 7 + j]);
      pk.dense(e->head, wp, *b, D, C, 7);
      e->head_state_off = soff;                             // bf16 [6][C_last]
      soff += align_up(6LL * C * 2, 16);
    }
  }
  if (pk.status != VV_OK) return pk.status;
  d->state_slot_bytes = align_up(soff, 256);
  d->staged.clear();
  d->staged_shape.clear();
  d->finalized = true;
  return VV_OK;
}

extern "C" int32_t vv_encoder_set_option(vv_encoder* e, const char* key, int64_t value) {
  if (!e) { set_error("null encoder"); return VV_ERR_INVALID_ARG; }
  return vv_decoder_set_option(&e->core, key, value);
}
extern "C" int64_t vv_encoder_param_count(const vv_encoder* e) { return e ? e->core.param_count : 0; }
extern "C" int32_t vv_encoder_hop(const vv_encoder* e) { return e ? e->core.hop : 0; }
// Causal receptive field in latent frames beyond a frame's own hop of samples.  The encoder is the decoder's mirror: a
// stride-1 k7 layer adds 6 rows of left context at its rate, a strided conv (k = 2r) adds r input rows = 1 output row,
// which is term by term the sum init_core evaluates for the decoder (58 frames for the shipped geometry).
extern "C" int32_t vv_encoder_halo_frames(const vv_encoder* e) { return e ? e->core.halo : 0; }
extern "C" int64_t vv_encoder_frames(const vv_encoder* e, int64_t L) {
  if (!e || L < 1) return 0;
  long long rows[VV_MAX_STAGES];
  enc_rows(e->core.cfg, L, rows);
  return rows[e->core.cfg.n_stages - 1];
}
extern "C" int64_t vv_encoder_take_launch_count(vv_encoder* e) { return e ? vv_decoder_take_launch_count(&e->core) : 0; }
extern "C" int32_t vv_encoder_set_profiling(vv_encoder* e, int32_t enabled) {
  if (!e) { set_error("null encoder"); return VV_ERR_INVALID_ARG; }
  return vv_decoder_set_profiling(&e->core, enabled);
}
extern "C" int32_t vv_encoder_read_profile(vv_encoder* e, vv_profile_record* out, int32_t cap) {
  if (!e) return 0;
  return vv_decoder_read_profile(&e->core, out, cap);
}

extern "C" int64_t vv_encoder_workspace_bytes(const vv_encoder* e, int32_t B, int64_t L) {
  if (!e || !e->core.finalized || B < 1 || L < 1) return 0;
  return enc_pass_bytes(e, B, L);
}

extern "C" int32_t vv_encoder_state_create(const vv_encoder* e, int32_t n_slots, vv_state** out) {
  if (!e) { set_error("null encoder"); return VV_ERR_INVALID_ARG; }
  return vv_state_create(&e->core, n_slots, out);
}

extern "C" int32_t vv_encode(vv_encoder* e, const void* audio, int32_t audio_dtype, int32_t B, int64_t L, float* mean,
                             void* workspace, int64_t workspace_bytes, vv_state* state, const int32_t* slot_ids,
                             void* cuda_stream) {
  if (!e || !audio || !mean || !workspace) { set_error("null argument"); return VV_ERR_INVALID_ARG; }
  vv_decoder* d = &e->core;
  if (!d->finalized) { set_error("vv_encoder_finalize has not succeeded"); return VV_ERR_NOT_FINALIZED; }
  if (B < 1 || L < 1 || L >= (1LL << 31)) { set_error("B=%d L=%lld", B, (long long)L); return VV_ERR_BAD_SHAPE; }
  if (audio_dtype != VV_F32 && audio_dtype != VV_BF16) { set_error("audio dtype must be f32 or bf16"); return VV_ERR_BAD_DTYPE; }
  if (((uintptr_t)workspace & 1023) != 0) { set_error("workspace must be 1 KB aligned"); return VV_ERR_INVALID_ARG; }
  if (state != nullptr && slot_ids == nullptr) { set_error("state given without slot_ids"); return VV_ERR_INVALID_ARG; }
  if (state != nullptr && state->dec != d) { set_error("state belongs to another encoder / decoder"); return VV_ERR_INVALID_ARG; }
  if (state != nullptr) { const int32_t prev = state_take_error(state); if (prev != VV_OK) return prev; }
  PdlScope pdl(state != nullptr && (long long)B * (L / d->hop) <= 2048);
  if (state != nullptr && L % d->hop != 0) {
    // a ragged chunk would be right-padded by every strided conv in mid-stream (:127-133): not the same signal
    set_error("streaming encode takes whole hops: L=%lld is not a multiple of %d", (long long)L, d->hop);
    return VV_ERR_BAD_SHAPE;
  }
  VV_CUDA(cudaSetDevice(d->cfg.device));
  cudaStream_t s = reinterpret_cast<cudaStream_t>(cuda_stream);
  uint8_t* ws = reinterpret_cast<uint8_t*>(workspace);
  const size_t asz = audio_dtype == VV_F32 ? 4 : 2;
  const long long T = vv_encoder_frames(e, L);
  int32_t rc = VV_OK;
  if (enc_pass_bytes(e, 1, L) > workspace_bytes) {
    // one utterance must fit; a caller with less memory feeds whole-hop chunks through a stream state
    set_error("workspace of %lld bytes cannot hold one utterance of %lld samples (%lld needed)",
              (long long)workspace_bytes, (long long)L, enc_pass_bytes(e, 1, L));
    rc = VV_ERR_WORKSPACE;
  } else {
    int nb = 1;                                          // split by utterance (independent items)
    while (nb < B && enc_pass_bytes(e, nb + 1, L) <= workspace_bytes) ++nb;
    for (int b0 = 0; b0 < B && rc == VV_OK; b0 += nb) {
      const int n = std::min(nb, B - b0);
      rc = enc_run_pass(e, reinterpret_cast<const uint8_t*>(audio) + (size_t)b0 * L * asz, audio_dtype == VV_BF16, L, n, L,
                        mean + (size_t)b0 * T * d->cfg.vae_dim, ws, state, state ? slot_ids + b0 : nullptr, s);
    }
  }
  prof_collect(d, s);
  return rc;
}

// ---- per-kernel entry points of the encoder path (offline form; tests time / check these alone) ----------------------
extern "C" int32_t vv_kernel_enc_stem(vv_encoder* e, const void* audio, int32_t audio_dtype, float* x, int32_t B,
                                      int64_t L, void* cuda_stream) {
  if (!e || !e->core.finalized) { set_error("encoder not finalized"); return VV_ERR_NOT_FINALIZED; }
  if (!audio || !x || B < 1 || L < 1 || L >= (1LL << 31)) { set_error("bad argument"); return VV_ERR_INVALID_ARG; }
  if (audio_dtype != VV_F32 && audio_dtype != VV_BF16) { set_error("audio dtype must be f32 or bf16"); return VV_ERR_BAD_DTYPE; }
  VV_CUDA(cudaSetDevice(e->core.cfg.device));
  cudaStream_t s = reinterpret_cast<cudaStream_t>(cuda_stream);
  EncStemArgs p;
  p.audio = audio; p.audio_is_bf16 = audio_dtype == VV_BF16; p.audio_bstride = L;
  p.w = e->stem_w; p.b = e->stem_b; p.x = x; p.nbatch = B; p.L = (int)L;
  p.st = make_state_ref(nullptr, nullptr, 0);
  const long long threads = (L + ENC_STEM_ROWS - 1) / ENC_STEM_ROWS * B * 8;
  launch_k(enc_stem_kernel, dim3((unsigned)std::min<long long>((threads + 255) / 256, 148LL * 64)), dim3(256), 0, s, p);
  e->core.launches++;
  VV_LAUNCH_OK();
  return VV_OK;
}

extern "C" int32_t vv_kernel_downsample(vv_encoder* e, int32_t stage, void* xb, float* y, int32_t B, int32_t rows,
                                        void* cuda_stream) {
  if (!e || !e->core.finalized) { set_error("encoder not finalized"); return VV_ERR_NOT_FINALIZED; }
  vv_decoder* d = &e->core;
  if (stage < 1 || stage >= (int)d->stages.size()) { set_error("stage %d has no strided conv", stage); return VV_ERR_INVALID_ARG; }
  if (!xb || !y || B < 1 || rows < 1) { set_error("bad argument"); return VV_ERR_INVALID_ARG; }
  VV_CUDA(cudaSetDevice(d->cfg.device));
  cudaStream_t s = reinterpret_cast<cudaStream_t>(cuda_stream);
  const StageW& sw = d->stages[stage];
  const int r = sw.ratio, Cin = sw.C / 2;
  const int rows_out = (rows + r - 1) / r;
  EncContextArgs cp;                                   // zero the r context rows and the ragged tail
  cp.xb = reinterpret_cast<__nv_bfloat16*>(xb); cp.nbatch = B; cp.rows = rows; cp.off = r; cp.C = Cin;
  cp.rpb = (long long)(rows_out + 1) * r;
  cp.st = make_state_ref(nullptr, nullptr, 0);
  const long long cells = (cp.rpb - rows) * B * (Cin / 8);
  launch_k(enc_context_kernel, dim3((unsigned)std::max<long long>(1, std::min<long long>((cells + 255) / 256, 148LL * 16))), dim3(256), 0, s, cp);
  d->launches++;
  VV_LAUNCH_OK();
  EpiOut eo;
  eo.epi = EPI_BIAS_F32; eo.out_f32 = y;
  return launch_gemm(d, reinterpret_cast<const __nv_bfloat16*>(xb), rows_out + 1, rows_out, B, sw.up, eo, s);
}

extern "C" int32_t vv_kernel_enc_head(vv_encoder* e, void* xb, float* mean, int32_t B, int32_t rows, void* cuda_stream) {
  if (!e || !e->core.finalized) { set_error("encoder not finalized"); return VV_ERR_NOT_FINALIZED; }
  vv_decoder* d = &e->core;
  if (!xb || !mean || B < 1 || rows < 1) { set_error("bad argument"); return VV_ERR_INVALID_ARG; }
  VV_CUDA(cudaSetDevice(d->cfg.device));
  cudaStream_t s = reinterpret_cast<cudaStream_t>(cuda_stream);
  const int C = d->stages.back().C;
  EncContextArgs cp;                                   // zero the 6 context rows
  cp.xb = reinterpret_cast<__nv_bfloat16*>(xb); cp.nbatch = B; cp.rows = rows; cp.off = 6; cp.C = C; cp.rpb = rows + 6;
  cp.st = make_state_ref(nullptr, nullptr, 0);
  const long long cells = 6LL * B * (C / 8);
  launch_k(enc_context_kernel, dim3((unsigned)std::max<long long>(1, std::min<long long>((cells + 255) / 256, 148LL * 16))), dim3(256), 0, s, cp);
  d->launches++;
  VV_LAUNCH_OK();
  EpiOut eo;
  eo.epi = EPI_BIAS_F32; eo.out_f32 = mean;
  return launch_gemm(d, reinterpret_cast<const __nv_bfloat16*>(xb), rows + 6, rows, B, e->head, eo, s);
}
