"""Host-side mirror of the reference's decode interface over the C ABI (include/vvdec.h).

Mirrors, by name and argument meaning:
  * ``VibeVoiceAcousticTokenizerModel.decode(latents, cache=None, sample_indices=None,
    use_cache=False, debug=False)``          — vibevoice/modular/modular_vibevoice_tokenizer.py:1099-1108
  * ``VibeVoiceTokenizerStreamingCache``      — :193-256 (``set_to_zero``, ``clear``)
PyTorch is used for device memory, streams and nothing else; all arithmetic runs in the CUDA library.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Dict, Iterable, Optional, Sequence, Tuple

import torch

from . import _capi


@dataclass(frozen=True)
class Geometry:
    """The fields of VibeVoiceAcousticTokenizerConfig that fix shapes on the decode path
    (configuration_vibevoice.py:16-44; parsed at modular_vibevoice_tokenizer.py:1011-1058)."""
    vae_dim: int = 64
    n_filters: int = 32
    ratios: Tuple[int, ...] = (8, 5, 5, 4, 2, 2)
    depths: Tuple[int, ...] = (8, 3, 3, 3, 3, 3, 3)      # decoder order (already reversed, :1028)
    kernel_size: int = 7
    eps: float = 1e-5

    @property
    def hop(self) -> int:
        h = 1
        for r in self.ratios:
            h *= r
        return h

    def stage_channels(self, i: int) -> int:
        return self.n_filters * (2 ** (len(self.depths) - 1 - i))

    @staticmethod
    def _check_supported(cfg) -> None:
        """Reject every config variant outside the path with ValueError, like the reference's own
        constructors do (:649, :728).  A missing attribute means the reference default, which is the supported value."""
        def want(name, ok):
            v = getattr(cfg, name, ok[0])
            if v not in ok:
                raise ValueError(f"unsupported tokenizer config: {name}={v!r} (this path implements {ok})")
        want("causal", (True,))
        want("mixer_layer", ("depthwise_conv",))
        want("conv_norm", ("none",))
        want("pad_mode", ("constant",))
        want("layernorm", ("RMSNorm",))
        want("conv_bias", (True,))
        want("disable_last_norm", (True,))
        want("channels", (1,))
        # Block1D builds its norms affine whatever this says (:629-630), the encoder / decoder final norm is Identity
        # (disable_last_norm): the flag has no effect on the path, but a config that sets it False announces weights
        # without norm scales, which the packer would then miss -- reject early with the reason
        want("layernorm_elementwise_affine", (True,))

    @staticmethod
    def _ints(v):
        return [int(d) for d in v.split("-")] if isinstance(v, str) else [int(d) for d in v]

    @classmethod
    def from_config(cls, cfg) -> "Geometry":
        """DECODER geometry from the reference's VibeVoiceAcousticTokenizerConfig (configuration_vibevoice.py:16-44),
        the way VibeVoiceAcousticTokenizerModel.__init__ derives it (:1019-1045): decoder depths = the encoder's
        reversed unless decoder_depths is given; ratios / n_filters from the decoder_* fields."""
        cls._check_supported(cfg)
        dd = getattr(cfg, "decoder_depths", None)
        depths = cls._ints(dd) if dd is not None else list(reversed(cls._ints(cfg.encoder_depths)))   # :1024-1028
        ratios = getattr(cfg, "decoder_ratios", None) or cfg.encoder_ratios
        nf = getattr(cfg, "decoder_n_filters", None) or cfg.encoder_n_filters
        return cls(vae_dim=int(cfg.vae_dim), n_filters=int(nf), ratios=tuple(int(r) for r in ratios),
                   depths=tuple(depths), kernel_size=7, eps=float(cfg.layernorm_eps))

    @classmethod
    def from_encoder_config(cls, cfg) -> "Geometry":
        """ENCODER geometry (acoustic or semantic tokenizer config), from the encoder_* fields only (:1030-1037,
        :1130-1150), expressed in the decoder order this struct uses: TokenizerEncoder reverses encoder_ratios (:701) and
        walks encoder_depths as given (:702), so the mirror decoder has ratios = encoder_ratios and depths = reversed."""
        cls._check_supported(cfg)
        depths = list(reversed(cls._ints(cfg.encoder_depths)))
        return cls(vae_dim=int(cfg.vae_dim), n_filters=int(cfg.encoder_n_filters),
                   ratios=tuple(int(r) for r in cfg.encoder_ratios), depths=tuple(depths), kernel_size=7,
                   eps=float(cfg.layernorm_eps))


def _dtype_code(dt: torch.dtype) -> int:
    if dt == torch.float32:
        return _capi.VV_F32
    if dt == torch.bfloat16:
        return _capi.VV_BF16
    raise ValueError(f"dtype {dt} not supported on this path (float32 or bfloat16)")


class B200StreamingCache:
    """Device-resident per-stream conv state (replaces VibeVoiceTokenizerStreamingCache, :193-256).  ``decoder`` is a
    B200AcousticDecoder or a B200AcousticEncoder (each keeps its own state: the reference keys its dict by layer id).

    One slot per ``sample_idx``; every stateful layer's causal context lives in one allocation and is
    updated in place by the decode kernels.  ``set_to_zero`` / ``clear`` keep the reference's meaning.
    """

    def __init__(self, decoder: "B200AcousticDecoder", n_slots: int = 256):
        self._dec = decoder
        self._lib = _capi.load()
        h = C.c_void_p()
        create = getattr(self._lib, getattr(decoder, "_STATE_CREATE", "vv_state_create"))   # encoder: vv_encoder_state_create
        _capi.check(create(decoder._h, int(n_slots), C.byref(h)), "vv_state_create")
        self._h = h
        self.n_slots = int(n_slots)
        self.seen = set()            # sample indices that own live state (host view, for the strict quirk)

    # -- reference API -----------------------------------------------------------------------------
    def set_to_zero(self, sample_indices):
        """:234-241 — zero the listed streams' state, keeping them allocated."""
        idx = self._dec._indices_tensor(sample_indices)
        if idx.numel() == 0:
            return
        _capi.check(self._lib.vv_state_zero(self._h, C.c_void_p(idx.data_ptr()), int(idx.numel()),
                                            C.c_void_p(torch.cuda.current_stream(idx.device).cuda_stream)),
                    "vv_state_zero")
        self._keep = idx             # keep the index tensor alive until the stream has consumed it

    def clear(self, layer_id=None, sample_indices=None):
        """:243-256 — forget state.  Per-layer clearing has no meaning for one fused state; a layer_id
        clears the listed (or all) streams entirely, which is what every reference caller does."""
        if sample_indices is None:
            _capi.check(self._lib.vv_state_zero(self._h, None, -1,
                                                C.c_void_p(torch.cuda.current_stream(self._dec.device).cuda_stream)),
                        "vv_state_zero")
            self.seen.clear()
        else:
            self.set_to_zero(sample_indices)
            for i in self._dec._indices_list(sample_indices):
                self.seen.discard(i)

    def check(self) -> None:
        """Raise IndexError if an earlier call used a sample index outside [0, n_slots) (indices given as CUDA tensors are
        consumed on the device without a host sync, so the call that carried the bad index could not fail itself; that
        row was decoded with a zero context and saved no state).  Synchronise first to cover calls still in flight."""
        _capi.check(self._lib.vv_state_check(self._h), "vv_state_check")

    # -- checkpoint --------------------------------------------------------------------------------
    def export_state(self, sample_idx: int) -> bytes:
        n = self._lib.vv_state_bytes_per_slot(self._h)
        buf = (C.c_uint8 * n)()
        _capi.check(self._lib.vv_state_export(self._h, int(sample_idx), buf, n), "vv_state_export")
        return bytes(buf)

    def import_state(self, sample_idx: int, blob: bytes):
        n = self._lib.vv_state_bytes_per_slot(self._h)
        buf = (C.c_uint8 * n).from_buffer_copy(blob)
        _capi.check(self._lib.vv_state_import(self._h, int(sample_idx), buf, n), "vv_state_import")
        self.seen.add(int(sample_idx))

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                self._lib.vv_state_destroy(self._h)
                self._h = None
        except Exception:
            pass


class B200AcousticDecoder:
    """Drop-in for the decode half of VibeVoiceAcousticTokenizerModel on one B200."""

    def __init__(self, geometry: Geometry = Geometry(), device=None, gemm_path: int = 0,
                 max_workspace_bytes: int = 8 << 30, strict_reference_quirks: bool = False,
                 default_slots: int = 256):
        self._lib = _capi.load()
        if not torch.cuda.is_available():
            raise RuntimeError("vvdec needs a CUDA device (sm_100a); there is no CPU fallback")
        self.device = torch.device("cuda", torch.cuda.current_device() if device is None else
                                   (device if isinstance(device, int) else torch.device(device).index or 0))
        self.geometry = geometry
        cfg = _capi.VvConfig()
        cfg.vae_dim, cfg.n_filters, cfg.n_stages = geometry.vae_dim, geometry.n_filters, len(geometry.depths)
        if len(geometry.ratios) != len(geometry.depths) - 1 or len(geometry.depths) > _capi.VV_MAX_STAGES:
            raise ValueError("geometry: len(ratios) must be len(depths) - 1 and at most 7")
        for i, r in enumerate(geometry.ratios):
            cfg.ratios[i] = r
        for i, d in enumerate(geometry.depths):
            cfg.depths[i] = d
        cfg.kernel_size, cfg.eps, cfg.device, cfg.gemm_path = geometry.kernel_size, geometry.eps, self.device.index, gemm_path
        h = C.c_void_p()
        _capi.check(self._lib.vv_decoder_create(C.byref(cfg), C.byref(h)), "vv_decoder_create")
        self._h = h
        self.max_workspace_bytes = int(max_workspace_bytes)
        self.strict_reference_quirks = bool(strict_reference_quirks)
        self.default_slots = int(default_slots)
        self._ws: Optional[torch.Tensor] = None
        self.vae_dim = geometry.vae_dim
        self.hop = geometry.hop

    # -- weights -----------------------------------------------------------------------------------
    def load_state_dict(self, sd: Dict[str, "torch.Tensor"], prefix: str = "") -> "B200AcousticDecoder":
        """``sd`` uses the reference's names (``decoder.upsample_layers.0.0.conv.conv.weight`` ...,
        SURVEY.md App. B), optionally under ``prefix`` (e.g. ``model.acoustic_tokenizer.``).  Values may be
        torch tensors (any device) or numpy arrays."""
        n = 0
        for k, v in sd.items():
            if not k.startswith(prefix + "decoder."):
                continue
            t = torch.as_tensor(v)
            if t.dtype not in (torch.float32, torch.bfloat16):
                t = t.float()
            t = t.detach().contiguous()
            shape = (C.c_int64 * t.dim())(*t.shape)
            _capi.check(self._lib.vv_decoder_load_tensor(self._h, k[len(prefix):].encode(), C.c_void_p(t.data_ptr()),
                                                         _dtype_code(t.dtype), shape, t.dim(), 1 if t.is_cuda else 0),
                        f"vv_decoder_load_tensor({k})")
            n += 1
        if n == 0:
            raise ValueError(f"no '{prefix}decoder.*' tensors in the state dict")
        _capi.check(self._lib.vv_decoder_finalize(self._h), "vv_decoder_finalize")
        return self

    @classmethod
    def from_reference(cls, model, **kw) -> "B200AcousticDecoder":
        """Build from a reference VibeVoiceAcousticTokenizerModel instance (weights are copied)."""
        dec = cls(Geometry.from_config(model.config), **kw)
        return dec.load_state_dict(model.state_dict())

    @property
    def param_count(self) -> int:
        return int(self._lib.vv_decoder_param_count(self._h))

    @property
    def halo_frames(self) -> int:
        return int(self._lib.vv_decoder_halo_frames(self._h))

    def set_profiling(self, enabled: bool) -> None:
        """Per-launch CUDA-event timing inside vv_decode (synchronises each call while enabled)."""
        _capi.check(self._lib.vv_decoder_set_profiling(self._h, 1 if enabled else 0), "vv_decoder_set_profiling")

    def read_profile(self):
        """-> list of dicts {kind, stage, launches, ms}; clears the accumulator."""
        n = self._lib.vv_decoder_read_profile(self._h, None, 0)
        buf = (_capi.VvProfileRecord * max(n, 1))()
        n = min(n, self._lib.vv_decoder_read_profile(self._h, buf, n))
        return [dict(kind=_capi.KERNEL_KINDS[buf[i].kind], stage=int(buf[i].stage), launches=int(buf[i].launches),
                     ms=float(buf[i].ms)) for i in range(n)]

    def take_launch_count(self) -> int:
        return int(self._lib.vv_decoder_take_launch_count(self._h))

    def set_option(self, key: str, value: int) -> None:
        """vv_decoder_set_option: "gemm_bn256" (0 never / 1 auto / 2 always), "gemm_bn256_waves", "gemm_bn64",
        "gemm_bn64_small", "block_kernel", "block_ws", "ffn_stream256" (include/vvdec.h)."""
        _capi.check(self._lib.vv_decoder_set_option(self._h, key.encode(), int(value)), f"vv_decoder_set_option({key})")

    # -- helpers -----------------------------------------------------------------------------------
    def _indices_list(self, sample_indices) -> Sequence[int]:
        if isinstance(sample_indices, torch.Tensor):
            return [int(i) for i in sample_indices.tolist()]
        return [int(i) for i in sample_indices]

    def _indices_tensor(self, sample_indices) -> torch.Tensor:
        """int32 device tensor of slot ids; a CUDA tensor is consumed without a host sync."""
        if isinstance(sample_indices, torch.Tensor):
            return sample_indices.to(device=self.device, dtype=torch.int32).contiguous()
        return torch.tensor(list(sample_indices), dtype=torch.int32, device=self.device)

    def workspace_bytes(self, B: int, T: int) -> int:
        return int(self._lib.vv_decoder_workspace_bytes(self._h, int(B), int(T)))

    def _workspace(self, B: int, T: int) -> torch.Tensor:
        need = min(self.workspace_bytes(B, T), self.max_workspace_bytes)
        if self._ws is None or self._ws.numel() < need:
            self._ws = None
            self._ws = torch.empty(need + 1024, dtype=torch.uint8, device=self.device)
        return self._ws

    def new_cache(self, n_slots: Optional[int] = None) -> B200StreamingCache:
        return B200StreamingCache(self, n_slots or self.default_slots)

    def _state_for(self, cache, batch: int = 0) -> B200StreamingCache:
        """Accept our cache, or ANY object the caller uses as the reference cache (e.g. a genuine
        VibeVoiceTokenizerStreamingCache built at modeling_vibevoice_inference.py:386): the device state hangs
        off it and its set_to_zero / clear are forwarded, so generate() runs unmodified."""
        if isinstance(cache, B200StreamingCache):
            return cache
        st = getattr(cache, "_b200_state", None)
        if st is None or st._dec is not self:
            # generate() passes batch positions as sample_indices (:571), at most the generation batch; the first
            # streaming call may carry only a subset of it, so keep the default as the floor
            st = self.new_cache(max(self.default_slots, int(batch)))
            cache._b200_state = st
            for name in ("set_to_zero", "clear"):
                orig = getattr(cache, name, None)

                def fwd(*a, _orig=orig, _name=name, **k):
                    getattr(st, _name)(*a, **k)
                    return _orig(*a, **k) if _orig is not None else None
                setattr(cache, name, fwd)
        return st

    # -- the reference entry point --------------------------------------------------------------------
    @torch.no_grad()
    def decode(self, latents, cache=None, sample_indices=None, use_cache=False, debug=False, *, workspace=None):
        """Same contract as VibeVoiceAcousticTokenizerModel.decode (:1099-1108).
        ``workspace`` (keyword only, not in the reference): a caller-owned uint8 CUDA tensor used instead of the decoder's
        own scratch -- a CUDA graph that captured this call keeps reading it on every replay (vvdec.FrameStep).

        latents: [B, T, vae_dim] or [B, vae_dim, T] (shape[1] == vae_dim means the latter, :1102-1105),
        float32 or bfloat16, on this decoder's device.  Returns audio [B, 1, T*hop] in the latents' dtype.
        ``debug`` is accepted and ignored, as in the reference's non-debug build of the path.
        """
        if not isinstance(latents, torch.Tensor) or latents.dim() != 3:
            raise ValueError("latents must be a 3-D tensor")
        if not latents.is_cuda:
            raise RuntimeError("latents must live on the CUDA device; vvdec has no CPU path")
        if latents.dtype not in (torch.float32, torch.bfloat16):
            latents = latents.float()
        D = self.vae_dim
        if latents.shape[1] == D:                       # already [B, D, T]
            layout, B, T = _capi.VV_LAYOUT_BDT, latents.shape[0], latents.shape[2]
        elif latents.shape[2] == D:
            layout, B, T = _capi.VV_LAYOUT_BTD, latents.shape[0], latents.shape[1]
        else:
            raise ValueError(f"latents shape {tuple(latents.shape)} has no axis of size vae_dim={D}")
        if B == 0 or T == 0:
            return latents.new_zeros((B, 1, T * self.hop))
        latents = latents.contiguous()
        state_h, slots = None, None
        if use_cache and cache is not None:             # :317, :469
            if sample_indices is None:
                raise AssertionError("sample_indices must be provided for streaming mode")       # :321
            st = self._state_for(cache, B)
            slots = self._indices_tensor(sample_indices)
            if slots.numel() != B:
                raise AssertionError("len(sample_indices) must match batch size")                 # :322-323
            host_view = not (isinstance(sample_indices, torch.Tensor) and sample_indices.is_cuda)
            if self.strict_reference_quirks or host_view:
                idx = self._indices_list(sample_indices)
                if max(idx) >= st.n_slots or min(idx) < 0:
                    raise IndexError(f"sample index out of range for a cache of {st.n_slots} slots")
                if self.strict_reference_quirks and any(i not in st.seen for i in idx):
                    st.set_to_zero(slots)               # all-or-nothing get(): everyone restarts (:204-207)
                st.seen.update(idx)
            state_h = st._h
        out = torch.empty((B, 1, T * self.hop), dtype=latents.dtype, device=self.device)
        ws = workspace if workspace is not None else self._workspace(B, T)
        ws_ptr = (ws.data_ptr() + 1023) & ~1023
        ws_bytes = ws.numel() - (ws_ptr - ws.data_ptr())
        stream = torch.cuda.current_stream(self.device).cuda_stream
        rc = self._lib.vv_decode(self._h, C.c_void_p(latents.data_ptr()), _dtype_code(latents.dtype), layout, int(B), int(T),
                                 C.c_void_p(out.data_ptr()), _dtype_code(out.dtype), C.c_void_p(ws_ptr), int(ws_bytes),
                                 state_h, C.c_void_p(slots.data_ptr()) if slots is not None else None,
                                 C.c_void_p(stream))
        _capi.check(rc, "vv_decode")
        self._keep = (latents, slots)                   # alive until the stream has consumed them
        return out

    __call__ = decode

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                self._lib.vv_decoder_destroy(self._h)
                self._h = None
        except Exception:
            pass


def patch(model, encode: bool = False, **kw) -> B200AcousticDecoder:
    """Swap ``decode`` of a reference tokenizer (or of ``model.model.acoustic_tokenizer`` inside the full
    VibeVoice model) for the B200 path, in place; ``generate()`` (modeling_vibevoice_inference.py:635-641)
    then runs unmodified.  Returns the decoder object.

    ``encode=True`` also swaps ``encode`` (the offline voice-prompt prefill, modeling_vibevoice_inference.py:154)
    for the B200 encoder (``tok._b200_encoder``).  Streaming encode calls (``use_cache=True`` with a cache) must
    feed whole hops (multiples of 3200 samples), which is what ``generate()`` does (:656-662)."""
    tok = model
    for attr in ("model", "acoustic_tokenizer"):
        if hasattr(tok, attr) and not hasattr(tok, "decoder"):
            tok = getattr(tok, attr)
    if not hasattr(tok, "decoder"):
        raise ValueError("could not find an acoustic tokenizer with a .decoder on the given model")
    dev = next(tok.parameters()).device
    dec = B200AcousticDecoder.from_reference(tok, device=dev if dev.type == "cuda" else None, **kw)
    tok._b200_decoder = dec
    tok.decode = dec.decode
    if encode:
        if not hasattr(tok, "encoder"):
            raise ValueError("the tokenizer has no .encoder to replace")
        from .encoder import B200AcousticEncoder
        ekw = {k: v for k, v in kw.items() if k in ("gemm_path",)}
        enc = B200AcousticEncoder.from_reference(tok, device=dev if dev.type == "cuda" else None, **ekw)
        tok._b200_encoder = enc
        tok.encode = enc.encode
    return dec
