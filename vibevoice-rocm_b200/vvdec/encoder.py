"""Host-side mirror of the reference's ENCODE interface over the C ABI (include/vvdec.h, vv_encoder_* / vv_encode).

Mirrors, by name and argument meaning:
  * ``VibeVoiceAcousticTokenizerModel.encode(audio, cache=None, sample_indices=None, use_cache=False, debug=False)``
    — vibevoice/modular/modular_vibevoice_tokenizer.py:1081-1085
  * ``VibeVoiceTokenizerEncoderOutput`` (mean / std, ``sample``, ``kl``, ``mode``) — :954-1000
  * ``VibeVoiceAcousticTokenizerModel.sampling`` — :1087-1097
PyTorch is used for device memory, streams and the random draws of ``sample``; the encoder arithmetic runs in the
CUDA library.  Streaming (``use_cache=True`` with a cache and ``sample_indices``, SConv1d._forward_streaming :327-382)
takes chunks that are whole hops (multiples of 3200 samples — what generate() feeds the semantic tokenizer,
modeling_vibevoice_inference.py:656-662); ragged streaming chunks raise ValueError.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Dict, Optional, Union

import torch

from . import _capi
from .decoder import B200StreamingCache, Geometry, _dtype_code


@dataclass
class B200TokenizerEncoderOutput:
    """Gaussian with fixed std around ``mean`` [B, T, vae_dim] (VibeVoiceTokenizerEncoderOutput, :954-1000)."""
    mean: torch.Tensor
    std: Optional[Union[float, torch.Tensor]] = None

    def sample(self, dist_type: str = "fix"):
        if dist_type == "fix":                                                   # :976-978
            return self.mean + self.std * torch.randn_like(self.mean), self.std
        if dist_type == "gaussian":                                              # :979-988
            value = self.std / 0.8
            std = torch.randn(self.mean.size(0), device=self.mean.device, dtype=self.mean.dtype) * value
            while std.dim() < self.mean.dim():
                std = std.unsqueeze(-1)
            return self.mean + std * torch.randn_like(self.mean), std
        return self.mean, self.std                                               # :989-990

    def kl(self):
        return self.mean * self.mean                                             # mse against zeros, no reduction (:992-995)

    def mode(self):
        return self.mean


class B200AcousticEncoder:
    """Drop-in for the encode half of VibeVoiceAcousticTokenizerModel (and, with vae_dim=128, for
    VibeVoiceSemanticTokenizerModel.encode, :1118-1186 — the same TokenizerEncoder) on one B200."""

    _STATE_CREATE = "vv_encoder_state_create"

    def __init__(self, geometry: Geometry = Geometry(), device=None, gemm_path: int = 0,
                 max_workspace_bytes: int = 24 << 30, fix_std: float = 0.5, std_dist_type: str = "gaussian",
                 default_slots: int = 256, strict_reference_quirks: bool = False):
        self._lib = _capi.load()
        if not torch.cuda.is_available():
            raise RuntimeError("vvdec needs a CUDA device (sm_100a); there is no CPU fallback")
        self.device = torch.device("cuda", torch.cuda.current_device() if device is None else
                                   (device if isinstance(device, int) else torch.device(device).index or 0))
        self.geometry = geometry
        if len(geometry.ratios) != len(geometry.depths) - 1 or len(geometry.depths) > _capi.VV_MAX_STAGES:
            raise ValueError("geometry: len(ratios) must be len(depths) - 1 and at most 7")
        cfg = _capi.VvConfig()
        cfg.vae_dim, cfg.n_filters, cfg.n_stages = geometry.vae_dim, geometry.n_filters, len(geometry.depths)
        for i, r in enumerate(geometry.ratios):
            cfg.ratios[i] = r
        for i, d in enumerate(geometry.depths):
            cfg.depths[i] = d
        cfg.kernel_size, cfg.eps, cfg.device, cfg.gemm_path = geometry.kernel_size, geometry.eps, self.device.index, gemm_path
        h = C.c_void_p()
        _capi.check(self._lib.vv_encoder_create(C.byref(cfg), C.byref(h)), "vv_encoder_create")
        self._h = h
        self.max_workspace_bytes = int(max_workspace_bytes)
        self.fix_std = fix_std                       # config.fix_std (:1014)
        self.std_dist_type = std_dist_type           # config.std_dist_type (:1015)
        self.default_slots = int(default_slots)
        self.strict_reference_quirks = bool(strict_reference_quirks)
        self._ws: Optional[torch.Tensor] = None
        self.vae_dim = geometry.vae_dim
        self.hop = geometry.hop

    # -- weights -----------------------------------------------------------------------------------
    def load_state_dict(self, sd: Dict[str, "torch.Tensor"], prefix: str = "") -> "B200AcousticEncoder":
        """``sd`` uses the reference's names (``encoder.downsample_layers.0.0.conv.conv.weight`` ...), optionally
        under ``prefix``; torch tensors (any device) or numpy arrays."""
        n = 0
        for k, v in sd.items():
            if not k.startswith(prefix + "encoder."):
                continue
            t = torch.as_tensor(v)
            if t.dtype not in (torch.float32, torch.bfloat16):
                t = t.float()
            t = t.detach().contiguous()
            shape = (C.c_int64 * t.dim())(*t.shape)
            _capi.check(self._lib.vv_encoder_load_tensor(self._h, k[len(prefix):].encode(), C.c_void_p(t.data_ptr()),
                                                         _dtype_code(t.dtype), shape, t.dim(), 1 if t.is_cuda else 0),
                        f"vv_encoder_load_tensor({k})")
            n += 1
        if n == 0:
            raise ValueError(f"no '{prefix}encoder.*' tensors in the state dict")
        _capi.check(self._lib.vv_encoder_finalize(self._h), "vv_encoder_finalize")
        return self

    @classmethod
    def from_reference(cls, model, **kw) -> "B200AcousticEncoder":
        """Build from a reference VibeVoiceAcousticTokenizerModel or VibeVoiceSemanticTokenizerModel instance (weights
        are copied).  The semantic model returns ``std=None`` (:1168-1169) and samples with 'none' (:1174)."""
        semantic = not hasattr(model, "fix_std")
        kw.setdefault("fix_std", None if semantic else getattr(model, "fix_std"))
        kw.setdefault("std_dist_type", "none" if semantic else getattr(model, "std_dist_type", "gaussian"))
        enc = cls(Geometry.from_encoder_config(model.config), **kw)
        return enc.load_state_dict(model.state_dict())

    @property
    def param_count(self) -> int:
        return int(self._lib.vv_encoder_param_count(self._h))

    @property
    def halo_frames(self) -> int:
        return int(self._lib.vv_encoder_halo_frames(self._h))

    def frames(self, L: int) -> int:
        return int(self._lib.vv_encoder_frames(self._h, int(L)))

    def workspace_bytes(self, B: int, L: int) -> int:
        return int(self._lib.vv_encoder_workspace_bytes(self._h, int(B), int(L)))

    def set_profiling(self, enabled: bool) -> None:
        _capi.check(self._lib.vv_encoder_set_profiling(self._h, 1 if enabled else 0), "vv_encoder_set_profiling")

    def read_profile(self):
        n = self._lib.vv_encoder_read_profile(self._h, None, 0)
        buf = (_capi.VvProfileRecord * max(n, 1))()
        n = min(n, self._lib.vv_encoder_read_profile(self._h, buf, n))
        return [dict(kind=_capi.KERNEL_KINDS[buf[i].kind], stage=int(buf[i].stage), launches=int(buf[i].launches),
                     ms=float(buf[i].ms)) for i in range(n)]

    def take_launch_count(self) -> int:
        return int(self._lib.vv_encoder_take_launch_count(self._h))

    def set_option(self, key: str, value: int) -> None:
        """vv_encoder_set_option (same keys as the decoder's)."""
        _capi.check(self._lib.vv_encoder_set_option(self._h, key.encode(), int(value)), f"vv_encoder_set_option({key})")

    # -- streaming state ------------------------------------------------------------------------------
    def _indices_list(self, sample_indices):
        if isinstance(sample_indices, torch.Tensor):
            return [int(i) for i in sample_indices.tolist()]
        return [int(i) for i in sample_indices]

    def _indices_tensor(self, sample_indices) -> torch.Tensor:
        if isinstance(sample_indices, torch.Tensor):
            return sample_indices.to(device=self.device, dtype=torch.int32).contiguous()
        return torch.tensor(list(sample_indices), dtype=torch.int32, device=self.device)

    def new_cache(self, n_slots: Optional[int] = None) -> B200StreamingCache:
        return B200StreamingCache(self, n_slots or self.default_slots)

    def _state_for(self, cache, batch: int = 0) -> B200StreamingCache:
        """Our cache, or any object the caller uses as the reference cache (a genuine VibeVoiceTokenizerStreamingCache):
        the encoder's device state hangs off it and its set_to_zero / clear are forwarded (same as the decoder)."""
        if isinstance(cache, B200StreamingCache):
            if cache._dec is not self:
                raise ValueError("this cache belongs to another decoder / encoder")
            return cache
        st = getattr(cache, "_b200_enc_state", None)
        if st is None or st._dec is not self:
            st = self.new_cache(max(self.default_slots, int(batch)))
            cache._b200_enc_state = st
            for name in ("set_to_zero", "clear"):
                orig = getattr(cache, name, None)

                def fwd(*a, _orig=orig, _name=name, **k):
                    getattr(st, _name)(*a, **k)
                    return _orig(*a, **k) if _orig is not None else None
                setattr(cache, name, fwd)
        return st

    def _workspace(self, B: int, L: int) -> torch.Tensor:
        need = min(self.workspace_bytes(B, L), max(self.max_workspace_bytes, self.workspace_bytes(1, L)))
        if self._ws is None or self._ws.numel() < need + 1024:
            self._ws = None
            self._ws = torch.empty(need + 1024, dtype=torch.uint8, device=self.device)
        return self._ws

    # -- the reference entry points -------------------------------------------------------------------
    @torch.no_grad()
    def encode(self, audio, cache=None, sample_indices=None, use_cache=False, debug=False, *, workspace=None):
        """Same contract as VibeVoiceAcousticTokenizerModel.encode (:1081-1085).
        ``workspace`` (keyword only): caller-owned uint8 CUDA scratch instead of the encoder's own (CUDA-graph capture).

        audio: [B, 1, L] (or [B, L]) float32 / bfloat16 on this encoder's device.  Returns
        B200TokenizerEncoderOutput(mean [B, ceil(L / hop), vae_dim] in the audio's dtype, std=fix_std).
        With ``use_cache=True`` and a cache the call continues the streams named by ``sample_indices``
        (L must be a multiple of the hop).
        """
        if not isinstance(audio, torch.Tensor) or audio.dim() not in (2, 3):
            raise ValueError("audio must be a [B, 1, L] or [B, L] tensor")
        if audio.dim() == 3:
            if audio.shape[1] != 1:
                raise ValueError(f"audio has {audio.shape[1]} channels; the tokenizer takes mono (channels=1)")
            audio = audio[:, 0, :]
        if not audio.is_cuda:
            raise RuntimeError("audio must live on the CUDA device; vvdec has no CPU path")
        if audio.dtype not in (torch.float32, torch.bfloat16):
            audio = audio.float()
        B, L = audio.shape
        T = self.frames(L) if L > 0 else 0
        if B == 0 or L == 0:
            return B200TokenizerEncoderOutput(mean=audio.new_zeros((B, T, self.vae_dim)), std=self.fix_std)
        audio = audio.contiguous()
        state_h, slots = None, None
        if use_cache and cache is not None:             # :317
            if sample_indices is None:
                raise AssertionError("sample_indices must be provided for streaming mode")       # :321
            st = self._state_for(cache, B)
            slots = self._indices_tensor(sample_indices)
            if slots.numel() != B:
                raise AssertionError("len(sample_indices) must match batch size")                 # :322-323
            if L % self.hop != 0:
                raise ValueError(f"streaming encode takes whole hops: L={L} is not a multiple of {self.hop}")
            host_view = not (isinstance(sample_indices, torch.Tensor) and sample_indices.is_cuda)
            if self.strict_reference_quirks or host_view:
                idx = self._indices_list(sample_indices)
                if max(idx) >= st.n_slots or min(idx) < 0:
                    raise IndexError(f"sample index out of range for a cache of {st.n_slots} slots")
                if self.strict_reference_quirks and any(i not in st.seen for i in idx):
                    st.set_to_zero(slots)               # all-or-nothing get(): everyone restarts (:204-207)
                st.seen.update(idx)
            state_h = st._h
        mean = torch.empty((B, T, self.vae_dim), dtype=torch.float32, device=self.device)
        ws = workspace if workspace is not None else self._workspace(B, L)
        ws_ptr = (ws.data_ptr() + 1023) & ~1023
        ws_bytes = ws.numel() - (ws_ptr - ws.data_ptr())
        stream = torch.cuda.current_stream(self.device).cuda_stream
        rc = self._lib.vv_encode(self._h, C.c_void_p(audio.data_ptr()), _dtype_code(audio.dtype), int(B), int(L),
                                 C.c_void_p(mean.data_ptr()), C.c_void_p(ws_ptr), int(ws_bytes), state_h,
                                 C.c_void_p(slots.data_ptr()) if slots is not None else None, C.c_void_p(stream))
        _capi.check(rc, "vv_encode")
        self._keep = (audio, slots)                     # alive until the stream has consumed them
        return B200TokenizerEncoderOutput(mean=mean if audio.dtype == torch.float32 else mean.to(audio.dtype),
                                          std=self.fix_std)

    __call__ = encode

    @torch.no_grad()
    def sampling(self, encoder_output, dist_type=None):
        """VibeVoiceAcousticTokenizerModel.sampling (:1087-1097)."""
        if self.std_dist_type == "none":                 # semantic tokenizer (:1172-1174): always the mean
            return encoder_output.sample(dist_type="none")
        dist_type = dist_type or self.std_dist_type
        if dist_type not in ("fix", "gaussian"):
            raise ValueError(f"Unsupported dist_type: {dist_type}, expected 'fix' or 'gaussian'")
        return encoder_output.sample(dist_type=dist_type)

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                self._lib.vv_encoder_destroy(self._h)
                self._h = None
        except Exception:
            pass


def patch_encoder(tokenizer, **kw) -> B200AcousticEncoder:
    """Swap ``encode`` of a reference tokenizer that has an ``.encoder`` (VibeVoiceAcousticTokenizerModel or
    VibeVoiceSemanticTokenizerModel, e.g. ``model.model.semantic_tokenizer``) for the B200 path, in place; the
    streaming calls of ``generate()`` (modeling_vibevoice_inference.py:656-662) then run unmodified as long as they feed
    whole hops.  Returns the encoder object (also kept as ``tokenizer._b200_encoder``)."""
    if not hasattr(tokenizer, "encoder"):
        raise ValueError("the given tokenizer has no .encoder")
    dev = next(tokenizer.parameters()).device
    enc = B200AcousticEncoder.from_reference(tokenizer, device=dev if dev.type == "cuda" else None, **kw)
    tokenizer._b200_encoder = enc
    tokenizer.encode = enc.encode
    return enc

