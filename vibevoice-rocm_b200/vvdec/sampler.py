"""Host-side mirror of the reference's per-frame diffusion sampling over the C ABI (include/vvdec.h, vv_sampler_*):
SURVEY.md 8f N4, the step right before the acoustic decoder.

Mirrors ``VibeVoiceForConditionalGenerationInference.sample_speech_tokens(condition, neg_condition, cfg_scale=3.0)``
(vibevoice/modular/modeling_vibevoice_inference.py:695-708), which runs ``ddpm_num_inference_steps`` passes of
``VibeVoiceDiffusionHead`` (modular_vibevoice_diffusion_head.py:254-280) on the conditioned + unconditioned rows, applies
classifier-free guidance and advances ``DPMSolverMultistepScheduler`` (vibevoice/schedule/dpm_solver.py:935-1022).
PyTorch is used for device memory, streams and the initial ``torch.randn`` draw (:698); the arithmetic runs in the CUDA
library.  No CPU / PyTorch fallback.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Dict, Optional

import torch

from . import _capi
from .decoder import _dtype_code


@dataclass(frozen=True)
class SamplerGeometry:
    """The fields of VibeVoiceDiffusionHeadConfig that fix shapes (configuration_vibevoice.py:129-160)."""
    hidden: int = 1536
    layers: int = 4
    ffn_ratio: float = 3.0
    latent: int = 64
    eps: float = 1e-5
    train_steps: int = 1000
    inference_steps: int = 20

    @property
    def ffn(self) -> int:
        return int(self.hidden * self.ffn_ratio)                       # :228

    @classmethod
    def from_config(cls, cfg) -> "SamplerGeometry":
        """Accepts the reference's VibeVoiceDiffusionHeadConfig; rejects what the path does not implement."""
        if getattr(cfg, "prediction_type", "v_prediction") != "v_prediction":
            raise ValueError(f"unsupported prediction_type {cfg.prediction_type!r} (this path implements 'v_prediction')")
        if getattr(cfg, "ddpm_beta_schedule", "cosine") not in ("cosine", "squaredcos_cap_v2"):
            raise ValueError(f"unsupported ddpm_beta_schedule {cfg.ddpm_beta_schedule!r} (this path implements 'cosine')")
        return cls(hidden=int(cfg.hidden_size), layers=int(cfg.head_layers), ffn_ratio=float(cfg.head_ffn_ratio),
                   latent=int(cfg.latent_size), eps=float(cfg.rms_norm_eps), train_steps=int(cfg.ddpm_num_steps),
                   inference_steps=int(cfg.ddpm_num_inference_steps))


class B200DiffusionSampler:
    """Drop-in for ``sample_speech_tokens`` on one B200."""

    def __init__(self, geometry: SamplerGeometry = SamplerGeometry(), device=None):
        self._lib = _capi.load()
        if not torch.cuda.is_available():
            raise RuntimeError("vvdec needs a CUDA device (sm_100a); there is no CPU fallback")
        self.device = torch.device("cuda", torch.cuda.current_device() if device is None else
                                   (device if isinstance(device, int) else torch.device(device).index or 0))
        self.geometry = geometry
        cfg = _capi.VvSamplerConfig(hidden=geometry.hidden, layers=geometry.layers, ffn=geometry.ffn, latent=geometry.latent,
                                    eps=geometry.eps, train_steps=geometry.train_steps, device=self.device.index)
        h = C.c_void_p()
        _capi.check(self._lib.vv_sampler_create(C.byref(cfg), C.byref(h)), "vv_sampler_create")
        self._h = h
        self.ddpm_inference_steps = geometry.inference_steps            # modeling_vibevoice_inference.py:82
        self._ws: Optional[torch.Tensor] = None

    # -- weights -----------------------------------------------------------------------------------
    def load_state_dict(self, sd: Dict[str, "torch.Tensor"], prefix: str = "") -> "B200DiffusionSampler":
        """``sd`` uses the head's own names (``noisy_images_proj.weight``, ``layers.0.ffn.gate_proj.weight`` ...), optionally
        under ``prefix`` (``model.prediction_head.`` in the full model); torch tensors (any device) or numpy arrays."""
        n = 0
        for k, v in sd.items():
            if not k.startswith(prefix):
                continue
            t = torch.as_tensor(v)
            if t.dtype not in (torch.float32, torch.bfloat16):
                t = t.float()
            t = t.detach().contiguous()
            shape = (C.c_int64 * t.dim())(*t.shape)
            _capi.check(self._lib.vv_sampler_load_tensor(self._h, k[len(prefix):].encode(), C.c_void_p(t.data_ptr()),
                                                         _dtype_code(t.dtype), shape, t.dim(), 1 if t.is_cuda else 0),
                        f"vv_sampler_load_tensor({k})")
            n += 1
        if n == 0:
            raise ValueError(f"no '{prefix}*' tensors in the state dict")
        _capi.check(self._lib.vv_sampler_finalize(self._h), "vv_sampler_finalize")
        return self

    @classmethod
    def from_reference(cls, prediction_head, **kw) -> "B200DiffusionSampler":
        """Build from a reference VibeVoiceDiffusionHead instance (``model.model.prediction_head``); weights are copied."""
        return cls(SamplerGeometry.from_config(prediction_head.config), **kw).load_state_dict(prediction_head.state_dict())

    @property
    def param_count(self) -> int:
        return int(self._lib.vv_sampler_param_count(self._h))

    def take_launch_count(self) -> int:
        return int(self._lib.vv_sampler_take_launch_count(self._h))

    def set_ddpm_inference_steps(self, num_steps=None) -> None:
        """modeling_vibevoice_inference.py:146-147."""
        self.ddpm_inference_steps = int(num_steps or self.geometry.inference_steps)

    def prepare(self, steps: Optional[int] = None) -> None:
        """Build (once) the solver table and timestep embeddings of a step count; call before capturing a CUDA graph."""
        _capi.check(self._lib.vv_sampler_prepare(self._h, int(steps or self.ddpm_inference_steps)), "vv_sampler_prepare")

    def _workspace(self, n: int) -> torch.Tensor:
        need = int(self._lib.vv_sampler_workspace_bytes(self._h, int(n)))
        if self._ws is None or self._ws.numel() < need + 1024:
            self._ws = None
            self._ws = torch.empty(need + 1024, dtype=torch.uint8, device=self.device)
        return self._ws

    def _ws_args(self, ws):
        ptr = (ws.data_ptr() + 1023) & ~1023
        return C.c_void_p(ptr), int(ws.numel() - (ptr - ws.data_ptr()))

    # -- the reference entry point -----------------------------------------------------------------------
    @torch.no_grad()
    def sample_speech_tokens(self, condition, neg_condition, cfg_scale=3.0, *, noise=None, workspace=None):
        """Same contract as the reference method (:695-708): condition, neg_condition [n, hidden] on this device ->
        speech latents [n, latent] in the conditions' dtype.  ``noise`` (keyword only, not in the reference): the initial
        latent [n, latent]; default = ``torch.randn`` on the device, the draw the reference makes at :698 (of whose 2n rows
        only the first n ever reach the result)."""
        if not (isinstance(condition, torch.Tensor) and isinstance(neg_condition, torch.Tensor)):
            raise ValueError("condition and neg_condition must be tensors")
        if not (condition.is_cuda and neg_condition.is_cuda):
            raise RuntimeError("conditions must live on the CUDA device; vvdec has no CPU path")
        g = self.geometry
        if condition.dim() != 2 or condition.shape[1] != g.hidden or neg_condition.shape != condition.shape:
            raise ValueError(f"condition / neg_condition must both be [n, {g.hidden}]")
        n = int(condition.shape[0])
        dt = condition.dtype
        if dt not in (torch.float32, torch.bfloat16):
            condition, neg_condition = condition.float(), neg_condition.float()
        neg_condition = neg_condition.to(condition.dtype)
        if n == 0:
            return condition.new_zeros((0, g.latent))
        condition, neg_condition = condition.contiguous(), neg_condition.contiguous()
        if noise is None:
            noise = torch.randn(n, g.latent, device=self.device, dtype=torch.float32)
        noise = noise.to(device=self.device, dtype=torch.float32).contiguous()
        if noise.shape != (n, g.latent):
            raise ValueError(f"noise must be [{n}, {g.latent}]")
        out = torch.empty((n, g.latent), dtype=torch.float32, device=self.device)
        ws_ptr, ws_bytes = self._ws_args(workspace if workspace is not None else self._workspace(n))
        stream = torch.cuda.current_stream(self.device).cuda_stream
        rc = self._lib.vv_sampler_sample(self._h, C.c_void_p(condition.data_ptr()), C.c_void_p(neg_condition.data_ptr()),
                                         _dtype_code(condition.dtype), n, C.c_void_p(noise.data_ptr()), C.c_float(float(cfg_scale)),
                                         int(self.ddpm_inference_steps), C.c_void_p(out.data_ptr()), ws_ptr, ws_bytes,
                                         C.c_void_p(stream))
        _capi.check(rc, "vv_sampler_sample")
        self._keep = (condition, neg_condition, noise)
        return out if dt == torch.float32 else out.to(dt)

    @torch.no_grad()
    def head(self, noisy, timestep: float, condition):
        """One ``VibeVoiceDiffusionHead.forward(noisy, t, condition)`` (:254-280) on an even number of rows (test hook)."""
        R = int(noisy.shape[0])
        noisy = noisy.to(device=self.device, dtype=torch.float32).contiguous()
        condition = condition.to(self.device).contiguous()
        if condition.dtype not in (torch.float32, torch.bfloat16):
            condition = condition.float()
        out = torch.empty((R, self.geometry.latent), dtype=torch.float32, device=self.device)
        ws_ptr, ws_bytes = self._ws_args(self._workspace(max(1, R // 2)))
        rc = self._lib.vv_sampler_head(self._h, C.c_void_p(noisy.data_ptr()), C.c_float(float(timestep)),
                                       C.c_void_p(condition.data_ptr()), _dtype_code(condition.dtype), R,
                                       C.c_void_p(out.data_ptr()), ws_ptr, ws_bytes,
                                       C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream))
        _capi.check(rc, "vv_sampler_head")
        return out

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                self._lib.vv_sampler_destroy(self._h)
                self._h = None
        except Exception:
            pass


def patch_sampler(model, **kw) -> B200DiffusionSampler:
    """Swap ``sample_speech_tokens`` of a reference ``VibeVoiceForConditionalGenerationInference`` for the B200 path, in place
    (``generate()`` calls it at modeling_vibevoice_inference.py:627).  Returns the sampler (also ``model._b200_sampler``)."""
    head = model.model.prediction_head
    dev = next(head.parameters()).device
    smp = B200DiffusionSampler.from_reference(head, device=dev if dev.type == "cuda" else None, **kw)
    smp.ddpm_inference_steps = int(getattr(model, "ddpm_inference_steps", smp.ddpm_inference_steps))
    model._b200_sampler = smp
    model.sample_speech_tokens = smp.sample_speech_tokens
    return smp
