"""The reference path written with the SAME torch ATen calls the reference modules make
(``F.pad`` + ``F.conv1d``, ``F.conv_transpose1d`` + right trim, ``F.linear``, ``F.gelu`` (erf), fp32-statistics RMSNorm),
offline form, as plain functions over a state dict with the reference's names.

TEST INFRASTRUCTURE / CPU BASELINE ONLY.  Like ``vv_oracle.py`` this is checker code: ``tests/``, ``smoke()`` and
``bench.py`` (``cpu_baseline``, ``--impl reference``, the bench-side parity samples and the ``torch_eager_gpu`` context
number) may import it; the product package never does.  It exists because the numpy port is ~4x slower than what the
reference really runs on a CPU (PyTorch/ATen with oneDNN + OpenMP): this module IS that call sequence, so timing it on the
box's host cores is the honest "reference CPU decoder" figure, and it is fast enough to check one K2 / K5 / K3-sized
utterance inside ``bench.py``.  ``/root/reference`` itself cannot travel to the GPU box (and ``generate()`` needs
``diffusers``, absent), so the modules are restated, not imported.

Parity pin: ``tests/test_oracle_torch_ref.py`` checks these functions against the committed golden outputs of the
UNMODIFIED reference (tests/golden/*.npz, made by oracle/gen_golden.py) and against the numpy oracle.

Reference lines (vibevoice/modular/modular_vibevoice_tokenizer.py):
  RMSNorm / ConvRMSNorm :65-66, :81-91          SConv1d (non-streaming) :384-418, padding :127-133, :282
  SConvTranspose1d (non-streaming) :551-576     FFN :588-596      Block1D :665-682
  TokenizerDecoder.forward_features :914-951    TokenizerEncoder.forward_features :775-813
"""
from __future__ import annotations

import math
from typing import Dict

import torch
import torch.nn.functional as F


def rms_norm_channels(x: torch.Tensor, weight: torch.Tensor, eps: float) -> torch.Tensor:
    """ConvRMSNorm (:81-91): transpose to [B,T,C], normalise in fp32, cast back, THEN scale by weight (in x's dtype)."""
    xt = x.transpose(1, 2)
    xf = xt.float()
    xn = (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)).to(x.dtype) * weight
    return xn.transpose(1, 2)


def sconv1d(x, w, b, stride: int = 1, groups: int = 1):
    """SConv1d._forward_non_streaming (:384-418): causal left pad of (k-1) - (stride-1) zeros, right pad so that the last
    window is complete (get_extra_padding_for_conv1d :127-133), nn.Conv1d."""
    k = w.shape[-1]
    pt = (k - 1) - (stride - 1)
    L = x.shape[-1]
    n_frames = (L - k + pt) / stride + 1
    ideal = (math.ceil(n_frames) - 1) * stride + (k - pt)
    return F.conv1d(F.pad(x, (pt, ideal - L)), w, b, stride=stride, groups=groups)


def sconv_transpose1d(x, w, b, stride: int):
    """SConvTranspose1d non-streaming (:551-576): nn.ConvTranspose1d then trim k - stride samples, all on the right
    (causal, trim_right_ratio = 1)."""
    y = F.conv_transpose1d(x, w, b, stride=stride)
    return y[..., :y.shape[-1] - (w.shape[-1] - stride)]


def block1d(x, w: Dict[str, torch.Tensor], p: str, eps: float):
    """Block1D.forward (:665-682)."""
    C = x.shape[1]
    h = rms_norm_channels(x, w[p + "norm.weight"], eps)
    h = sconv1d(h, w[p + "mixer.conv.conv.conv.weight"], w[p + "mixer.conv.conv.conv.bias"], groups=C)
    x = x + h * w[p + "gamma"][None, :, None]
    h = rms_norm_channels(x, w[p + "ffn_norm.weight"], eps).transpose(1, 2)
    h = F.linear(F.gelu(F.linear(h, w[p + "ffn.linear1.weight"], w[p + "ffn.linear1.bias"])),
                 w[p + "ffn.linear2.weight"], w[p + "ffn.linear2.bias"]).transpose(1, 2)
    return x + h * w[p + "ffn_gamma"][None, :, None]


@torch.no_grad()
def decode(w: Dict[str, torch.Tensor], geom, latents: torch.Tensor) -> torch.Tensor:
    """VibeVoiceAcousticTokenizerModel.decode (:1099-1108) -> TokenizerDecoder.forward (:914-951), offline.
    latents [B, T, D] (or [B, D, T]: shape[1] == vae_dim) -> audio [B, 1, T*hop].  ``geom`` needs vae_dim, ratios,
    depths, eps (oracle.geometry.DecoderGeometry or vvdec.Geometry)."""
    x = latents if latents.shape[1] == geom.vae_dim else latents.transpose(1, 2)
    x = sconv1d(x, w["decoder.upsample_layers.0.0.conv.conv.weight"], w["decoder.upsample_layers.0.0.conv.conv.bias"])
    for i in range(len(geom.depths)):
        if i > 0:
            p = f"decoder.upsample_layers.{i}.0.convtr.convtr."
            x = sconv_transpose1d(x, w[p + "weight"], w[p + "bias"], geom.ratios[i - 1])
        for j in range(geom.depths[i]):
            x = block1d(x, w, f"decoder.stages.{i}.{j}.", geom.eps)
    return sconv1d(x, w["decoder.head.conv.conv.weight"], w["decoder.head.conv.conv.bias"])


@torch.no_grad()
def encode(w: Dict[str, torch.Tensor], geom, audio: torch.Tensor) -> torch.Tensor:
    """VibeVoiceAcousticTokenizerModel.encode (:1081-1085) -> TokenizerEncoder.forward (:775-813), offline.
    audio [B, 1, L] (or [B, L]) -> mean [B, ceil(L / hop), vae_dim].  ``geom`` is in decoder order (ratios as configured,
    depths reversed), like everywhere in this repo: encoder stage i uses ratio ratios[n-1-i] and depth depths[n-1-i]."""
    x = audio if audio.dim() == 3 else audio[:, None, :]
    n = len(geom.depths)
    for i in range(n):
        p = f"encoder.downsample_layers.{i}.0.conv.conv."
        x = sconv1d(x, w[p + "weight"], w[p + "bias"], stride=1 if i == 0 else geom.ratios[n - 1 - i])
        for j in range(geom.depths[n - 1 - i]):
            x = block1d(x, w, f"encoder.stages.{i}.{j}.", geom.eps)
    x = sconv1d(x, w["encoder.head.conv.conv.weight"], w["encoder.head.conv.conv.bias"])
    return x.transpose(1, 2).contiguous()


def to_weights(sd, device="cpu", dtype=torch.float32) -> Dict[str, torch.Tensor]:
    """state dict (torch tensors or numpy arrays, reference names) -> tensors on ``device`` in ``dtype``."""
    return {k: torch.as_tensor(v).to(device=device, dtype=dtype) for k, v in sd.items()}


def use_all_host_threads() -> int:
    """torchrun exports OMP_NUM_THREADS=1 to every rank; the CPU arm runs on rank 0 alone and must use the box's cores.
    Returns the thread count in effect."""
    import os
    n = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    torch.set_num_threads(max(1, n))
    return torch.get_num_threads()
