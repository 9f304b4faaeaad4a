"""Seeded random-init decoder weights with the reference's state-dict names (there is no checkpoint in
the reference tree and no network).  "stress" init makes every layer load-bearing (the
reference's own layer-scale init of 1e-6 makes the Block1D stack numerically invisible), values are
bf16-representable so that a fp32 reference and this bf16 path hold literally the same weights."""
from __future__ import annotations

import math
from typing import Dict

import torch

from .decoder import Geometry


def tensor_specs(g: Geometry):
    k = g.kernel_size
    n = len(g.depths)
    c0 = g.stage_channels(0)
    yield "decoder.upsample_layers.0.0.conv.conv.weight", (c0, g.vae_dim, k), "w", g.vae_dim * k
    yield "decoder.upsample_layers.0.0.conv.conv.bias", (c0,), "b", 0
    for i in range(n):
        c = g.stage_channels(i)
        if i > 0:
            cin, r = g.stage_channels(i - 1), g.ratios[i - 1]
            p = f"decoder.upsample_layers.{i}.0.convtr.convtr."
            yield p + "weight", (cin, c, 2 * r), "w", 2 * cin
            yield p + "bias", (c,), "b", 0
        for j in range(g.depths[i]):
            p = f"decoder.stages.{i}.{j}."
            yield p + "gamma", (c,), "g", 0
            yield p + "ffn_gamma", (c,), "g", 0
            yield p + "norm.weight", (c,), "g", 0
            yield p + "ffn_norm.weight", (c,), "g", 0
            yield p + "mixer.conv.conv.conv.weight", (c, 1, k), "w", k
            yield p + "mixer.conv.conv.conv.bias", (c,), "b", 0
            yield p + "ffn.linear1.weight", (4 * c, c), "w", c
            yield p + "ffn.linear1.bias", (4 * c,), "b", 0
            yield p + "ffn.linear2.weight", (c, 4 * c), "w", 4 * c
            yield p + "ffn.linear2.bias", (c,), "b", 0
    cl = g.stage_channels(n - 1)
    yield "decoder.head.conv.conv.weight", (1, cl, k), "w", cl * k
    yield "decoder.head.conv.conv.bias", (1,), "b", 0


def encoder_tensor_specs(g: Geometry):
    """The mirror stack (TokenizerEncoder, modular_vibevoice_tokenizer.py:687-813): stage i has n_filters * 2^i
    channels, the decoder's depth n-1-i and a strided conv of the decoder's ratio n-1-i in front."""
    k = g.kernel_size
    n = len(g.depths)
    yield "encoder.downsample_layers.0.0.conv.conv.weight", (g.n_filters, 1, k), "w", k
    yield "encoder.downsample_layers.0.0.conv.conv.bias", (g.n_filters,), "b", 0
    for i in range(n):
        c = g.n_filters * 2 ** i
        if i > 0:
            r = g.ratios[n - 1 - i]
            p = f"encoder.downsample_layers.{i}.0.conv.conv."
            yield p + "weight", (c, c // 2, 2 * r), "w", (c // 2) * 2 * r
            yield p + "bias", (c,), "b", 0
        for j in range(g.depths[n - 1 - i]):
            p = f"encoder.stages.{i}.{j}."
            yield p + "gamma", (c,), "g", 0
            yield p + "ffn_gamma", (c,), "g", 0
            yield p + "norm.weight", (c,), "g", 0
            yield p + "ffn_norm.weight", (c,), "g", 0
            yield p + "mixer.conv.conv.conv.weight", (c, 1, k), "w", k
            yield p + "mixer.conv.conv.conv.bias", (c,), "b", 0
            yield p + "ffn.linear1.weight", (4 * c, c), "w", c
            yield p + "ffn.linear1.bias", (4 * c,), "b", 0
            yield p + "ffn.linear2.weight", (c, 4 * c), "w", 4 * c
            yield p + "ffn.linear2.bias", (c,), "b", 0
    cl = g.n_filters * 2 ** (n - 1)
    yield "encoder.head.conv.conv.weight", (g.vae_dim, cl, k), "w", cl * k
    yield "encoder.head.conv.conv.bias", (g.vae_dim,), "b", 0


def random_encoder_state_dict(g: Geometry = Geometry(), seed: int = 4321) -> Dict[str, torch.Tensor]:
    """Encoder counterpart of random_state_dict (same distributions)."""
    return _draw(encoder_tensor_specs(g), seed)


def random_state_dict(g: Geometry = Geometry(), seed: int = 1234) -> Dict[str, torch.Tensor]:
    """fp32 CPU tensors, bf16-representable: weights ~ N(0, 1/fan_in), biases ~ N(0, 0.1^2),
    layer scales and norm weights ~ U(0.5, 1.5)."""
    return _draw(tensor_specs(g), seed)


def _draw(specs, seed: int) -> Dict[str, torch.Tensor]:
    gen = torch.Generator().manual_seed(seed)
    sd = {}
    for name, shape, kind, fan_in in specs:
        if kind == "w":
            t = torch.randn(shape, generator=gen) / math.sqrt(fan_in)
        elif kind == "b":
            t = torch.randn(shape, generator=gen) * 0.1
        else:
            t = torch.rand(shape, generator=gen) + 0.5
        sd[name] = t.to(torch.bfloat16).float()
    return sd
