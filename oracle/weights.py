"""Seeded synthetic decoder weights + latents (TEST INFRASTRUCTURE ONLY — see geometry.py).

There is no checkpoint in the reference tree and no network, so both sides of
every parity test load the SAME synthetic tensors, drawn here with numpy's counter-based
Philox generator (bit-reproducible across machines) and rounded to bf16-representable fp32
(SURVEY.md §8c tolerance protocol: "weights & latents drawn in fp32, rounded to bf16, and
those rounded values loaded into the fp32 reference").

Two inits:
  * ``default`` — the distributions of the reference's own ``_init_weights``
    (modular_vibevoice_tokenizer.py:1067-1079: N(0, 0.01) for Linear/Conv1d, zero bias;
    ConvTranspose1d keeps torch's default U(+-1/sqrt(fan_in)) for weight and bias), norm weights 1,
    gamma = layer_scale_init_value = 1e-6 (:658-660).  With it the Block1D stack is numerically
    invisible, so it is kept only as the literal "random-init reference" case.
  * ``stress``  — gamma, ffn_gamma, norm weights ~ U(0.5, 1.5); biases ~ N(0, 0.1^2);
    conv/linear weights ~ N(0, 1/fan_in) (transposed conv: fan_in = 2*Cin, two taps per phase).
    Every kernel class is load-bearing under this init.
"""
from __future__ import annotations

from typing import Dict

import numpy as np

from .geometry import DecoderGeometry, FULL


def round_to_bf16(x: np.ndarray) -> np.ndarray:
    """Round-to-nearest-even fp32 -> bf16, returned as fp32 (bf16-representable)."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    u = x.view(np.uint32)
    r = (u + np.uint32(0x7FFF) + ((u >> np.uint32(16)) & np.uint32(1))) & np.uint32(0xFFFF0000)
    return r.view(np.float32)


def _rng(seed: int, idx: int) -> np.random.Generator:
    return np.random.Generator(np.random.Philox(key=[np.uint64(seed), np.uint64(idx)]))


def make_weights(geom: DecoderGeometry = FULL, init: str = "stress", seed: int = 1234,
                 bf16_round: bool = True) -> Dict[str, np.ndarray]:
    """name -> fp32 ndarray for every decoder parameter of the reference state-dict."""
    assert init in ("stress", "default", "mild")
    out: Dict[str, np.ndarray] = {}
    for idx, (name, shape, kind) in enumerate(geom.tensor_specs()):
        g = _rng(seed, idx)
        n = int(np.prod(shape))
        if kind in ("conv_w", "linear_w", "convtr_w"):
            if kind == "linear_w":
                fan_in = shape[1]
            elif kind == "conv_w":
                fan_in = shape[1] * shape[2]
            else:  # convtr [Cin, Cout, k]; each output phase sees 2 taps x Cin inputs
                fan_in = 2 * shape[0]
            if init == "default":
                if kind == "convtr_w":
                    bound = 1.0 / np.sqrt(shape[1] * shape[2])      # torch default for ConvTranspose1d
                    w = g.uniform(-bound, bound, size=n).astype(np.float32)
                else:
                    w = g.standard_normal(n, dtype=np.float32) * np.float32(0.01)
            else:
                w = g.standard_normal(n, dtype=np.float32) * np.float32(1.0 / np.sqrt(fan_in))
        elif kind == "bias":
            if init == "default":
                if ".convtr." in name:
                    cin_k = None
                    # bound uses the weight's fan_in = Cout*k; recover k from the ratio schedule
                    i = int(name.split(".")[2])
                    bound = 1.0 / np.sqrt(shape[0] * 2 * geom.ratios[i - 1])
                    w = g.uniform(-bound, bound, size=n).astype(np.float32)
                else:
                    w = np.zeros(n, dtype=np.float32)
            else:
                w = g.standard_normal(n, dtype=np.float32) * np.float32(0.1)
        elif kind == "norm_w":
            w = np.ones(n, dtype=np.float32) if init == "default" else \
                g.uniform(0.5, 1.5, size=n).astype(np.float32)
        elif kind == "gamma":
            if init == "default":
                w = np.full(n, 1e-6, dtype=np.float32)
            elif init == "mild":
                w = g.uniform(0.05, 0.3, size=n).astype(np.float32)
            else:
                w = g.uniform(0.5, 1.5, size=n).astype(np.float32)
        else:  # pragma: no cover
            raise ValueError(kind)
        w = w.reshape(shape)
        out[name] = round_to_bf16(w) if bf16_round else w
    return out


def make_encoder_weights(geom: DecoderGeometry = FULL, init: str = "stress", seed: int = 4321,
                         bf16_round: bool = True) -> Dict[str, np.ndarray]:
    """name -> fp32 ndarray for every ENCODER parameter (same distributions as make_weights)."""
    assert init in ("stress", "default", "mild")
    out: Dict[str, np.ndarray] = {}
    for idx, (name, shape, kind) in enumerate(geom.encoder_tensor_specs()):
        g = _rng(seed, idx)
        n = int(np.prod(shape))
        if kind in ("conv_w", "linear_w"):
            fan_in = shape[1] if kind == "linear_w" else shape[1] * shape[2]
            scale = 0.01 if init == "default" else 1.0 / np.sqrt(fan_in)
            w = g.standard_normal(n, dtype=np.float32) * np.float32(scale)
        elif kind == "bias":
            w = np.zeros(n, dtype=np.float32) if init == "default" else \
                g.standard_normal(n, dtype=np.float32) * np.float32(0.1)
        elif kind == "norm_w":
            w = np.ones(n, dtype=np.float32) if init == "default" else g.uniform(0.5, 1.5, size=n).astype(np.float32)
        else:  # gamma
            w = np.full(n, 1e-6, dtype=np.float32) if init == "default" else \
                (g.uniform(0.05, 0.3, size=n) if init == "mild" else g.uniform(0.5, 1.5, size=n)).astype(np.float32)
        w = w.reshape(shape)
        out[name] = round_to_bf16(w) if bf16_round else w
    return out


def make_audio(B: int, L: int, seed: int = 5, bf16_round: bool = True) -> np.ndarray:
    """[B, 1, L] synthetic audio: 0.1 * N(0,1), lightly low-passed."""
    g = _rng(seed, 1 << 21)
    x = g.standard_normal(B * (L + 4), dtype=np.float32).reshape(B, L + 4) * np.float32(0.1)
    x = (x[:, 0:L] + 2 * x[:, 1:L + 1] + 3 * x[:, 2:L + 2] + 2 * x[:, 3:L + 3] + x[:, 4:L + 4]) / np.float32(9.0)
    x = x.reshape(B, 1, L).astype(np.float32)
    return round_to_bf16(x) if bf16_round else x


def make_latents(B: int, T: int, dim: int = 64, seed: int = 1, bf16_round: bool = True) -> np.ndarray:
    """[B, T, dim] N(0,1) latents (SURVEY.md §8d synthetic inputs)."""
    g = _rng(seed, 1 << 20)
    x = g.standard_normal(B * T * dim, dtype=np.float32).reshape(B, T, dim)
    return round_to_bf16(x) if bf16_round else x
