"""CPU restatement of the reference's PCM16 output step.  TEST INFRASTRUCTURE ONLY (tests/, smoke(), bench.py's
cpu_baseline leg) — the product path never imports this.

Reference: ``convert_to_16_bit_wav`` (main.py:2956-2979; the same function at demo/gradio_demo.py:1205), applied to each
streamed chunk of one sample (main.py:937-951) after ``bfloat16 -> float32`` (main.py:939-942).
Parity: PINNED — tests/golden/pcm16_*.npz hold outputs of the reference's own function (oracle/gen_golden_pcm16.py runs
its source, lifted out of main.py with ``ast`` because main.py itself imports gradio), and tests/test_oracle_pcm16.py
checks this restatement against them bit for bit.
"""
from __future__ import annotations

import numpy as np


def pcm16_row(x: np.ndarray) -> np.ndarray:
    """One chunk / utterance: main.py:2972-2978.  fp32 in, int16 out."""
    x = np.asarray(x, dtype=np.float32)
    peak = np.max(np.abs(x)) if x.size else np.float32(0)
    if peak > 1.0:                                   # :2973-2974  "Normalize to range [-1, 1] if it's not already"
        x = x / peak
    return (x * 32767).astype(np.int16)              # :2977  fp32 multiply, C cast (truncation toward zero)


def pcm16_rows(x: np.ndarray) -> tuple[np.ndarray, np.ndarray]:
    """[B, L] -> (int16 [B, L], peaks fp32 [B]); every row converted on its own, as the streamer hands out one chunk per
    sample (vibevoice/modular/streamer.py:50-55)."""
    x = np.asarray(x, dtype=np.float32)
    out = np.stack([pcm16_row(r) for r in x]) if x.shape[0] else np.zeros(x.shape, np.int16)
    peaks = np.abs(x).max(axis=1).astype(np.float32) if x.size else np.zeros((x.shape[0],), np.float32)
    return out, peaks


def make_wave(B: int, L: int, seed: int, scale: float, bf16: bool = False) -> np.ndarray:
    """Seeded test waveform [B, L] fp32: row b is N(0, (scale * (b + 1) / B)^2) with a few exact edge values planted
    (0, +-1, values just past +-1, a x.5/32767 tie) so both branches and the truncation are exercised."""
    rng = np.random.default_rng(seed)
    x = rng.standard_normal((B, L)).astype(np.float32)
    x *= (scale * (np.arange(B, dtype=np.float32) + 1) / B)[:, None]
    edge = np.array([0.0, 1.0, -1.0, 1.0000001, -1.0000001, 0.5 / 32767, -0.5 / 32767, 16383.5 / 32767, 0.99999994],
                    dtype=np.float32)
    n = min(L, edge.size)
    x[0, :n] = edge[:n]
    if bf16:                                          # bf16-representable values (what a bf16 decode() returns)
        u = x.view(np.uint32)
        u = ((u + 0x7FFF + ((u >> 16) & 1)) & 0xFFFF0000).astype(np.uint32)
        x = u.view(np.float32)
    return x
