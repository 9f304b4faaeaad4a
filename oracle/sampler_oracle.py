"""CPU restatement (numpy) of the reference's per-frame diffusion sampling: the step right BEFORE the
acoustic decoder -- it produces the 64-d latent the decoder consumes.

TEST INFRASTRUCTURE ONLY (same rule as vv_oracle.py: tests/, smoke() and bench.py's CPU legs may import it, the product
package never does).  Pinned against outputs of the UNMODIFIED reference classes run in the build container
(oracle/gen_golden_sampler.py through oracle/ref_shim.import_reference_sampler; fixtures tests/golden/samp_*.npz).

What it restates, with the reference lines:
  * VibeVoiceForConditionalGenerationInference.sample_speech_tokens   modeling_vibevoice_inference.py:695-708
  * VibeVoiceDiffusionHead.forward and its layers                     modular_vibevoice_diffusion_head.py:20-45 (RMSNorm,
        modulate), :48-93 (TimestepEmbedder), :96-124 (SwiGLU FFN), :127-161 (HeadLayer), :164-190 (FinalLayer), :254-280
  * DPMSolverMultistepScheduler (a copy of diffusers' scheduler)      vibevoice/schedule/dpm_solver.py:28-85 (cosine betas),
        :228-296 (tables), :321-424 (set_timesteps), :483-487, :528-590 (v-prediction -> x0), :627-695 (first order),
        :696-818 (second order, midpoint), :935-1022 (step)
    with the configuration the model builds it with (modeling_vibevoice.py:138-142: num_train_timesteps = ddpm_num_steps,
    beta_schedule 'cosine', prediction_type 'v_prediction'; every other argument at its default: solver_order 2,
    'dpmsolver++', 'midpoint', lower_order_final, final_sigmas_type 'zero', timestep_spacing 'linspace').
The random draw of the initial noise (torch.randn, :698) is an INPUT here: parity is defined for a given noise tensor.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Dict, List, Tuple

import numpy as np

F32 = np.float32


@dataclass(frozen=True)
class SamplerGeometry:
    """VibeVoiceDiffusionHeadConfig (configuration_vibevoice.py:129-160); defaults = the 1.5B model's JSON."""
    hidden: int = 1536
    layers: int = 4
    ffn_ratio: float = 3.0
    latent: int = 64
    eps: float = 1e-5
    train_steps: int = 1000
    freq_dim: int = 256               # TimestepEmbedder.frequency_embedding_size (:56)

    @property
    def ffn(self) -> int:
        return int(self.hidden * self.ffn_ratio)

    def tensor_specs(self) -> List[Tuple[str, Tuple[int, ...], str, int]]:
        """(name, shape, kind, fan_in) in the reference's state-dict order of names (no biases anywhere: bias=False)."""
        H, Fd, L = self.hidden, self.ffn, self.latent
        out = [("noisy_images_proj.weight", (H, L), "w", L), ("cond_proj.weight", (H, H), "w", H),
               ("t_embedder.mlp.0.weight", (H, self.freq_dim), "w", self.freq_dim), ("t_embedder.mlp.2.weight", (H, H), "w", H)]
        for i in range(self.layers):
            p = f"layers.{i}."
            out += [(p + "ffn.gate_proj.weight", (Fd, H), "w", H), (p + "ffn.up_proj.weight", (Fd, H), "w", H),
                    (p + "ffn.down_proj.weight", (H, Fd), "w", Fd), (p + "norm.weight", (H,), "g", 0),
                    (p + "adaLN_modulation.1.weight", (3 * H, H), "a", H)]
        out += [("final_layer.linear.weight", (L, H), "w", H), ("final_layer.adaLN_modulation.1.weight", (2 * H, H), "a", H)]
        return out


def make_sampler_weights(g: SamplerGeometry, seed: int = 2024, bf16_round: bool = True) -> Dict[str, np.ndarray]:
    """Seeded "stress" init: the reference's own init zeroes every adaLN matrix and the output matrix (:237-250), which
    makes the head return exactly 0 -- a parity test would pass with the kernels deleted.  Here every matrix is
    N(0, 1/fan_in) (adaLN: half that std, so that shift / scale / gate are O(0.5)), norm weights U(0.5, 1.5); values are
    bf16-representable so that the fp32 reference and the bf16 path hold literally the same weights."""
    from .weights import round_to_bf16
    w = {}
    for idx, (name, shape, kind, fan_in) in enumerate(g.tensor_specs()):
        rng = np.random.Generator(np.random.Philox(key=[np.uint64(seed), np.uint64(idx)]))
        if kind == "g":
            t = rng.uniform(0.5, 1.5, size=shape).astype(F32)
        else:
            t = (rng.standard_normal(size=shape, dtype=F32) * F32((0.5 if kind == "a" else 1.0) / math.sqrt(fan_in))).astype(F32)
        w[name] = round_to_bf16(t) if bf16_round else t
    return w


# ---------------------------------------------------------------------------------------------------------------------
# scheduler
# ---------------------------------------------------------------------------------------------------------------------
def cosine_sigmas(train_steps: int) -> np.ndarray:
    """betas_for_alpha_bar(..., 'cosine') (:28-85) -> alphas_cumprod (:252-253) -> sigmas (:264), float32 like torch."""
    def abar(t):
        return math.cos((t + 0.008) / 1.008 * math.pi / 2) ** 2
    betas = np.array([min(1 - abar((i + 1) / train_steps) / abar(i / train_steps), 0.999) for i in range(train_steps)], dtype=F32)
    ac = np.cumprod((F32(1.0) - betas).astype(F32), dtype=F32)
    return (((F32(1.0) - ac) / ac) ** F32(0.5)).astype(F32)


def set_timesteps(train_steps: int, n: int) -> Tuple[np.ndarray, np.ndarray]:
    """set_timesteps (:321-424) for 'linspace' spacing, lambda_min_clipped = -inf (clipped_idx 0), final sigma 0.
    -> (timesteps int64 [n], sigmas float32 [n + 1])."""
    ts = np.linspace(0, train_steps - 1, n + 1).round()[::-1][:-1].copy().astype(np.int64)
    sig = cosine_sigmas(train_steps)
    s = np.interp(ts, np.arange(0, len(sig)), sig)
    return ts, np.concatenate([s, [0.0]]).astype(F32)


def _alpha_sigma(sigma: np.float32) -> Tuple[np.float32, np.float32]:
    """_sigma_to_alpha_sigma_t (:483-487) in float32 (the sigmas are 0-d float32 tensors in the reference)."""
    alpha = F32(1.0) / F32(np.sqrt(F32(sigma * sigma + F32(1.0))))
    return alpha, F32(sigma * alpha)


def step_coefficients(sigmas: np.ndarray) -> np.ndarray:
    """Per step i the solver update is linear in (x, m0, m1) with scalar coefficients:
         m0     = alpha_s * x - sigma_s * v                       convert_model_output, v_prediction (:578-581)
         x_next = A * x + B * m0 + D * (m0 - m1)                  first order (D = 0, :667-674) or second order midpoint
                                                                  (:745-768): D = -0.5 * alpha_t * (exp(-h) - 1) / r0
       First order at step 0 (lower_order_nums < 1) and at the LAST step (final_sigmas_type == 'zero' makes
       lower_order_final true, :977-982); second order otherwise (solver_order 2).
    -> float32 [n, 5] = (alpha_s, sigma_s, A, B, D)."""
    n = len(sigmas) - 1
    out = np.zeros((n, 5), dtype=F32)
    with np.errstate(divide="ignore"):
        for i in range(n):
            a_s, s_s = _alpha_sigma(sigmas[i])
            a_t, s_t = _alpha_sigma(sigmas[i + 1])
            lam_t = F32(np.log(a_t)) - F32(np.log(s_t))          # +inf at the last step (sigma 0)
            lam_s = F32(np.log(a_s)) - F32(np.log(s_s))
            h = F32(lam_t - lam_s)
            em = F32(np.exp(-h)) - F32(1.0)
            A, B, D = F32(s_t / s_s), F32(-(a_t * em)), F32(0.0)
            if not (i == 0 or i == n - 1):
                a_p, s_p = _alpha_sigma(sigmas[i - 1])
                lam_p = F32(np.log(a_p)) - F32(np.log(s_p))
                r0 = F32((lam_s - lam_p) / h)
                D = F32(F32(-0.5) * (a_t * em) / r0)
            out[i] = (a_s, s_s, A, B, D)
    return out


# ---------------------------------------------------------------------------------------------------------------------
# head
# ---------------------------------------------------------------------------------------------------------------------
def silu(x):
    return (x / (F32(1.0) + np.exp(-x))).astype(F32)


def rms_norm(x, weight, eps):
    """RMSNorm (:20-41): x * rsqrt(mean(x^2) + eps) [* weight]."""
    y = x * (F32(1.0) / np.sqrt(np.mean(x * x, axis=-1, keepdims=True, dtype=F32) + F32(eps)))
    return (y * weight if weight is not None else y).astype(F32)


def timestep_embedding(t: np.ndarray, dim: int, max_period: float = 10000.0) -> np.ndarray:
    """TimestepEmbedder.timestep_embedding (:65-88): [cos(t f_k), sin(t f_k)], f_k = exp(-ln(max_period) k / half)."""
    half = dim // 2
    freqs = np.exp(-math.log(max_period) * np.arange(half, dtype=F32) / F32(half)).astype(F32)
    args = t.astype(F32)[:, None] * freqs[None]
    return np.concatenate([np.cos(args), np.sin(args)], axis=-1).astype(F32)


class OracleSampler:
    def __init__(self, weights: Dict[str, np.ndarray], geom: SamplerGeometry = SamplerGeometry()):
        self.g = geom
        self.w = {k: np.ascontiguousarray(v, dtype=F32) for k, v in weights.items()}

    def t_embed(self, t: np.ndarray) -> np.ndarray:
        """TimestepEmbedder.forward (:90-93): Linear -> SiLU -> Linear, no biases."""
        w = self.w
        e = timestep_embedding(t, self.g.freq_dim)
        return silu(e @ w["t_embedder.mlp.0.weight"].T) @ w["t_embedder.mlp.2.weight"].T

    def head(self, noisy: np.ndarray, t: np.ndarray, condition: np.ndarray) -> np.ndarray:
        """VibeVoiceDiffusionHead.forward (:254-280).  noisy [N, latent], t [N], condition [N, hidden] -> [N, latent]."""
        g, w = self.g, self.w
        x = noisy @ w["noisy_images_proj.weight"].T
        c = condition @ w["cond_proj.weight"].T + self.t_embed(t)
        sc = silu(c)                                                                 # adaLN_modulation[0] (:155, :180)
        for i in range(g.layers):                                                    # HeadLayer.forward (:158-161)
            p = f"layers.{i}."
            shift, scale, gate = np.split(sc @ w[p + "adaLN_modulation.1.weight"].T, 3, axis=-1)
            hmod = rms_norm(x, w[p + "norm.weight"], g.eps) * (F32(1.0) + scale) + shift           # modulate (:43-45)
            f = (silu(hmod @ w[p + "ffn.gate_proj.weight"].T) * (hmod @ w[p + "ffn.up_proj.weight"].T)) @ w[p + "ffn.down_proj.weight"].T
            x = x + gate * f
        shift, scale = np.split(sc @ w["final_layer.adaLN_modulation.1.weight"].T, 2, axis=-1)    # FinalLayer (:186-190)
        x = rms_norm(x, None, g.eps) * (F32(1.0) + scale) + shift
        return (x @ w["final_layer.linear.weight"].T).astype(F32)

    def sample(self, condition: np.ndarray, neg_condition: np.ndarray, noise: np.ndarray, cfg_scale: float = 3.0,
               steps: int = 20) -> np.ndarray:
        """sample_speech_tokens (:695-708) for a given initial noise [2n, latent] (the reference draws 2n rows and uses
        only the first n: `half = speech[: len(speech) // 2]`, then carries [half_next, half_next'] -- both halves of
        `speech` evolve from `half` + their own previous value through the scheduler's linear update, so the second half
        is dead weight that never reaches the result; it is kept out of this restatement)."""
        n = condition.shape[0]
        ts, sigmas = set_timesteps(self.g.train_steps, steps)
        co = step_coefficients(sigmas)
        cond = np.concatenate([condition, neg_condition], 0).astype(F32)
        x = noise[:n].astype(F32)
        m_prev = None
        for i, t in enumerate(ts):
            v = self.head(np.concatenate([x, x], 0), np.full((2 * n,), t, dtype=F32), cond)
            eps = v[n:] + F32(cfg_scale) * (v[:n] - v[n:])                          # :704
            a_s, s_s, A, B, D = co[i]
            m0 = a_s * x - s_s * eps
            x = A * x + B * m0 + (D * (m0 - m_prev) if m_prev is not None and D != 0 else F32(0.0))
            m_prev = m0
        return x.astype(F32)
