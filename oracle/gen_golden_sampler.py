"""Generate tests/golden/samp_*.npz from the UNMODIFIED reference sampler (build container only):
DPMSolverMultistepScheduler (vibevoice/schedule/dpm_solver.py), VibeVoiceDiffusionHead
(modular_vibevoice_diffusion_head.py) and the code of sample_speech_tokens (modeling_vibevoice_inference.py:695-708),
imported / lifted by oracle/ref_shim.import_reference_sampler.  fp32 on CPU, seeded stress weights (bf16-representable).

    python -m oracle.gen_golden_sampler

Each fixture holds: the scheduler's timesteps and sigmas, one head forward (noisy, t, condition -> v), the initial noise
(replayed from the torch seed the reference function consumed) and the sampled latents."""
from __future__ import annotations

import json
import os
import types

import numpy as np
import torch

from .ref_shim import import_reference_sampler
from .sampler_oracle import SamplerGeometry, make_sampler_weights
from .weights import round_to_bf16

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")

CASES = [  # name, geometry, n, steps, cfg_scale, weight seed, input seed
    ("samp_tiny_n3_s20", SamplerGeometry(hidden=256, layers=2), 3, 20, 3.0, 2024, 7),
    ("samp_tiny_n5_s10", SamplerGeometry(hidden=256, layers=2), 5, 10, 1.3, 2025, 8),
    ("samp_full_n2_s20", SamplerGeometry(), 2, 20, 3.0, 2026, 9),
]


def inputs(g: SamplerGeometry, n: int, seed: int):
    rng = np.random.Generator(np.random.Philox(key=[np.uint64(31), np.uint64(seed)]))
    cond = round_to_bf16(rng.standard_normal((n, g.hidden), dtype=np.float32))
    neg = round_to_bf16(rng.standard_normal((n, g.hidden), dtype=np.float32))
    return cond, neg


def main():
    Sched, Head, HeadCfg, sample_fn = import_reference_sampler()
    for name, g, n, steps, cfg_scale, wseed, iseed in CASES:
        w = make_sampler_weights(g, wseed)
        head = Head(HeadCfg(hidden_size=g.hidden, head_layers=g.layers, head_ffn_ratio=g.ffn_ratio, rms_norm_eps=g.eps,
                            latent_size=g.latent)).eval()
        sd = head.state_dict()
        assert set(sd) == set(w), (sorted(set(sd) ^ set(w)))
        head.load_state_dict({k: torch.from_numpy(v) for k, v in w.items()})
        sched = Sched(num_train_timesteps=g.train_steps, beta_schedule="cosine", prediction_type="v_prediction")  # modeling_vibevoice.py:138-142
        cond, neg = inputs(g, n, iseed)
        stub = types.SimpleNamespace(model=types.SimpleNamespace(noise_scheduler=sched, prediction_head=head),
                                     ddpm_inference_steps=steps, config=types.SimpleNamespace(acoustic_vae_dim=g.latent))
        with torch.no_grad():
            torch.manual_seed(1000 + iseed)
            out = sample_fn(stub, torch.from_numpy(cond), torch.from_numpy(neg), cfg_scale=cfg_scale).numpy()
            torch.manual_seed(1000 + iseed)
            noise = torch.randn(2 * n, g.latent).numpy()                       # the draw the function made (:698)
            t_probe = torch.full((2 * n,), float(sched.timesteps[3]))
            v = head(torch.from_numpy(noise), t_probe, condition=torch.from_numpy(np.concatenate([cond, neg], 0))).numpy()
        meta = dict(kind="sampler", geom=dict(hidden=g.hidden, layers=g.layers, ffn_ratio=g.ffn_ratio, latent=g.latent, eps=g.eps,
                                              train_steps=g.train_steps), n=n, steps=steps, cfg_scale=cfg_scale, wseed=wseed,
                    iseed=iseed, t_probe=float(sched.timesteps[3]))
        np.savez_compressed(os.path.join(OUT, name + ".npz"), meta=json.dumps(meta), out=out, noise=noise, head_v=v,
                            timesteps=sched.timesteps.numpy(), sigmas=sched.sigmas.numpy())
        print(name, out.shape, "rms", float(np.sqrt((out ** 2).mean())), "peak", float(np.abs(out).max()))


if __name__ == "__main__":
    main()
