"""Parity of the diffusion-head sampler (vv_sampler_* through the Python mirror of sample_speech_tokens)
against (a) outputs of the UNMODIFIED reference scheduler / head / sample_speech_tokens (tests/golden/samp_*.npz) and
(b) the numpy oracle on seeded inputs at sizes the fixtures do not hold.

Tolerance: the path's own bound (BASELINE.json north_star), SNR >= 40 dB and max-abs <= 2e-2 of the reference peak, for one
head pass AND for the sampled latent.  The sampled latent is the end of a 10-20 step recursion with classifier-free guidance
(the difference of two head outputs times cfg_scale), so amplification of the bf16 operand noise was a concern; measured on
B200 it is not there: one head pass 51 dB, sampled latents 55-59 dB (fp32 in / out) and 51-54 dB (bf16 in / out) on the
three reference fixtures, 55.9 dB on the 70-stream oracle case -- the solver update is contractive."""
import json
import os

import numpy as np
import pytest

from oracle.gen_golden_sampler import inputs
from oracle.sampler_oracle import OracleSampler, SamplerGeometry, make_sampler_weights
from oracle.vv_oracle import parity_metrics

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
SAMPLE_SNR_MIN, SAMPLE_REL_MAX = 40.0, 2e-2


def load(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    meta = json.loads(str(z["meta"]))
    return z, meta, SamplerGeometry(**meta["geom"])


def build(g, w, steps):
    import vvdec
    smp = vvdec.B200DiffusionSampler(vvdec.SamplerGeometry(hidden=g.hidden, layers=g.layers, ffn_ratio=g.ffn_ratio,
                                                            latent=g.latent, eps=g.eps, train_steps=g.train_steps,
                                                            inference_steps=steps), device=0)
    return smp.load_state_dict({k: torch.from_numpy(v) for k, v in w.items()})


@pytest.mark.parametrize("name", ["samp_tiny_n3_s20", "samp_tiny_n5_s10", "samp_full_n2_s20"])
def test_sampler_matches_reference(name):
    z, meta, g = load(name)
    w = make_sampler_weights(g, meta["wseed"])
    smp = build(g, w, meta["steps"])
    assert smp.param_count == sum(int(np.prod(s)) for _, s, _, _ in g.tensor_specs())
    cond, neg = inputs(g, meta["n"], meta["iseed"])
    n = meta["n"]
    # (1) one head forward, same inputs as the fixture's probe
    v = smp.head(torch.from_numpy(z["noise"]).cuda(), meta["t_probe"], torch.from_numpy(np.concatenate([cond, neg], 0)).cuda())
    mh = parity_metrics(v.cpu().numpy(), z["head_v"])
    print(name, "head", mh)
    assert mh["snr_db"] >= 40.0 and mh["rel_max"] <= 2e-2, mh
    # (2) the whole sampling loop from the reference's own noise draw
    for dt in (torch.float32, torch.bfloat16):
        out = smp.sample_speech_tokens(torch.from_numpy(cond).cuda().to(dt), torch.from_numpy(neg).cuda().to(dt),
                                       cfg_scale=meta["cfg_scale"], noise=torch.from_numpy(z["noise"][:n]).cuda())
        assert out.shape == z["out"].shape and out.dtype == dt
        m = parity_metrics(out.float().cpu().numpy(), z["out"])
        print(name, dt, "sample", m, "launches", smp.take_launch_count())
        assert m["snr_db"] >= SAMPLE_SNR_MIN and m["rel_max"] <= SAMPLE_REL_MAX, m


def test_sampler_matches_oracle_larger_batch_and_is_deterministic():
    """n = 70 streams (2n = 140 rows: a second, ragged row tile in every GEMM), 12 steps, against the numpy oracle."""
    g = SamplerGeometry(hidden=384, layers=3)
    w = make_sampler_weights(g, 77)
    smp = build(g, w, 12)
    n = 70
    cond, neg = inputs(g, n, 21)
    noise = np.random.Generator(np.random.Philox(key=[5, 5])).standard_normal((n, g.latent), dtype=np.float32)
    ref = OracleSampler(w, g).sample(cond, neg, noise, cfg_scale=2.0, steps=12)
    args = (torch.from_numpy(cond).cuda(), torch.from_numpy(neg).cuda())
    out = smp.sample_speech_tokens(*args, cfg_scale=2.0, noise=torch.from_numpy(noise).cuda())
    m = parity_metrics(out.cpu().numpy(), ref)
    print("n=70", m)
    assert m["snr_db"] >= SAMPLE_SNR_MIN and m["rel_max"] <= SAMPLE_REL_MAX, m
    again = smp.sample_speech_tokens(*args, cfg_scale=2.0, noise=torch.from_numpy(noise).cuda())
    assert torch.equal(out, again)                                   # same kernels, same order: bit-equal
    # rows are independent: a sub-batch reproduces its rows (same tiles per row -> bit-equal is not guaranteed across tile
    # shapes, so compare within rounding)
    sub = smp.sample_speech_tokens(args[0][:9], args[1][:9], cfg_scale=2.0, noise=torch.from_numpy(noise[:9]).cuda())
    ms = parity_metrics(sub.cpu().numpy(), out[:9].cpu().numpy())
    assert ms["snr_db"] >= 60.0, ms


@pytest.mark.parametrize("hidden,n", [(2048, 5), (3584, 3)])
def test_sampler_wide_heads(hidden, n):
    """Heads wider than the 1.5B model's 1536 (3584 = the 7B model's): K >= 2048 makes the GEMMs take their long-K (DEEP)
    instantiations -- gemm_tc2_kernel<EPI_SWIGLU_BF16, true, 256> for [gate; up], the bias epilogue for adaLN / cond_proj --
    and the down projection (K = 3 x hidden) its K-slice planes with more k-blocks per slice.  One layer, 4 steps."""
    g = SamplerGeometry(hidden=hidden, layers=1)
    w = make_sampler_weights(g, 300 + hidden)
    smp = build(g, w, 4)
    cond, neg = inputs(g, n, 40 + n)
    noise = np.random.Generator(np.random.Philox(key=[9, hidden])).standard_normal((n, g.latent), dtype=np.float32)
    ref = OracleSampler(w, g).sample(cond, neg, noise, cfg_scale=1.7, steps=4)
    out = smp.sample_speech_tokens(torch.from_numpy(cond).cuda(), torch.from_numpy(neg).cuda(), cfg_scale=1.7,
                                   noise=torch.from_numpy(noise).cuda())
    m = parity_metrics(out.cpu().numpy(), ref)
    print(hidden, m)
    assert m["snr_db"] >= SAMPLE_SNR_MIN and m["rel_max"] <= SAMPLE_REL_MAX, m


def test_sampler_errors_and_default_noise():
    import vvdec
    with pytest.raises(ValueError):
        vvdec.B200DiffusionSampler(vvdec.SamplerGeometry(hidden=100), device=0)       # not a multiple of 128
    g = SamplerGeometry(hidden=256, layers=2)
    smp = build(g, make_sampler_weights(g, 2024), 20)
    with pytest.raises(RuntimeError):
        smp.sample_speech_tokens(torch.zeros(2, 256), torch.zeros(2, 256))
    with pytest.raises(ValueError):
        smp.sample_speech_tokens(torch.zeros(2, 128, device="cuda"), torch.zeros(2, 128, device="cuda"))
    c = torch.randn(4, 256, device="cuda")
    torch.manual_seed(3)
    a = smp.sample_speech_tokens(c, -c)                               # draws its own noise like the reference (:698)
    torch.manual_seed(3)
    b = smp.sample_speech_tokens(c, -c)
    assert a.shape == (4, 64) and torch.isfinite(a).all() and torch.equal(a, b)
    assert smp.sample_speech_tokens(c[:0], c[:0]).shape == (0, 64)
