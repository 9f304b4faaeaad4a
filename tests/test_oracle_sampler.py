"""Pin the numpy restatement of the diffusion-head sampler (oracle/sampler_oracle.py) against outputs of
the UNMODIFIED reference scheduler / head / sample_speech_tokens (tests/golden/samp_*.npz, oracle/gen_golden_sampler.py).
CPU only."""
import json
import os

import numpy as np
import pytest

from oracle.gen_golden_sampler import inputs
from oracle.sampler_oracle import OracleSampler, SamplerGeometry, make_sampler_weights, set_timesteps, step_coefficients
from oracle.vv_oracle import parity_metrics

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    meta = json.loads(str(z["meta"]))
    return z, meta, SamplerGeometry(**meta["geom"])


@pytest.mark.parametrize("name", ["samp_tiny_n3_s20", "samp_tiny_n5_s10", "samp_full_n2_s20"])
def test_sampler_oracle_matches_reference(name):
    z, meta, g = load(name)
    ts, sig = set_timesteps(g.train_steps, meta["steps"])
    assert np.array_equal(ts, z["timesteps"])                                  # integer work: exact
    # the reference builds alphas_cumprod as a FLOAT32 running product of 1000 terms (dpm_solver.py:252-253): torch and numpy
    # accumulate it in different order, worth up to ~1e-5 relative on the small sigmas (1000 x 2^-24 worst case 6e-5)
    assert np.allclose(sig, z["sigmas"], rtol=5e-5, atol=0)
    orc = OracleSampler(make_sampler_weights(g, meta["wseed"]), g)
    cond, neg = inputs(g, meta["n"], meta["iseed"])
    n = meta["n"]
    v = orc.head(z["noise"], np.full((2 * n,), meta["t_probe"], dtype=np.float32), np.concatenate([cond, neg], 0))
    mh = parity_metrics(v, z["head_v"])
    assert mh["snr_db"] > 100.0, mh                                            # fp32 vs fp32: summation order
    out = orc.sample(cond, neg, z["noise"], cfg_scale=meta["cfg_scale"], steps=meta["steps"])
    assert out.shape == z["out"].shape
    m = parity_metrics(out, z["out"])
    assert m["snr_db"] > 80.0 and m["rel_max"] < 1e-3, m                       # 10-20 dependent steps of fp32 noise


def test_step_coefficients_shape_and_limits():
    _, sig = set_timesteps(1000, 20)
    co = step_coefficients(sig)
    assert co.shape == (20, 5) and np.isfinite(co).all()
    assert co[0, 4] == 0 and co[-1, 4] == 0 and (co[1:-1, 4] != 0).all()        # first order at both ends, second between
    # last step lands on the data prediction: sigma_t = 0 -> x_next = m0  (A = 0, B = 1)
    assert co[-1, 2] == 0 and abs(co[-1, 3] - 1.0) < 1e-6
