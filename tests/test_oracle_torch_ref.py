"""Pin oracle/torch_ref.py (the reference's ATen call sequence, used as the CPU arm of bench.py and as its parity
checker at K2 / K5 / K3 sizes) against the committed outputs of the UNMODIFIED reference and the numpy oracle.  CPU only."""
import numpy as np
import pytest

from conftest import load_golden, geom_from_meta
from oracle import torch_ref
from oracle.vv_oracle import parity_metrics
from oracle.weights import make_audio, make_encoder_weights, make_latents

torch = pytest.importorskip("torch")


@pytest.mark.parametrize("name", ["full_stress_T6", "full_default_T6", "full_mild_T8", "tiny_stress_T12",
                                  "tiny_stress_T16_ambiguous", "full_stress_K1_T75"])
def test_decode_matches_reference_outputs(name, weight_cache):
    z, meta = load_golden(name)
    geom = geom_from_meta(meta)
    w = torch_ref.to_weights(weight_cache(geom, meta["init"], meta["wseed"]))
    lat = torch.from_numpy(make_latents(meta["B"], meta["T"], geom.vae_dim, meta["lseed"]))
    y = torch_ref.decode(w, geom, lat).numpy()
    assert y.shape == z["audio"].shape
    m = parity_metrics(y, z["audio"])
    assert m["snr_db"] > 100.0 and m["rel_max"] < 1e-4, m          # same ATen kernels as the reference: rounding only


@pytest.mark.parametrize("name", ["enc_full_stress_L6500", "enc_full_stress_L1", "enc_full_default_L9999",
                                  "enc_full_mild_B2_L12800", "enc_tiny_stress_L10117", "enc_sem128_stress_L7001"])
def test_encode_matches_reference_outputs(name):
    z, meta = load_golden(name)
    geom = geom_from_meta(meta)
    w = torch_ref.to_weights(make_encoder_weights(geom, meta["init"], meta["wseed"]))
    audio = torch.from_numpy(make_audio(meta["B"], meta["L"], meta["aseed"]))
    y = torch_ref.encode(w, geom, audio).numpy()
    assert y.shape == z["mean"].shape
    m = parity_metrics(y, z["mean"])
    assert m["snr_db"] > 100.0 and m["rel_max"] < 1e-4, m


def test_thread_helper_uses_the_host_cores(monkeypatch):
    monkeypatch.setenv("OMP_NUM_THREADS", "1")                       # what torchrun exports to every rank
    before = torch.get_num_threads()
    try:
        n = torch_ref.use_all_host_threads()
        assert n >= 1 and n == torch.get_num_threads()
    finally:
        torch.set_num_threads(before)
