"""Caller replica of the generation loop's tokenizer traffic (modeling_vibevoice_inference.py:634-662): every step a
subset of streams (``diffusion_indices``) decodes ONE latent frame through the acoustic decoder's streaming cache and the
3200 samples that come out go straight (on the device) into the semantic tokenizer's streaming encode with its own cache;
``set_to_zero`` on both caches at a speech end (:540-544).  ``generate()`` itself needs ``diffusers`` (absent), so the
boundary is exercised with the same calls, the same argument kinds (reference-style cache objects, device index tensors)
and checked step by step against the numpy oracle doing the same walk."""
import numpy as np
import pytest

from oracle.geometry import DecoderGeometry
from oracle.vv_oracle import OracleDecoder, OracleEncoder, OracleStreamingCache, parity_metrics
from oracle.weights import make_encoder_weights, make_latents, make_weights, round_to_bf16

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


class RefStyleCache:
    """Stands in for VibeVoiceTokenizerStreamingCache: constructed with no arguments by the caller (:386-387), mutated
    through set_to_zero (:543-544); the B200 path hangs its device state off the object."""

    def __init__(self):
        self.cache = {}

    def set_to_zero(self, sample_indices):
        self.zero_calls = getattr(self, "zero_calls", 0) + 1


def test_decode_then_semantic_encode_per_frame():
    import vvdec
    ag = DecoderGeometry(depths=(1, 1, 1, 1, 1, 1, 1))                     # acoustic decoder (full channel ladder)
    sg = DecoderGeometry(vae_dim=128, depths=(1, 1, 1, 1, 1, 1, 1))        # semantic encoder: same class, vae_dim 128
    wd, we = make_weights(ag, "mild", 3), make_encoder_weights(sg, "mild", 4)
    dec = vvdec.B200AcousticDecoder(vvdec.Geometry(depths=ag.depths), device=0)
    dec.load_state_dict({k: torch.from_numpy(v) for k, v in wd.items()})
    enc = vvdec.B200AcousticEncoder(vvdec.Geometry(vae_dim=128, depths=sg.depths), device=0, fix_std=None, std_dist_type="none")
    enc.load_state_dict({k: torch.from_numpy(v) for k, v in we.items()})
    od, oe = OracleDecoder(wd, ag), OracleEncoder(we, sg)

    n_streams, steps = 4, 7
    lat = make_latents(steps * n_streams, 1, ag.vae_dim, 50).reshape(steps, n_streams, 1, ag.vae_dim)
    active = [[0, 1, 2, 3], [0, 1, 2, 3], [0, 2, 3], [0, 2, 3], [0, 1, 2, 3], [1, 3], [0, 1, 2, 3]]
    zero_after = {3: [1]}                                                   # stream 1 ends a speech segment after step 3
    a_cache, s_cache = RefStyleCache(), RefStyleCache()                     # what generate() constructs
    oa_cache, os_cache = OracleStreamingCache(), OracleStreamingCache()
    worst_a, worst_s = 1e9, 1e9
    for t in range(steps):
        idx = active[t]
        didx = torch.tensor(idx, device="cuda")                             # diffusion_indices: a device tensor
        x = torch.from_numpy(lat[t, idx]).cuda().to(torch.bfloat16)         # [n, 1, 64] bf16, like scaled_latent
        chunk = dec.decode(x, cache=a_cache, sample_indices=didx, use_cache=True)        # :635-641
        assert chunk.shape == (len(idx), 1, 3200) and chunk.dtype == torch.bfloat16
        sem = enc.encode(chunk, cache=s_cache, sample_indices=didx, use_cache=True).mean  # :656-662
        assert sem.shape == (len(idx), 1, 128)
        # oracle: fp32 walk; its semantic encoder is fed the SAME bf16 audio the device produced (the hand-over tensor)
        ref_chunk = od.decode(lat[t, idx], cache=oa_cache, sample_indices=idx, use_cache=True)
        ma = parity_metrics(chunk.float().cpu().numpy(), ref_chunk)
        ref_sem = oe.encode(chunk.float().cpu().numpy(), cache=os_cache, sample_indices=idx, use_cache=True)
        ms = parity_metrics(sem.float().cpu().numpy(), ref_sem)
        assert ma["snr_db"] >= 40.0 and ma["rel_max"] <= 2e-2, (t, ma)
        assert ms["snr_db"] >= 40.0 and ms["rel_max"] <= 2e-2, (t, ms)
        worst_a, worst_s = min(worst_a, ma["snr_db"]), min(worst_s, ms["snr_db"])
        if t in zero_after:
            z = torch.tensor(zero_after[t], device="cuda")
            a_cache.set_to_zero(z)
            s_cache.set_to_zero(z)
            oa_cache.set_to_zero(zero_after[t])
            os_cache.set_to_zero(zero_after[t])
    assert a_cache.zero_calls == 1 and s_cache.zero_calls == 1              # the caller's own method still ran
    assert hasattr(a_cache, "_b200_state") and hasattr(s_cache, "_b200_enc_state")
    print(f"worst SNR over {steps} steps: decode {worst_a:.1f} dB, semantic encode {worst_s:.1f} dB")
