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


@pytest.mark.parametrize("use_graph", [True, False])
def test_frame_step_graph_matches_oracle_without_host_syncs(use_graph):
    """the same call pattern through vvdec.FrameStep -- decode -> semantic encode -> PCM16 as ONE CUDA graph
    per stream-count bucket (steps with 3 or 2 active streams are padded to the bucket of 4 with slot id -1), sample indices
    consumed on the device.  The whole loop runs under torch.cuda.set_sync_debug_mode("error"): any host synchronisation
    raises.  Outputs are compared afterwards, step by step, with the numpy oracle doing the same walk and with the eager
    per-call path (bit-equal: same kernels)."""
    import vvdec
    from oracle.pcm16 import pcm16_rows
    ag = DecoderGeometry(depths=(1, 1, 1, 1, 1, 1, 1))
    sg = DecoderGeometry(vae_dim=128, depths=(1, 1, 1, 1, 1, 1, 1))
    wd, we = make_weights(ag, "mild", 3), make_encoder_weights(sg, "mild", 4)
    dec = vvdec.B200AcousticDecoder(vvdec.Geometry(depths=ag.depths), device=0)
    dec.load_state_dict({k: torch.from_numpy(v) for k, v in wd.items()})
    enc = vvdec.B200AcousticEncoder(vvdec.Geometry(vae_dim=128, depths=sg.depths), device=0, fix_std=None, std_dist_type="none")
    enc.load_state_dict({k: torch.from_numpy(v) for k, v in we.items()})
    od, oe = OracleDecoder(wd, ag), OracleEncoder(we, sg)
    n_streams, steps = 4, 7
    lat = make_latents(steps * n_streams, 1, ag.vae_dim, 50).reshape(steps, n_streams, 1, ag.vae_dim)
    active = [[0, 1, 2, 3], [0, 1, 2, 3], [0, 2, 3], [0, 2, 3], [0, 1, 2, 3], [1, 3], [0, 1, 2, 3]]
    zero_after = {3: [1]}
    fs = vvdec.FrameStep(dec, enc, pcm=True, n_slots=8, bucket=4, use_graph=use_graph)
    xs = [torch.from_numpy(lat[t, active[t]]).cuda().to(torch.bfloat16) for t in range(steps)]
    ds = [torch.tensor(active[t], device="cuda") for t in range(steps)]
    zs = {t: torch.tensor(v, device="cuda") for t, v in zero_after.items()}
    fs(xs[0], ds[0])                                    # builds the bucket's buffers / graph (allocations, capture)
    fs.set_to_zero(torch.arange(n_streams, device="cuda"))
    torch.cuda.synchronize()
    outs = []
    torch.cuda.set_sync_debug_mode("error")
    try:
        for t in range(steps):
            outs.append(fs(xs[t], ds[t]))
            if t in zs:
                fs.set_to_zero(zs[t])
    finally:
        torch.cuda.set_sync_debug_mode("default")
    torch.cuda.synchronize()
    fs.check()
    if use_graph:
        assert fs.graph_launches == steps + 1 and len(fs._plans) == 1
    # eager per-call walk with fresh caches: the graph replays the same kernels, so results are bit-equal
    a_cache, s_cache = dec.new_cache(8), enc.new_cache(8)
    oa_cache, os_cache = OracleStreamingCache(), OracleStreamingCache()
    for t in range(steps):
        idx = active[t]
        chunk = dec.decode(xs[t], cache=a_cache, sample_indices=ds[t], use_cache=True)
        sem = enc.encode(chunk, cache=s_cache, sample_indices=ds[t], use_cache=True).mean
        o = outs[t]
        assert o.audio.shape == (len(idx), 1, 3200) and o.semantic.shape == (len(idx), 1, 128) and o.pcm.shape == (len(idx), 3200)
        assert torch.equal(o.audio, chunk) and torch.equal(o.semantic, sem), t
        ref_chunk = od.decode(lat[t, idx], cache=oa_cache, sample_indices=idx, use_cache=True)
        ma = parity_metrics(o.audio.float().cpu().numpy(), ref_chunk)
        ref_sem = oe.encode(o.audio.float().cpu().numpy(), cache=os_cache, sample_indices=idx, use_cache=True)
        ms = parity_metrics(o.semantic.float().cpu().numpy(), ref_sem)
        assert ma["snr_db"] >= 40.0 and ma["rel_max"] <= 2e-2, (t, ma)
        assert ms["snr_db"] >= 40.0 and ms["rel_max"] <= 2e-2, (t, ms)
        want, _ = pcm16_rows(o.audio.float().cpu().numpy()[:, 0])
        assert np.array_equal(o.pcm.cpu().numpy(), want), t              # integer output: bit-exact
        if t in zero_after:
            a_cache.set_to_zero(zs[t]); s_cache.set_to_zero(zs[t])
            oa_cache.set_to_zero(zero_after[t]); os_cache.set_to_zero(zero_after[t])


def test_out_of_range_slot_ids_are_safe_and_reported():
    """indices given as CUDA tensors are never seen by the host.  An id outside [0, n_slots) must not index the
    state arena; the row decodes like a first chunk, other streams are untouched, and the NEXT call reports IndexError."""
    import vvdec
    ag = DecoderGeometry(depths=(1, 1, 1, 1, 1, 1, 1))
    wd = make_weights(ag, "mild", 3)
    dec = vvdec.B200AcousticDecoder(vvdec.Geometry(depths=ag.depths), device=0)
    dec.load_state_dict({k: torch.from_numpy(v) for k, v in wd.items()})
    lat = torch.from_numpy(make_latents(2, 3, ag.vae_dim, 9)).cuda()
    cache = dec.new_cache(4)
    good = torch.tensor([1, 2], device="cuda")
    a0 = dec.decode(lat[:, :1], cache=cache, sample_indices=good, use_cache=True)
    bad = torch.tensor([1, 99], device="cuda")
    a1 = dec.decode(lat[:, 1:2], cache=cache, sample_indices=bad, use_cache=True)      # row 1 has no slot
    torch.cuda.synchronize()
    ref = dec.decode(lat[:1, :2])                                                       # stream 1 continued normally
    assert torch.equal(torch.cat([a0[:1], a1[:1]], -1), ref)
    assert torch.equal(a1[1:], dec.decode(lat[1:, 1:2]))                                # the bad row == a first chunk
    with pytest.raises(IndexError):
        dec.decode(lat[:, 2:3], cache=cache, sample_indices=good, use_cache=True)       # reported once, by the next call
    a2 = dec.decode(lat[:, 2:3], cache=cache, sample_indices=good, use_cache=True)      # and cleared
    assert a2.shape == (2, 1, 3200)
    pad = torch.tensor([1, -1], device="cuda")                                          # -1 = padding row: no report
    dec.decode(lat[:, 2:3], cache=cache, sample_indices=pad, use_cache=True)
    torch.cuda.synchronize()
    cache.check()
