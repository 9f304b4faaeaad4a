"""The host wrapper against the GENUINE reference objects (build container only: /root/reference is imported through
oracle/ref_shim.py; skipped where the tree is absent, e.g. on the GPU box).  The C library is mocked at the ctypes
boundary, so this runs on CPU and checks the wiring only: config parsing, state-dict name matching, patch() and the
forwarding of a real VibeVoiceTokenizerStreamingCache's set_to_zero / clear to the device state.
Reference: modular_vibevoice_tokenizer.py:193-256 (cache), :1003-1108 (model), configuration_vibevoice.py:13-127."""
import ctypes as C

import pytest

from oracle.ref_shim import import_reference, reference_available

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.skipif(not reference_available(), reason="reference tree not present (GPU box)")


class FakeLib:
    """Stands in for libvvdec_b200.so: records every call, hands out fake handles, returns VV_OK."""

    def __init__(self):
        self.calls = []

    def __getattr__(self, name):
        def fn(*args):
            self.calls.append((name, args))
            if name in ("vv_decoder_create", "vv_encoder_create", "vv_state_create", "vv_encoder_state_create", "vv_sampler_create"):
                args[-1]._obj.value = 0x1000 + len(self.calls)
            if name in ("vv_decoder_halo_frames", "vv_encoder_halo_frames"):
                return 58
            if name in ("vv_last_error", "vv_status_string"):
                return b""
            return 0
        return fn

    def names(self, fn):
        return [a[1].decode() for n, a in self.calls if n == fn]


@pytest.fixture()
def fake(monkeypatch):
    import vvdec
    from vvdec import _capi
    lib = FakeLib()
    monkeypatch.setattr(_capi, "load", lambda: lib)
    monkeypatch.setattr(torch.cuda, "is_available", lambda: True)
    monkeypatch.setattr(torch.cuda, "current_device", lambda: 0)

    class _S:
        cuda_stream = 0
    monkeypatch.setattr(torch.cuda, "current_stream", lambda *a, **k: _S())
    for cls in (vvdec.B200AcousticDecoder, vvdec.B200AcousticEncoder):       # index tensors stay on the CPU here
        monkeypatch.setattr(cls, "_indices_tensor",
                            lambda self, si: torch.as_tensor(si if not isinstance(si, torch.Tensor) else si.tolist(),
                                                             dtype=torch.int32))
    return lib


def test_geometry_from_the_genuine_configs():
    from vvdec.decoder import Geometry
    Cfg, _, _ = import_reference()
    from vibevoice.modular.configuration_vibevoice import VibeVoiceSemanticTokenizerConfig
    g = Geometry.from_config(Cfg())                       # defaults == the tokenizer block of both shipped JSONs
    assert g == Geometry() and g.hop == 3200
    assert Geometry.from_encoder_config(Cfg()) == Geometry()
    import json
    with open("/root/reference/vibevoice/configs/qwen2.5_1.5b_64k.json") as f:     # the shipped semantic block: vae_dim 128
        sem = {k: v for k, v in json.load(f)["semantic_tokenizer_config"].items() if k != "model_type"}
    sg = Geometry.from_encoder_config(VibeVoiceSemanticTokenizerConfig(**sem))
    assert sg.vae_dim == 128 and sg.depths == g.depths and sg.ratios == g.ratios and sg.n_filters == 32
    # a decoder that is NOT the encoder's mirror: each side reads its own fields
    c = Cfg(decoder_depths="2-1-1-1-1-1-1", decoder_ratios=[8, 5, 5, 4, 2, 2], encoder_depths="3-3-3-3-3-3-8")
    assert Geometry.from_config(c).depths == (2, 1, 1, 1, 1, 1, 1)
    assert Geometry.from_encoder_config(c).depths == (8, 3, 3, 3, 3, 3, 3)
    for field, bad in (("causal", False), ("mixer_layer", "conv"), ("layernorm", "LN"), ("pad_mode", "reflect"),
                       ("conv_norm", "weight_norm"), ("layernorm_elementwise_affine", False), ("conv_bias", False)):
        with pytest.raises(ValueError):
            Geometry.from_config(Cfg(**{field: bad}))


def test_state_dict_names_and_shapes_match_the_genuine_model():
    """Every decoder.* / encoder.* parameter of the genuine VibeVoiceAcousticTokenizerModel(Config()) is a tensor the packer
    expects, with the same shape, and nothing is missing.  (Built for real, in bf16 to halve the memory: its constructor
    calls .item(), so the meta device does not work.)"""
    from vvdec.decoder import Geometry
    from vvdec.synthetic import encoder_tensor_specs, tensor_specs
    Cfg, Model, _ = import_reference()
    prev = torch.get_default_dtype()
    torch.set_default_dtype(torch.bfloat16)
    try:
        model = Model(Cfg())
    finally:
        torch.set_default_dtype(prev)
    sd = {k: tuple(v.shape) for k, v in model.state_dict().items()}
    ours = {n: tuple(s) for n, s, _, _ in tensor_specs(Geometry())}
    ours.update({n: tuple(s) for n, s, _, _ in encoder_tensor_specs(Geometry())})
    assert {k: v for k, v in sd.items() if k.startswith(("decoder.", "encoder."))} == ours
    assert set(sd) - set(ours) <= {"fix_std"}             # the only non-layer entry, if registered as a buffer


def _tiny_reference():
    Cfg, Model, Cache = import_reference()
    cfg = Cfg(vae_dim=8, encoder_n_filters=4, decoder_n_filters=4, encoder_ratios=[2, 2], encoder_depths="1-1-2")
    torch.manual_seed(0)
    return Model(cfg).eval(), Cache


def test_patch_and_streaming_cache_forwarding_with_genuine_objects(fake):
    import vvdec
    model, Cache = _tiny_reference()
    dec = vvdec.patch(model, encode=True)
    # every decoder parameter of the genuine state dict went through vv_decoder_load_tensor under its own name
    want = sorted(k for k in model.state_dict() if k.startswith("decoder."))
    assert sorted(fake.names("vv_decoder_load_tensor")) == want
    assert sorted(fake.names("vv_encoder_load_tensor")) == sorted(k for k in model.state_dict() if k.startswith("encoder."))
    assert [n for n, _ in fake.calls].count("vv_decoder_finalize") == 1
    cfg = [a for n, a in fake.calls if n == "vv_decoder_create"][0][0]._obj
    assert (cfg.vae_dim, cfg.n_filters, cfg.n_stages, list(cfg.ratios)[:2], list(cfg.depths)[:3]) == (8, 4, 3, [2, 2], [2, 1, 1])
    assert model.decode == dec.decode and model._b200_decoder is dec and model.encode == model._b200_encoder.encode
    assert model.fix_std == model._b200_encoder.fix_std and model._b200_encoder.std_dist_type == model.std_dist_type
    # a genuine cache, constructed the way generate() does (modeling_vibevoice_inference.py:386), gets the device state
    cache = Cache()
    st = dec._state_for(cache, 3)
    assert isinstance(st, vvdec.B200StreamingCache) and cache._b200_state is st and dec._state_for(cache) is st
    assert st.n_slots == max(dec.default_slots, 3)
    n0 = len(fake.calls)
    cache.set_to_zero(torch.tensor([1, 2]))               # :540-544; the reference's own method still runs (empty dict: no-op)
    z = [a for n, a in fake.calls[n0:] if n == "vv_state_zero"]
    assert len(z) == 1 and z[0][2] == 2                   # forwarded: two slot ids
    cache.clear()                                         # :243-256
    z = [a for n, a in fake.calls[n0:] if n == "vv_state_zero"]
    assert len(z) == 2 and z[1][2] == -1                  # all slots
    assert cache.cache == {}                              # the reference object is intact
    # the encoder hangs ITS state off the same kind of object under another attribute
    est = model._b200_encoder._state_for(Cache(), 2)
    assert est._dec is model._b200_encoder
    assert any(n == "vv_encoder_state_create" for n, _ in fake.calls)


def test_patch_finds_the_tokenizer_inside_a_wrapping_model(fake):
    """generate() reaches the tokenizer as self.model.acoustic_tokenizer (modeling_vibevoice.py:119-142)."""
    import vvdec
    tok, _ = _tiny_reference()

    class Inner(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.acoustic_tokenizer = tok

    class Outer(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.model = Inner()
    outer = Outer()
    dec = vvdec.patch(outer)
    assert outer.model.acoustic_tokenizer.decode == dec.decode


def test_sampler_geometry_names_and_patch_with_genuine_objects(fake):
    """the genuine VibeVoiceDiffusionHeadConfig (shipped JSON block) -> SamplerGeometry; a genuine
    VibeVoiceDiffusionHead's state-dict names / shapes == what the packer expects; patch_sampler() swaps
    sample_speech_tokens on a model-shaped object and loads every tensor under its own name."""
    import json
    import types
    import vvdec
    from oracle.ref_shim import import_reference_sampler
    from oracle.sampler_oracle import SamplerGeometry as OG
    _, Head, HeadCfg, _ = import_reference_sampler()
    with open("/root/reference/vibevoice/configs/qwen2.5_1.5b_64k.json") as f:
        blk = {k: v for k, v in json.load(f)["diffusion_head_config"].items() if k != "model_type"}
    g = vvdec.SamplerGeometry.from_config(HeadCfg(**blk))
    assert (g.hidden, g.layers, g.ffn, g.latent, g.train_steps, g.inference_steps) == (1536, 4, 4608, 64, 1000, 20)
    assert g == vvdec.SamplerGeometry()                                       # the defaults ARE the 1.5B model's head
    for field, bad in (("prediction_type", "epsilon"), ("ddpm_beta_schedule", "linear")):
        with pytest.raises(ValueError):
            vvdec.SamplerGeometry.from_config(HeadCfg(**dict(blk, **{field: bad})))
    head = Head(HeadCfg(hidden_size=256, head_layers=2)).eval()
    want = {n: tuple(s) for n, s, _, _ in OG(hidden=256, layers=2).tensor_specs()}
    assert {k: tuple(v.shape) for k, v in head.state_dict().items()} == want
    model = types.SimpleNamespace(model=types.SimpleNamespace(prediction_head=head), ddpm_inference_steps=10)
    smp = vvdec.patch_sampler(model)
    assert sorted(fake.names("vv_sampler_load_tensor")) == sorted(want)
    assert model.sample_speech_tokens == smp.sample_speech_tokens and smp.ddpm_inference_steps == 10
    cfg = [a for n, a in fake.calls if n == "vv_sampler_create"][0][0]._obj
    assert (cfg.hidden, cfg.layers, cfg.ffn, cfg.latent, cfg.train_steps) == (256, 2, 768, 64, 1000)
