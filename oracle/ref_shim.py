"""Import the UNMODIFIED reference tokenizer from /root/reference (build container only).

TEST INFRASTRUCTURE ONLY.  /root/reference does not exist on the GPU box, so nothing in the
``-m gpu`` tests, ``smoke()`` or ``bench.py`` may call this at run time; it is used by
``oracle/gen_golden.py`` (fixture generation) and by CPU tests that skip when the tree is absent.

transformers >= 5 already owns model_type "vibevoice_acoustic_tokenizer", so the reference's
``AutoModel.register`` at modular_vibevoice_tokenizer.py:1188 raises; the shim passes
``exist_ok=True``.  The reference source is not touched.
"""
from __future__ import annotations

import os
import sys
from typing import Dict

import numpy as np

REFERENCE_ROOT = os.environ.get("VV_REFERENCE_ROOT", "/root/reference")


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "vibevoice", "modular", "modular_vibevoice_tokenizer.py"))


def import_reference():
    """-> (VibeVoiceAcousticTokenizerConfig, VibeVoiceAcousticTokenizerModel, VibeVoiceTokenizerStreamingCache)"""
    if not reference_available():
        raise RuntimeError(f"reference tree not found at {REFERENCE_ROOT}")
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    from transformers.models.auto import AutoModel as _AM
    if not getattr(_AM, "_vv_exist_ok_shim", False):
        _orig = _AM.register.__func__
        _AM.register = classmethod(lambda cls, c, m, exist_ok=False: _orig(cls, c, m, exist_ok=True))
        _AM._vv_exist_ok_shim = True
    from vibevoice.modular.configuration_vibevoice import VibeVoiceAcousticTokenizerConfig
    from vibevoice.modular.modular_vibevoice_tokenizer import (
        VibeVoiceAcousticTokenizerModel, VibeVoiceTokenizerStreamingCache)
    return VibeVoiceAcousticTokenizerConfig, VibeVoiceAcousticTokenizerModel, VibeVoiceTokenizerStreamingCache


def build_reference_model(geom, weights: Dict[str, np.ndarray]):
    """Reference model (fp32, CPU, eval) for ``geom`` with ``weights`` loaded into its decoder."""
    import torch
    Cfg, Model, _ = import_reference()
    depths = "-".join(str(d) for d in reversed(geom.depths))        # encoder order; decoder reverses (:1028)
    cfg = Cfg(vae_dim=geom.vae_dim, encoder_n_filters=geom.n_filters, decoder_n_filters=geom.n_filters,
              encoder_ratios=list(geom.ratios), encoder_depths=depths, layernorm_eps=geom.eps)
    model = Model(cfg).eval()
    sd = model.state_dict()
    for k, v in weights.items():
        assert k in sd and tuple(sd[k].shape) == tuple(v.shape), (k, v.shape)
        sd[k] = torch.from_numpy(np.ascontiguousarray(v))
    model.load_state_dict(sd)
    return model


# --------------------------------------------------------------------------------------------------------------------
# the diffusion-head sampler.  vibevoice/schedule/dpm_solver.py needs `diffusers` (absent here, no network)
# for four base-class / helper names only; the stand-in below provides exactly those, so that the UNMODIFIED reference
# scheduler and head run in the build container.  modeling_vibevoice_inference.py does not import under transformers 5
# (it wants transformers.models.qwen2.tokenization_qwen2_fast), so sample_speech_tokens (:695-708) is lifted out of its
# source with `ast` and executed unmodified against a stub `self` (same technique as oracle/gen_golden_pcm16.py).
# --------------------------------------------------------------------------------------------------------------------
def install_diffusers_stub() -> None:
    """Minimal stand-in for the names dpm_solver.py:23-26 imports (no-op if a real diffusers is importable)."""
    import dataclasses
    import enum
    import functools
    import inspect
    import types
    try:
        import diffusers  # noqa: F401
        return
    except ImportError:
        pass
    import torch

    def mod(name, **kw):
        m = types.ModuleType(name)
        m.__dict__.update(kw)
        sys.modules[name] = m
        return m

    class ConfigMixin:
        config_name = "scheduler_config.json"

    def register_to_config(init):
        @functools.wraps(init)
        def wrapped(self, *a, **k):
            ba = inspect.signature(init).bind(self, *a, **k)
            ba.apply_defaults()
            self.config = types.SimpleNamespace(**{n: v for n, v in ba.arguments.items() if n != "self"})
            self.register_to_config = lambda **kw: self.config.__dict__.update(kw)
            init(self, *a, **k)
        return wrapped

    class SchedulerMixin:
        pass

    @dataclasses.dataclass
    class SchedulerOutput:
        prev_sample: object

    class KarrasDiffusionSchedulers(enum.Enum):
        DPMSolverMultistepScheduler = 1

    def randn_tensor(shape, generator=None, device=None, dtype=None):
        return torch.randn(shape, generator=generator, device=device, dtype=dtype)

    mod("diffusers")
    mod("diffusers.configuration_utils", ConfigMixin=ConfigMixin, register_to_config=register_to_config)
    mod("diffusers.utils", deprecate=lambda *a, **k: None)
    mod("diffusers.utils.torch_utils", randn_tensor=randn_tensor)
    mod("diffusers.schedulers")
    mod("diffusers.schedulers.scheduling_utils", KarrasDiffusionSchedulers=KarrasDiffusionSchedulers,
        SchedulerMixin=SchedulerMixin, SchedulerOutput=SchedulerOutput)


def import_reference_sampler():
    """-> (DPMSolverMultistepScheduler, VibeVoiceDiffusionHead, VibeVoiceDiffusionHeadConfig, sample_speech_tokens)
    where sample_speech_tokens is the reference METHOD's own code (modeling_vibevoice_inference.py:695-708), to be called
    as f(self_stub, condition, neg_condition, cfg_scale)."""
    import ast
    import_reference()
    install_diffusers_stub()
    from vibevoice.schedule.dpm_solver import DPMSolverMultistepScheduler
    from vibevoice.modular.configuration_vibevoice import VibeVoiceDiffusionHeadConfig
    from vibevoice.modular.modular_vibevoice_diffusion_head import VibeVoiceDiffusionHead
    path = os.path.join(REFERENCE_ROOT, "vibevoice", "modular", "modeling_vibevoice_inference.py")
    tree = ast.parse(open(path).read())
    fn = None
    for node in ast.walk(tree):
        if isinstance(node, ast.FunctionDef) and node.name == "sample_speech_tokens":
            fn = node
    if fn is None:
        raise RuntimeError("sample_speech_tokens not found in the reference")
    fn.decorator_list = []                                 # @torch.no_grad(): applied by the caller
    import torch
    ns = {"torch": torch}
    exec(compile(ast.Module(body=[fn], type_ignores=[]), path, "exec"), ns)
    return DPMSolverMultistepScheduler, VibeVoiceDiffusionHead, VibeVoiceDiffusionHeadConfig, ns["sample_speech_tokens"]
