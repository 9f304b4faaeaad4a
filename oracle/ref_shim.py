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
