"""CPU restatement (numpy, fp32) of the reference acoustic-tokenizer decode path.

TEST INFRASTRUCTURE ONLY.  This module is the *checker* for the CUDA path.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may
import it; the product package never does (and fails loudly without its CUDA library).

Parity pin: the reference ships no tests or golden vectors, so this restatement
is pinned against outputs of the reference itself, run in the build container by
``oracle/gen_golden.py`` (imports /root/reference) and committed under ``tests/golden/``;
``tests/test_oracle_golden.py`` re-checks it on every run.

The arithmetic the reference delegates to torch ATen (``nn.Conv1d``, ``nn.ConvTranspose1d``,
``nn.Linear``, ``F.pad``, erf-GELU; torch pinned 2.5.1 in the reference's pyproject.toml:22) is
restated from its published definition.  Every function cites the reference lines it follows;
all paths are in vibevoice/modular/modular_vibevoice_tokenizer.py unless stated.
Layout follows the reference: activations are [B, C, T].
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

from .geometry import DecoderGeometry, FULL

try:  # exact erf; scipy is in the image.  math.erf fallback keeps the oracle dependency-light.
    from scipy.special import erf as _erf
except Exception:  # pragma: no cover
    _erf = np.vectorize(math.erf, otypes=[np.float32])

F32 = np.float32


# --------------------------------------------------------------------------------------
# primitives
# --------------------------------------------------------------------------------------
def rms_norm_channels(x: np.ndarray, weight: np.ndarray, eps: float) -> np.ndarray:
    """ConvRMSNorm.forward (:81-91) -> RMSNorm._norm (:65-66).

    x [B, C, T]; statistics over C in fp32: x * rsqrt(mean_C(x^2) + eps), then * weight.
    """
    ms = np.mean(x.astype(F32) ** 2, axis=1, keepdims=True, dtype=F32)
    y = x * (F32(1.0) / np.sqrt(ms + F32(eps)))
    return (y * weight[None, :, None]).astype(F32)


def gelu_erf(x: np.ndarray) -> np.ndarray:
    """ACT2FN["gelu"] = exact erf GELU (:589)."""
    return (F32(0.5) * x * (F32(1.0) + _erf(x * F32(0.7071067811865476)).astype(F32))).astype(F32)


def conv1d_valid(x: np.ndarray, w: np.ndarray, b: Optional[np.ndarray], groups: int = 1) -> np.ndarray:
    """nn.Conv1d(stride=1, padding=0, dilation=1) as used at :168/:359/:413.

    x [B, Cin, L], w [Cout, Cin/groups, k] -> [B, Cout, L-k+1]
    """
    B, Cin, L = x.shape
    Cout, cpg, k = w.shape
    To = L - k + 1
    assert To >= 1
    if groups == 1:
        y = np.zeros((B, Cout, To), dtype=F32)
        for j in range(k):
            y += np.matmul(w[:, :, j], x[:, :, j:j + To])
    else:
        assert groups == Cin == Cout and cpg == 1, "only depthwise grouping is on the path (:640-647)"
        y = np.zeros((B, Cout, To), dtype=F32)
        for j in range(k):
            y += w[None, :, 0, j, None] * x[:, :, j:j + To]
    if b is not None:
        y += b[None, :, None]
    return y


def sconv1d(x: np.ndarray, w: np.ndarray, b: np.ndarray, groups: int = 1,
            state: Optional[np.ndarray] = None) -> Tuple[np.ndarray, Optional[np.ndarray]]:
    """SConv1d, stride 1, causal, pad_mode 'constant'.

    state is None  -> _forward_non_streaming (:384-418): left-pad (k-1) zeros (extra_padding = 0
                      for stride 1, :127-133), conv.
    state [B,C,k-1]-> _forward_streaming (:327-382): concat state, conv without padding, new state =
                      last k-1 columns of [state, x].
    """
    k = w.shape[2]
    ctx = k - 1                                                     # :282 with stride 1, dilation 1
    if state is None:
        xp = np.pad(x, ((0, 0), (0, 0), (ctx, 0)))
        return conv1d_valid(xp, w, b, groups), None
    xin = np.concatenate([state, x], axis=2)                        # :350
    y = conv1d_valid(xin, w, b, groups)                             # :359
    return y, xin[:, :, xin.shape[2] - ctx:].copy()                 # :370-372


def conv_transpose1d(x: np.ndarray, w: np.ndarray, b: np.ndarray, stride: int) -> np.ndarray:
    """nn.ConvTranspose1d(Cin, Cout, k, stride) (:183): y[co, q*s + j] += sum_ci x[ci, q] w[ci, co, j]."""
    B, Cin, T = x.shape
    _, Cout, k = w.shape
    L = (T - 1) * stride + k
    y = np.zeros((B, Cout, L), dtype=F32)
    for j in range(k):
        contrib = np.matmul(w[:, :, j].T, x)                        # [B, Cout, T]
        y[:, :, j:j + (T - 1) * stride + 1:stride] += contrib
    return y + b[None, :, None]


def sconv_transpose1d(x: np.ndarray, w: np.ndarray, b: np.ndarray, stride: int,
                      state: Optional[np.ndarray] = None, streaming: bool = False
                      ) -> Tuple[np.ndarray, Optional[np.ndarray]]:
    """SConvTranspose1d, causal, trim_right_ratio = 1.

    offline   (:551-576): convtr then drop the last (k - stride) samples (all trim on the right).
    streaming (:478-549): convtr over [cached inputs, x], trim, keep the last T*stride outputs
                          (everything on the first chunk, :522-524), cache the last k-1 inputs (:539-547).
    ``state`` is the cached input [B, Cin, <=k-1] or None for "no cache entry".
    """
    k = w.shape[2]
    trim = k - stride                                               # :443, :564-565 (ceil(pt*1.0))
    if not streaming:
        y = conv_transpose1d(x, w, b, stride)
        return y[:, :, :y.shape[2] - trim], None
    T = x.shape[2]
    cached = state if state is not None else np.zeros((x.shape[0], x.shape[1], 0), dtype=F32)
    full_in = np.concatenate([cached, x], axis=2)                   # :495
    y = conv_transpose1d(full_in, w, b, stride)
    y = y[:, :, :y.shape[2] - trim]                                 # :515-516
    if cached.shape[2] != 0:                                        # :522-533
        y = y[:, :, y.shape[2] - T * stride:]
    ctx = k - 1                                                     # :447
    new_state = full_in[:, :, max(0, full_in.shape[2] - ctx):].copy()
    return y, new_state


def ffn(x_btc: np.ndarray, w1, b1, w2, b2) -> np.ndarray:
    """FFN.forward (:592-596) on [B, T, C]: linear2(gelu(linear1(x)))."""
    h = np.matmul(x_btc, w1.T) + b1
    h = gelu_erf(h)
    return (np.matmul(h, w2.T) + b2).astype(F32)


# --------------------------------------------------------------------------------------
# streaming cache — same observable semantics as VibeVoiceTokenizerStreamingCache (:193-256)
# --------------------------------------------------------------------------------------
class OracleStreamingCache:
    def __init__(self):
        self.cache: Dict[Tuple[str, int], np.ndarray] = {}

    def get(self, layer_id: str, sample_indices: Sequence[int]):
        states = []
        for idx in sample_indices:                                  # :204-209
            key = (layer_id, int(idx))
            if key not in self.cache:
                return None                                         # all-or-nothing (:206-207)
            states.append(self.cache[key])
        mx = max(s.shape[-1] for s in states)
        states = [np.pad(s, ((0, 0), (mx - s.shape[-1], 0))) for s in states]   # :213-224 left zero-pad
        return np.stack(states, axis=0)

    def set(self, layer_id: str, sample_indices: Sequence[int], states: np.ndarray):
        for i, idx in enumerate(sample_indices):                    # :230-232
            self.cache[(layer_id, int(idx))] = states[i].copy()

    def set_to_zero(self, sample_indices: Sequence[int]):
        s = set(int(i) for i in sample_indices)
        for key in list(self.cache.keys()):                         # :236-241
            if key[1] in s:
                self.cache[key] = np.zeros_like(self.cache[key])

    def clear(self):
        self.cache.clear()


# --------------------------------------------------------------------------------------
# the decoder walk
# --------------------------------------------------------------------------------------
class OracleDecoder:
    """TokenizerDecoder.forward (:948-951) + VibeVoiceAcousticTokenizerModel.decode (:1099-1108)."""

    def __init__(self, weights: Dict[str, np.ndarray], geom: DecoderGeometry = FULL):
        self.g = geom
        self.w = {k: np.ascontiguousarray(v, dtype=F32) for k, v in weights.items()}

    # -- helpers ----------------------------------------------------------------
    def _sconv(self, layer_id, x, w, b, groups, cache, idx):
        if cache is None:
            return sconv1d(x, w, b, groups)[0]
        st = cache.get(layer_id, idx)
        if st is None:                                              # :337-340 zeros context
            st = np.zeros((x.shape[0], x.shape[1], w.shape[2] - 1), dtype=F32)
        y, ns = sconv1d(x, w, b, groups, state=st)
        cache.set(layer_id, idx, ns)                                # :380
        return y

    def _sconvtr(self, layer_id, x, w, b, stride, cache, idx):
        if cache is None:
            return sconv_transpose1d(x, w, b, stride)[0]
        st = cache.get(layer_id, idx)
        y, ns = sconv_transpose1d(x, w, b, stride, state=st, streaming=True)
        cache.set(layer_id, idx, ns)                                # :547
        return y

    def _block(self, i, j, x, cache, idx, taps):
        g, w = self.g, self.w
        p = f"decoder.stages.{i}.{j}."
        # mixer half (:927-932 / :667-672)
        r = x
        h = rms_norm_channels(x, w[p + "norm.weight"], g.eps)
        h = self._sconv(p + "mixer", h, w[p + "mixer.conv.conv.conv.weight"],
                        w[p + "mixer.conv.conv.conv.bias"], x.shape[1], cache, idx)
        x = r + h * w[p + "gamma"][None, :, None]
        if taps is not None:
            taps[p + "mixer_out"] = x
        # ffn half (:935-942 / :675-682)
        r = x
        h = rms_norm_channels(x, w[p + "ffn_norm.weight"], g.eps)
        h = ffn(h.transpose(0, 2, 1), w[p + "ffn.linear1.weight"], w[p + "ffn.linear1.bias"],
                w[p + "ffn.linear2.weight"], w[p + "ffn.linear2.bias"]).transpose(0, 2, 1)
        x = r + h * w[p + "ffn_gamma"][None, :, None]
        if taps is not None:
            taps[p + "out"] = x
        return x

    # -- public -----------------------------------------------------------------
    def decode(self, latents: np.ndarray, cache: Optional[OracleStreamingCache] = None,
               sample_indices: Optional[Sequence[int]] = None, use_cache: bool = False,
               taps: Optional[dict] = None) -> np.ndarray:
        g, w = self.g, self.w
        x = np.asarray(latents, dtype=F32)
        if x.shape[1] != g.vae_dim:                                 # :1102-1105 layout sniffing
            x = x.transpose(0, 2, 1)
        if not (use_cache and cache is not None):                   # :317, :469
            cache = None
        else:
            assert sample_indices is not None and len(sample_indices) == x.shape[0]   # :322-323
        idx = None if cache is None else [int(i) for i in sample_indices]
        for i in range(g.n_stages):                                 # :915-944
            if i == 0:
                x = self._sconv("stem", x, w["decoder.upsample_layers.0.0.conv.conv.weight"],
                                w["decoder.upsample_layers.0.0.conv.conv.bias"], 1, cache, idx)
            else:
                p = f"decoder.upsample_layers.{i}.0.convtr.convtr."
                x = self._sconvtr(f"up{i}", x, w[p + "weight"], w[p + "bias"], g.ratios[i - 1], cache, idx)
            if taps is not None:
                taps[f"up{i}"] = x
            for j in range(g.depths[i]):
                x = self._block(i, j, x, cache, idx, taps)
        # self.norm is Identity (disable_last_norm, :908-911); head (:950)
        x = self._sconv("head", x, w["decoder.head.conv.conv.weight"], w["decoder.head.conv.conv.bias"],
                        1, cache, idx)
        return x


# --------------------------------------------------------------------------------------
# the encoder walk (offline form)
# --------------------------------------------------------------------------------------
def sconv1d_strided(x: np.ndarray, w: np.ndarray, b: np.ndarray, stride: int) -> np.ndarray:
    """SConv1d with stride, causal, pad_mode 'constant', non-streaming (:384-418).

    padding_total = (k-1) - (stride-1) zeros on the left (:282), extra_padding zeros on the right so the last
    window is complete (get_extra_padding_for_conv1d :127-133); then nn.Conv1d(stride).  [B,Cin,L] -> [B,Cout,ceil]
    """
    B, Cin, L = x.shape
    Cout, _, k = w.shape
    pt = (k - 1) - (stride - 1)
    n_frames = (L - k + pt) / stride + 1
    ideal = (math.ceil(n_frames) - 1) * stride + (k - pt)
    extra = ideal - L
    xp = np.pad(x, ((0, 0), (0, 0), (pt, extra)))
    To = (xp.shape[2] - k) // stride + 1
    y = np.zeros((B, Cout, To), dtype=F32)
    for j in range(k):
        y += np.matmul(w[:, :, j], xp[:, :, j:j + (To - 1) * stride + 1:stride])
    return y + b[None, :, None]


def sconv1d_strided_streaming(x: np.ndarray, w: np.ndarray, b: np.ndarray, stride: int,
                              state: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    """SConv1d._forward_streaming (:327-382) with stride: context_size = (k-1) - (stride-1) cached inputs are put in
    front of x, the conv runs without padding (trailing samples that do not fill a window are dropped by nn.Conv1d,
    :359), the new state = last context_size columns of [state, x] (:370-372)."""
    k = w.shape[2]
    xin = np.concatenate([state, x], axis=2)
    To = (xin.shape[2] - k) // stride + 1
    y = np.zeros((x.shape[0], w.shape[0], max(To, 0)), dtype=F32)
    for j in range(k):
        if To > 0:
            y += np.matmul(w[:, :, j], xin[:, :, j:j + (To - 1) * stride + 1:stride])
    ctx = state.shape[2]
    return y + b[None, :, None], xin[:, :, xin.shape[2] - ctx:].copy()


class OracleEncoder:
    """TokenizerEncoder.forward (:810-813) + VibeVoiceAcousticTokenizerModel.encode (:1081-1085), offline and streaming."""

    def __init__(self, weights: Dict[str, np.ndarray], geom: DecoderGeometry = FULL):
        self.g = geom
        self.w = {k: np.ascontiguousarray(v, dtype=F32) for k, v in weights.items()}

    def _sconv(self, layer_id, x, w, b, groups, stride, cache, idx):
        if cache is None:
            return sconv1d(x, w, b, groups)[0] if stride == 1 else sconv1d_strided(x, w, b, stride)
        ctx = (w.shape[2] - 1) - (stride - 1)                        # :282
        st = cache.get(layer_id, idx)
        if st is None:                                              # :337-340 zeros context
            st = np.zeros((x.shape[0], x.shape[1], ctx), dtype=F32)
        if stride == 1:
            y, ns = sconv1d(x, w, b, groups, state=st)
        else:
            y, ns = sconv1d_strided_streaming(x, w, b, stride, st)
        cache.set(layer_id, idx, ns)                                # :380
        return y

    def _block(self, i, j, x, taps, cache=None, idx=None):
        g, w = self.g, self.w
        p = f"encoder.stages.{i}.{j}."
        r = x                                                       # mixer half (:788-794)
        h = rms_norm_channels(x, w[p + "norm.weight"], g.eps)
        h = self._sconv(p + "mixer", h, w[p + "mixer.conv.conv.conv.weight"], w[p + "mixer.conv.conv.conv.bias"],
                        x.shape[1], 1, cache, idx)
        x = r + h * w[p + "gamma"][None, :, None]
        r = x                                                       # ffn half (:796-804)
        h = rms_norm_channels(x, w[p + "ffn_norm.weight"], g.eps)
        h = ffn(h.transpose(0, 2, 1), w[p + "ffn.linear1.weight"], w[p + "ffn.linear1.bias"],
                w[p + "ffn.linear2.weight"], w[p + "ffn.linear2.bias"]).transpose(0, 2, 1)
        x = r + h * w[p + "ffn_gamma"][None, :, None]
        if taps is not None:
            taps[p + "out"] = x
        return x

    def encode(self, audio: np.ndarray, taps: Optional[dict] = None, cache: Optional[OracleStreamingCache] = None,
               sample_indices: Optional[Sequence[int]] = None, use_cache: bool = False) -> np.ndarray:
        """audio [B, 1, L] (or [B, L]) -> mean [B, frames, vae_dim]  (:1084 permutes to [B, T, D]); frames =
        ceil(L / hop) offline, floor((ctx + L - k) / stride) + 1 per strided layer when streaming."""
        g, w = self.g, self.w
        x = np.asarray(audio, dtype=F32)
        if x.ndim == 2:
            x = x[:, None, :]
        streaming = use_cache and cache is not None
        idx = [int(i) for i in sample_indices] if streaming else None
        if streaming:
            assert len(idx) == x.shape[0]
        c = cache if streaming else None
        for i in range(g.n_stages):                                 # :775-806
            p = f"encoder.downsample_layers.{i}.0.conv.conv."
            stride = 1 if i == 0 else g.enc_ratio(i)                 # stem k7 stride 1 (:732-734); k = 2r, stride r (:737-743)
            x = self._sconv(p, x, w[p + "weight"], w[p + "bias"], 1, stride, c, idx)
            if taps is not None:
                taps[f"down{i}"] = x
            for j in range(g.enc_depth(i)):
                x = self._block(i, j, x, taps, c, idx)
        # self.norm is Identity (disable_last_norm, :767-770); head k7 (:771, :812)
        x = self._sconv("encoder.head", x, w["encoder.head.conv.conv.weight"], w["encoder.head.conv.conv.bias"], 1, 1, c, idx)
        return np.ascontiguousarray(x.transpose(0, 2, 1))


# --------------------------------------------------------------------------------------
# parity metrics (SURVEY.md §8c tolerance protocol)
# --------------------------------------------------------------------------------------
def parity_metrics(test: np.ndarray, ref: np.ndarray) -> Dict[str, float]:
    t = np.asarray(test, dtype=np.float64).ravel()
    r = np.asarray(ref, dtype=np.float64).ravel()
    d = t - r
    peak = float(np.max(np.abs(r))) or 1.0
    num = float(np.sum(r * r))
    den = float(np.sum(d * d))
    snr = float("inf") if den == 0.0 else 10.0 * math.log10(max(num, 1e-300) / den)
    return {"rel_max": float(np.max(np.abs(d))) / peak, "snr_db": snr, "peak": peak,
            "max_abs": float(np.max(np.abs(d)))}
