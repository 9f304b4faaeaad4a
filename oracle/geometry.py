"""Decoder geometry shared by the oracle, the golden generator and the tests.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is imported by the product
package (``vibevoice-rocm_b200/vvdec``); only ``tests/``, ``__graft_entry__.smoke()``
and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs use it.

Restates the shape bookkeeping of the reference decoder:
  * hyper-parameters  — vibevoice/modular/configuration_vibevoice.py:16-44
  * depth reversal    — vibevoice/modular/modular_vibevoice_tokenizer.py:1024-1028
  * channel schedule  — modular_vibevoice_tokenizer.py:864-880 (stem, transposed convs),
                        :900-906 (stages), :912 (head)
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Tuple


@dataclass(frozen=True)
class DecoderGeometry:
    vae_dim: int = 64
    n_filters: int = 32
    ratios: Tuple[int, ...] = (8, 5, 5, 4, 2, 2)
    depths: Tuple[int, ...] = (8, 3, 3, 3, 3, 3, 3)   # already reversed (":1028")
    channels: int = 1
    kernel_size: int = 7
    eps: float = 1e-5
    ffn_expansion: int = 4

    @property
    def n_stages(self) -> int:
        return len(self.depths)

    def stage_channels(self, i: int) -> int:
        # modular_vibevoice_tokenizer.py:901
        return self.n_filters * (2 ** (self.n_stages - 1 - i))

    def stage_rate(self, i: int) -> int:
        """time-steps per latent frame at stage i"""
        r = 1
        for k in range(i):
            r *= self.ratios[k]
        return r

    @property
    def hop(self) -> int:
        return self.stage_rate(self.n_stages - 1)

    def tensor_specs(self) -> List[Tuple[str, Tuple[int, ...], str]]:
        """(state-dict name under the tokenizer model, shape, kind) in a canonical order.

        kinds: conv_w, convtr_w, linear_w, bias, norm_w, gamma
        Names verified against ``named_parameters()`` of the reference.
        """
        specs: List[Tuple[str, Tuple[int, ...], str]] = []
        k = self.kernel_size
        c0 = self.stage_channels(0)
        specs.append(("decoder.upsample_layers.0.0.conv.conv.weight", (c0, self.vae_dim, k), "conv_w"))
        specs.append(("decoder.upsample_layers.0.0.conv.conv.bias", (c0,), "bias"))
        for i in range(self.n_stages):
            c = self.stage_channels(i)
            if i > 0:
                cin = self.stage_channels(i - 1)
                r = self.ratios[i - 1]
                p = f"decoder.upsample_layers.{i}.0.convtr.convtr."
                specs.append((p + "weight", (cin, c, 2 * r), "convtr_w"))
                specs.append((p + "bias", (c,), "bias"))
            for j in range(self.depths[i]):
                p = f"decoder.stages.{i}.{j}."
                specs.append((p + "gamma", (c,), "gamma"))
                specs.append((p + "ffn_gamma", (c,), "gamma"))
                specs.append((p + "norm.weight", (c,), "norm_w"))
                specs.append((p + "ffn_norm.weight", (c,), "norm_w"))
                specs.append((p + "mixer.conv.conv.conv.weight", (c, 1, k), "conv_w"))
                specs.append((p + "mixer.conv.conv.conv.bias", (c,), "bias"))
                specs.append((p + "ffn.linear1.weight", (self.ffn_expansion * c, c), "linear_w"))
                specs.append((p + "ffn.linear1.bias", (self.ffn_expansion * c,), "bias"))
                specs.append((p + "ffn.linear2.weight", (c, self.ffn_expansion * c), "linear_w"))
                specs.append((p + "ffn.linear2.bias", (c,), "bias"))
        cl = self.stage_channels(self.n_stages - 1)
        specs.append(("decoder.head.conv.conv.weight", (self.channels, cl, k), "conv_w"))
        specs.append(("decoder.head.conv.conv.bias", (self.channels,), "bias"))
        return specs

    # ---- encoder (mirror stack, TokenizerEncoder :687-813) --------------------------------------------
    def enc_stage_channels(self, i: int) -> int:
        """channels of encoder stage i (:737-738, :765): n_filters * 2^i = decoder stage n-1-i"""
        return self.n_filters * (2 ** i)

    def enc_depth(self, i: int) -> int:
        """encoder depths are the decoder's reversed (:1019-1028)"""
        return self.depths[self.n_stages - 1 - i]

    def enc_ratio(self, i: int) -> int:
        """stride of the downsampling conv INTO encoder stage i >= 1: reversed(config.ratios)[i-1] (:701)"""
        return self.ratios[self.n_stages - 1 - i]

    def encoder_tensor_specs(self) -> List[Tuple[str, Tuple[int, ...], str]]:
        """(state-dict name, shape, kind) of every encoder parameter, canonical order (names verified against
        ``named_parameters()`` of the reference)."""
        specs: List[Tuple[str, Tuple[int, ...], str]] = []
        k = self.kernel_size
        specs.append(("encoder.downsample_layers.0.0.conv.conv.weight", (self.n_filters, self.channels, k), "conv_w"))
        specs.append(("encoder.downsample_layers.0.0.conv.conv.bias", (self.n_filters,), "bias"))
        for i in range(self.n_stages):
            c = self.enc_stage_channels(i)
            if i > 0:
                r = self.enc_ratio(i)
                p = f"encoder.downsample_layers.{i}.0.conv.conv."
                specs.append((p + "weight", (c, c // 2, 2 * r), "conv_w"))
                specs.append((p + "bias", (c,), "bias"))
            for j in range(self.enc_depth(i)):
                p = f"encoder.stages.{i}.{j}."
                specs.append((p + "gamma", (c,), "gamma"))
                specs.append((p + "ffn_gamma", (c,), "gamma"))
                specs.append((p + "norm.weight", (c,), "norm_w"))
                specs.append((p + "ffn_norm.weight", (c,), "norm_w"))
                specs.append((p + "mixer.conv.conv.conv.weight", (c, 1, k), "conv_w"))
                specs.append((p + "mixer.conv.conv.conv.bias", (c,), "bias"))
                specs.append((p + "ffn.linear1.weight", (self.ffn_expansion * c, c), "linear_w"))
                specs.append((p + "ffn.linear1.bias", (self.ffn_expansion * c,), "bias"))
                specs.append((p + "ffn.linear2.weight", (c, self.ffn_expansion * c), "linear_w"))
                specs.append((p + "ffn.linear2.bias", (c,), "bias"))
        cl = self.enc_stage_channels(self.n_stages - 1)
        specs.append(("encoder.head.conv.conv.weight", (self.vae_dim, cl, k), "conv_w"))
        specs.append(("encoder.head.conv.conv.bias", (self.vae_dim,), "bias"))
        return specs

    def receptive_field_frames(self) -> int:
        """Left context, in latent frames, that influences an output sample.

        Sum over layers of (left taps) / (layer input rate); every SConv1d on the decoder
        path is stride 1 with k-1 left taps (:282), every transposed conv has k = 2*stride
        so exactly one previous input column contributes (:875-878 with :563-571).
        """
        import math
        k1 = self.kernel_size - 1
        ctx = float(k1)                       # stem at rate 1
        for i in range(self.n_stages):
            rate = self.stage_rate(i)
            if i > 0:
                ctx += 1.0 / self.stage_rate(i - 1)
            ctx += self.depths[i] * k1 / rate
        ctx += k1 / self.hop                  # head
        return int(math.ceil(ctx - 1e-9))


FULL = DecoderGeometry()
# A reduced geometry the reference itself can build (decoder_n_filters=4, shallower) — used for
# fast CPU tests and for exercising generic-C code paths.  Same ratios so the hop stays 3200.
TINY = DecoderGeometry(vae_dim=16, n_filters=4, ratios=(8, 5, 5, 4, 2, 2), depths=(2, 1, 1, 1, 1, 1, 2))
