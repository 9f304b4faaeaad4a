"""Per-kernel parity of the kernels the ENCODER path adds, called alone through the C ABI against the numpy oracle of
the same reference op: the stem SConv1d(1 -> 32, k7), the strided SConv1d(C/2 -> C, k = 2r, stride r) as a 2-tap GEMM
over the padded bf16 operand copy (ragged lengths exercise the right padding), and the head SConv1d(2048 -> 64, k7)."""
import ctypes as C

import numpy as np
import pytest

from oracle.geometry import DecoderGeometry, FULL
from oracle import vv_oracle as O
from oracle.weights import make_audio, make_encoder_weights, round_to_bf16

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def _rng(seed):
    return np.random.Generator(np.random.Philox(key=[np.uint64(77), np.uint64(seed)]))


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


@pytest.fixture(scope="module")
def enc_pair():
    import vvdec
    geom = DecoderGeometry(depths=(1, 1, 1, 1, 1, 1, 1))           # the convs under test do not depend on the depths
    w = make_encoder_weights(geom, "stress", 31)
    enc = vvdec.B200AcousticEncoder(vvdec.Geometry(depths=geom.depths), device=0)
    enc.load_state_dict({k: torch.from_numpy(v) for k, v in w.items()})
    return enc, w


@pytest.mark.parametrize("B,L,dt", [(2, 5000, "f32"), (1, 1, "f32"), (3, 7, "bf16"), (1, 100003, "bf16")])
def test_encoder_stem(enc_pair, B, L, dt):
    from vvdec import _capi
    enc, w = enc_pair
    audio = make_audio(B, L, 60 + L % 17)                          # bf16-representable
    ref = O.sconv1d(audio, w["encoder.downsample_layers.0.0.conv.conv.weight"],
                    w["encoder.downsample_layers.0.0.conv.conv.bias"])[0]                 # [B, 32, L]
    a = torch.from_numpy(audio[:, 0]).cuda().to(torch.bfloat16 if dt == "bf16" else torch.float32).contiguous()
    x = torch.empty((B, L, 32), dtype=torch.float32, device="cuda")
    _capi.check(enc._lib.vv_kernel_enc_stem(enc._h, _ptr(a), 1 if dt == "bf16" else 0, _ptr(x), B, L, _stream()),
                "vv_kernel_enc_stem")
    torch.cuda.synchronize()
    m = O.parity_metrics(x.cpu().numpy().transpose(0, 2, 1), ref)
    assert m["rel_max"] < 1e-5, m                                  # fp32 FMAs on the same operands: order only


@pytest.mark.parametrize("stage,B,rows", [(1, 2, 300), (1, 1, 1), (2, 3, 33), (3, 2, 41), (4, 1, 130), (5, 2, 23),
                                          (6, 1, 17), (6, 2, 8), (3, 1, 4000)])
def test_strided_conv_gemm(enc_pair, stage, B, rows):
    from vvdec import _capi
    enc, w = enc_pair
    r, Cout = FULL.enc_ratio(stage), FULL.enc_stage_channels(stage)
    Cin = Cout // 2
    x = round_to_bf16(_rng(700 + stage * 11 + rows).standard_normal((B, rows, Cin)).astype(np.float32))
    p = f"encoder.downsample_layers.{stage}.0.conv.conv."
    ref = O.sconv1d_strided(x.transpose(0, 2, 1), w[p + "weight"], w[p + "bias"], r)     # [B, Cout, ceil(rows / r)]
    rows_out = -(-rows // r)
    assert ref.shape == (B, Cout, rows_out)
    xb = torch.full((B, (rows_out + 1) * r, Cin), 5.0, dtype=torch.bfloat16, device="cuda")   # pads: garbage on entry
    xb[:, r:r + rows] = torch.from_numpy(x).cuda().to(torch.bfloat16)
    y = torch.empty((B, rows_out, Cout), dtype=torch.float32, device="cuda")
    _capi.check(enc._lib.vv_kernel_downsample(enc._h, stage, _ptr(xb), _ptr(y), B, rows, _stream()), "vv_kernel_downsample")
    torch.cuda.synchronize()
    m = O.parity_metrics(y.cpu().numpy().transpose(0, 2, 1), ref)
    assert m["rel_max"] < 5e-5, m                                  # bf16 products are exact in fp32; order only


@pytest.mark.parametrize("B,rows", [(2, 9), (1, 1), (64, 3), (1, 300)])
def test_encoder_head(enc_pair, B, rows):
    from vvdec import _capi
    enc, w = enc_pair
    C_last = FULL.enc_stage_channels(6)
    x = round_to_bf16(_rng(900 + rows).standard_normal((B, rows, C_last)).astype(np.float32))
    ref = O.sconv1d(x.transpose(0, 2, 1), w["encoder.head.conv.conv.weight"], w["encoder.head.conv.conv.bias"])[0]
    xb = torch.full((B, rows + 6, C_last), -3.0, dtype=torch.bfloat16, device="cuda")
    xb[:, 6:] = torch.from_numpy(x).cuda().to(torch.bfloat16)
    mean = torch.empty((B, rows, 64), dtype=torch.float32, device="cuda")
    _capi.check(enc._lib.vv_kernel_enc_head(enc._h, _ptr(xb), _ptr(mean), B, rows, _stream()), "vv_kernel_enc_head")
    torch.cuda.synchronize()
    m = O.parity_metrics(mean.cpu().numpy().transpose(0, 2, 1), ref)
    assert m["rel_max"] < 5e-5, m
