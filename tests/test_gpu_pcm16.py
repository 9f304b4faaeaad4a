"""vv_pcm16 through the C ABI on the GPU: bit-exact against the reference's own outputs (golden) and
against the numpy restatement on aligned / unaligned / strided / long / bf16 inputs; the streamer's queue contract."""
import ctypes as C
import glob
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN, load_golden
from oracle.pcm16 import make_wave, pcm16_rows

pytestmark = pytest.mark.gpu
CASES = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "pcm16_*.npz")))


@pytest.fixture(scope="module")
def vv():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    import vvdec
    return vvdec


@pytest.mark.parametrize("name", CASES)
def test_golden_bit_exact(vv, name):
    z, meta = load_golden(name)
    x = torch.from_numpy(make_wave(meta["B"], meta["L"], meta["seed"], meta["scale"], meta["bf16"])).cuda()
    if meta["bf16"]:
        x = x.to(torch.bfloat16)                      # exact: the values are bf16-representable
    got, peak = vv.pcm16(x, return_peak=True)
    assert got.dtype == torch.int16
    assert np.array_equal(got.cpu().numpy(), z["pcm"])
    assert np.array_equal(peak.cpu().numpy(), np.abs(x.float().cpu().numpy()).max(axis=1))


@pytest.mark.parametrize("B,L,scale", [(1, 1, 3.0), (3, 7, 0.5), (2, 3199, 2.0), (5, 8192, 4.0), (2, 8193, 4.0),
                                        (3, 16384, 1.0), (2, 100001, 9.0), (7, 3200, 0.9)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_matches_oracle(vv, B, L, scale, dtype):
    x = make_wave(B, L, 100 + B + L, scale, bf16=(dtype == torch.bfloat16))
    want, peaks = pcm16_rows(x)
    got, peak = vv.pcm16(torch.from_numpy(x).cuda().to(dtype), return_peak=True)
    assert np.array_equal(got.cpu().numpy(), want)
    assert np.array_equal(peak.cpu().numpy(), peaks)


def test_layouts_and_unaligned_views(vv):
    x = make_wave(4, 6405, 7, 2.5)
    xt = torch.from_numpy(x).cuda()
    want, _ = pcm16_rows(x)
    assert np.array_equal(vv.pcm16(xt.unsqueeze(1)).cpu().numpy(), want[:, None, :])          # decode() output shape
    assert np.array_equal(vv.pcm16(xt[0]).cpu().numpy(), want[0])                             # [L]
    v = xt[:, 3:3203]                                                                          # strided rows, odd offset
    assert np.array_equal(vv.pcm16(v).cpu().numpy(), pcm16_rows(x[:, 3:3203])[0])
    big = torch.from_numpy(make_wave(2, 50000, 8, 3.0)).cuda()
    assert np.array_equal(vv.pcm16(big[:, 1:40002]).cpu().numpy(), pcm16_rows(big[:, 1:40002].cpu().numpy())[0])
    assert vv.pcm16(xt[:0]).shape == (0, 6405)


def test_errors(vv):
    with pytest.raises(RuntimeError):
        vv.pcm16(torch.zeros(4, 8))                                                            # CPU tensor: no fallback
    with pytest.raises(ValueError):
        vv.pcm16(torch.zeros(4, 8, device="cuda", dtype=torch.float16))
    lib = vv._capi.load()
    x = torch.zeros(2, 9000, device="cuda")
    out = torch.empty(2, 9000, device="cuda", dtype=torch.int16)
    rc = lib.vv_pcm16(C.c_void_p(x.data_ptr()), 0, 2, 9000, 9000, C.c_void_p(out.data_ptr()), None, None)
    assert rc == -1 and b"peak" in lib.vv_last_error()                                         # long rows need the peak buffer
    rc = lib.vv_pcm16(C.c_void_p(x.data_ptr()), 7, 2, 9000, 9000, C.c_void_p(out.data_ptr()), None, None)
    assert rc == -3


def test_full_size_rows(vv):
    """K2-sized rows (60 s = 1 440 000 samples) in bf16, 8 utterances: still bit-exact against numpy."""
    g = torch.Generator(device="cuda").manual_seed(5)
    x = (torch.randn(8, 1, 1_440_000, device="cuda", generator=g) * 4.0).to(torch.bfloat16)
    got = vv.pcm16(x).cpu().numpy()[:, 0]
    want, _ = pcm16_rows(x.float().cpu().numpy()[:, 0])
    assert np.array_equal(got, want)


def test_streamer_contract(vv):
    """AudioStreamer's queue semantics (streamer.py:41-83) with int16 chunks: routing by sample index, finished
    samples dropped, stop signal, per-sample and batch iteration."""
    st = vv.B200PcmStreamer(batch_size=3, stop_signal=None, timeout=5.0)
    x = make_wave(3, 3200, 21, 2.0)
    want, _ = pcm16_rows(x)
    xt = torch.from_numpy(x).cuda().to(torch.bfloat16)
    wantb, _ = pcm16_rows(xt.float().cpu().numpy())
    st.put(xt[[0, 2]].unsqueeze(1), torch.tensor([0, 2], device="cuda"))       # device-side indices, [n, 1, L] chunks
    st.put(torch.from_numpy(x).cuda()[[1]], [1])
    st.end([2])
    st.put(xt[[2, 0]], [2, 0])                                                  # sample 2 has ended: dropped
    st.end()
    a = list(st.get_stream(0))
    assert len(a) == 2 and a[0].dtype == np.int16 and np.array_equal(a[0], wantb[0]) and np.array_equal(a[1], wantb[0])
    b = list(st.get_stream(1))
    assert len(b) == 1 and np.array_equal(b[0], want[1])
    c = list(st.get_stream(2))
    assert len(c) == 1 and np.array_equal(c[0], wantb[2])
    assert st.puts == 3 and st.d2h_bytes == (2 + 1 + 2) * 3200 * 2
    with pytest.raises(ValueError):
        st.get_stream(3)
    st2 = vv.B200PcmStreamer(batch_size=2, timeout=5.0)
    st2.put(xt[:2], [0, 1])
    st2.end()
    batches = list(st2)
    assert len(batches) == 1 and sorted(batches[0]) == [0, 1] and np.array_equal(batches[0][1], wantb[1])
