"""vv_pcm16_static: the allocation-free PCM16 entry point (no counterpart in the reference)
against the existing vv_pcm16, through the C ABI, on two calls in a row.

Its per-row peak scratch is a file-scope static __device__ array that every call on the device shares and that keeps the
last call's contents.  The second call below therefore uses QUIET rows (peak < 1, the no-normalise branch) right after a
call with LOUD rows in the same scratch rows: if the entry point did not clear what the peak pass (an atomicMax) needs at
zero, the stale loud peaks would survive and every sample of the second call would be scaled wrongly.
Integer output, so the bar is the one tests/test_gpu_pcm16.py uses: bit-exact.

Kept in a file of its own that pytest collects LAST; it builds nothing that other tests
own and passes both alone in a fresh process and after the rest of the suite."""
import ctypes as C

import pytest

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

MAX_ROWS = 4096                                  # VV_PCM16_STATIC_MAX_ROWS (include/vvdec.h)


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _code(dt):
    from vvdec import _capi
    return _capi.VV_BF16 if dt == torch.bfloat16 else _capi.VV_F32


def _existing(lib, wave):
    """The existing entry point, with a caller-provided peak buffer."""
    from vvdec import _capi
    B, L = wave.shape
    pcm = torch.empty((B, L), dtype=torch.int16, device="cuda")
    peak = torch.empty((B,), dtype=torch.float32, device="cuda")
    _capi.check(lib.vv_pcm16(_ptr(wave), _code(wave.dtype), B, L, L, _ptr(pcm), _ptr(peak), _stream()), "vv_pcm16")
    return pcm


def _static(lib, wave, pcm):
    from vvdec import _capi
    B, L = wave.shape
    _capi.check(lib.vv_pcm16_static(_ptr(wave), _code(wave.dtype), B, L, L, _ptr(pcm), _stream()), "vv_pcm16_static")


def _wave(B, L, amp, seed, dt):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(B, L, generator=g) * amp).to(dt).cuda().contiguous()


@pytest.mark.parametrize("L,dt", [(40003, torch.float32),       # long rows, unaligned: peak pass + convert pass
                                  (65536, torch.bfloat16),       # long rows, vector path, bf16 input
                                  (3200, torch.bfloat16)])       # one streamed chunk: the one-launch form
def test_static_scratch_matches_existing_entry_point_on_two_calls_in_a_row(L, dt):
    import vvdec  # noqa: F401  (loads the library; fails loudly if it is missing)
    from vvdec import _capi
    lib = _capi.load()
    B = 5
    loud = _wave(B, L, 3.0, 100 + L % 97, dt)                    # peaks ~ 12: normalised
    quiet = _wave(B, L, 0.05, 200 + L % 97, dt)                  # peaks ~ 0.2: NOT normalised
    want_loud, want_quiet = _existing(lib, loud), _existing(lib, quiet)
    out = torch.full((B, L), -7, dtype=torch.int16, device="cuda")           # every output buffer exists before the calls
    out2 = torch.full((B, L), -7, dtype=torch.int16, device="cuda")
    torch.cuda.synchronize()
    _static(lib, loud, out)                                      # call 1 (also resolves the scratch address on this device)
    _static(lib, quiet, out2)                                    # call 2, same stream, same scratch rows, stale loud peaks
    torch.cuda.synchronize()
    assert torch.equal(out, want_loud)
    assert torch.equal(out2, want_quiet)
    assert int(want_quiet.abs().max()) < 32767 // 2 < int(want_loud.abs().max())   # the two calls took different branches
    # a third and fourth call of the same shapes: no device memory is taken by the entry point (nothing it could have
    # allocated lazily is left to allocate by now)
    free0, _ = torch.cuda.mem_get_info()
    _static(lib, loud, out)
    _static(lib, quiet, out2)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0, (free0, free1)
    assert torch.equal(out, want_loud) and torch.equal(out2, want_quiet)


def test_static_scratch_limit_is_enforced_with_a_clear_error():
    import vvdec  # noqa: F401
    from vvdec import _capi
    lib = _capi.load()
    L = 16
    wave = torch.zeros((MAX_ROWS + 1, L), dtype=torch.float32, device="cuda")
    pcm = torch.empty((MAX_ROWS + 1, L), dtype=torch.int16, device="cuda")
    rc = lib.vv_pcm16_static(_ptr(wave), _code(wave.dtype), MAX_ROWS + 1, L, L, _ptr(pcm), _stream())
    assert rc == -2                                              # VV_ERR_BAD_SHAPE
    msg = lib.vv_last_error().decode()
    assert str(MAX_ROWS) in msg and "static scratch" in msg, msg
    with pytest.raises(ValueError):
        _capi.check(rc, "vv_pcm16_static")
    # exactly the maximum is accepted, and agrees with the existing entry point
    w = _wave(MAX_ROWS, 8200, 2.0, 7, torch.float32)
    out = torch.empty((MAX_ROWS, 8200), dtype=torch.int16, device="cuda")
    _static(lib, w, out)
    torch.cuda.synchronize()
    assert torch.equal(out, _existing(lib, w))
