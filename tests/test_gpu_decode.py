"""End-to-end parity of the CUDA decode path (through the Python mirror of decode(), i.e. the C ABI)
against (a) the golden outputs of the UNMODIFIED reference (tests/golden, oracle/gen_golden.py) and
(b) size-independent properties at larger sizes.  Tolerance = BASELINE.json north_star:
max-abs <= 2e-2 of the reference peak and SNR >= 40 dB vs the fp32 reference."""
import numpy as np
import pytest

from conftest import load_golden, geom_from_meta
from oracle.geometry import FULL
from oracle.vv_oracle import parity_metrics
from oracle.weights import make_latents

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

SNR_MIN, REL_MAX = 40.0, 2e-2


def _assert_parity(y, ref, what=""):
    m = parity_metrics(y, ref)
    assert m["snr_db"] >= SNR_MIN and m["rel_max"] <= REL_MAX, (what, m)
    return m


def _decoder_for(meta, weight_cache, **kw):
    import vvdec
    geom = geom_from_meta(meta)
    w = weight_cache(geom, meta["init"], meta["wseed"])
    dec = vvdec.B200AcousticDecoder(vvdec.Geometry(vae_dim=geom.vae_dim, n_filters=geom.n_filters, ratios=geom.ratios,
                                                   depths=geom.depths, eps=geom.eps), device=0, **kw)
    dec.load_state_dict({k: torch.from_numpy(v) for k, v in w.items()})
    return dec, geom


@pytest.mark.parametrize("gemm_path", [1, 3, 2, 0])
def test_offline_stress_T6_matches_reference(weight_cache, gemm_path):
    """gemm_path=1 (CUDA-core validation GEMM) isolates every non-tensor-core kernel, 2 = unfused tcgen05 GEMMs,
    0 = the product path (fused FFN kernel at C <= 64)."""
    z, meta = load_golden("full_stress_T6")
    dec, geom = _decoder_for(meta, weight_cache, gemm_path=gemm_path)
    lat = make_latents(meta["B"], meta["T"], geom.vae_dim, meta["lseed"])
    for dt in (torch.float32, torch.bfloat16):
        y = dec.decode(torch.from_numpy(lat).cuda().to(dt))
        assert y.shape == z["audio"].shape and y.dtype == dt
        m = _assert_parity(y.float().cpu().numpy(), z["audio"], f"path{gemm_path} {dt}")
        print(f"gemm_path={gemm_path} {dt}: {m}")
    assert dec.param_count == 343_695_969


@pytest.mark.parametrize("bn256", [0, 2])
def test_offline_forced_tile_widths_match_reference(full_decoder, bn256):
    """The whole path with the 128 x 256 GEMM tile forced everywhere it exists (what the K2 headline routes to at its
    row counts: stem, every upsampler into stages 1-4, every FFN GEMM at C >= 256) and with it disabled."""
    dec, _ = full_decoder
    z, meta = load_golden("full_stress_T6")
    lat = make_latents(meta["B"], meta["T"], FULL.vae_dim, meta["lseed"])
    dec.set_option("gemm_bn256", bn256)
    try:
        y = dec.decode(torch.from_numpy(lat).cuda())
    finally:
        dec.set_option("gemm_bn256", 1)
    print(bn256, _assert_parity(y.cpu().numpy(), z["audio"], f"bn256={bn256}"))
    z2, meta2 = load_golden("full_stress_K1_T75")
    lat2 = make_latents(meta2["B"], meta2["T"], FULL.vae_dim, meta2["lseed"])
    dec.set_option("gemm_bn256", bn256)
    try:
        y2 = dec.decode(torch.from_numpy(lat2).cuda())
    finally:
        dec.set_option("gemm_bn256", 1)
    print(bn256, _assert_parity(y2.cpu().numpy(), z2["audio"], f"K1 bn256={bn256}"))


def test_offline_K1_T75_matches_reference(full_decoder):
    """BASELINE.json configs[0]: batch 1, 10 s (75 frames)."""
    dec, _ = full_decoder
    z, meta = load_golden("full_stress_K1_T75")
    lat = make_latents(meta["B"], meta["T"], FULL.vae_dim, meta["lseed"])
    y = dec.decode(torch.from_numpy(lat).cuda())
    print(_assert_parity(y.cpu().numpy(), z["audio"], "K1"))


def test_offline_default_and_mild_init_match_reference(weight_cache):
    for name in ("full_default_T6", "full_mild_T8"):
        z, meta = load_golden(name)
        dec, geom = _decoder_for(meta, weight_cache)
        lat = make_latents(meta["B"], meta["T"], geom.vae_dim, meta["lseed"])
        y = dec.decode(torch.from_numpy(lat).cuda())
        print(name, _assert_parity(y.cpu().numpy(), z["audio"], name))
        del dec


def test_layout_bdt_and_ambiguity(full_decoder):
    """[B, D, T] is accepted as is (:1102-1105); [B, T, D] with T == D is READ as [B, D, T], like the reference."""
    dec, _ = full_decoder
    lat = torch.from_numpy(make_latents(1, 5, 64, 31)).cuda()
    a = dec.decode(lat)
    b = dec.decode(lat.permute(0, 2, 1).contiguous())
    assert torch.equal(a, b)
    sq = torch.from_numpy(make_latents(1, 64, 64, 32)).cuda()          # T == vae_dim
    y = dec.decode(sq)
    y2 = dec.decode(sq.permute(0, 2, 1).contiguous())                   # the other reading
    assert y.shape == (1, 1, 64 * 3200) and not torch.equal(y, y2)
    assert torch.equal(y, dec.decode(sq.clone()))


def test_streaming_replay_matches_reference(full_decoder):
    """Replays the reference call pattern (late joiner -> all-or-nothing reset :204-207, set_to_zero :234-241,
    multi-frame chunk) recorded in tests/golden/full_stress_stream.npz, in strict-quirk mode."""
    dec, _ = full_decoder
    z, meta = load_golden("full_stress_stream")
    sched = meta["schedule"]
    n_dec = sum(1 for s in sched if s[0] == "decode")
    maxT = max(s[2] for s in sched if s[0] == "decode")
    lat = make_latents(n_dec * meta["n_streams"], maxT, FULL.vae_dim, meta["lseed"]).reshape(
        n_dec, meta["n_streams"], maxT, FULL.vae_dim)
    dec.strict_reference_quirks = True
    try:
        cache = dec.new_cache(8)
        d = 0
        for s, step in enumerate(sched):
            if step[0] == "zero":
                cache.set_to_zero(step[1])
                continue
            idx, T = step[1], step[2]
            x = torch.from_numpy(np.ascontiguousarray(lat[d, idx, :T])).cuda()
            y = dec.decode(x, cache=cache, sample_indices=torch.tensor(idx), use_cache=True)
            _assert_parity(y.cpu().numpy(), z[f"out:{s}"], f"step {s}")
            d += 1
    finally:
        dec.strict_reference_quirks = False


def test_streaming_equals_offline_and_uneven_chunks(full_decoder):
    dec, _ = full_decoder
    lat = torch.from_numpy(make_latents(3, 14, 64, 41)).cuda()
    off = dec.decode(lat)
    cache = dec.new_cache(16)
    idx = torch.tensor([5, 2, 9], device="cuda")
    outs, t = [], 0
    for n in (1, 1, 3, 2, 7):                       # frame-by-frame and multi-frame chunks (T < 6, = 7)
        outs.append(dec.decode(lat[:, t:t + n], cache=cache, sample_indices=idx, use_cache=True))
        t += n
    st = torch.cat(outs, -1)
    # identical kernels, identical summation order: streaming must equal offline bit for bit
    assert torch.equal(st, off), parity_metrics(st.cpu().numpy(), off.cpu().numpy())


def test_both_block_kernels_stream_equal_offline(full_decoder):
    """block_kernel 0 (block_fused_kernel) and 1 (block_tma_kernel: halo rows patched from / saved to the stream state by
    its S1 warps) each reproduce their own offline decode bit for bit when streamed in uneven chunks.  They do NOT agree
    with each other to fp32 rounding: block_tma folds ffn_norm_w into W1 and ffn_gamma into W2 BEFORE the bf16 rounding of
    the weights, so its GEMM operands differ from block_fused's by one bf16 rounding (2^-9 relative).  Measured on B200:
    54.1 dB between the two end-to-end outputs, each of which is 44-47 dB from the fp32 reference; the bound below is the
    tolerance of the path itself (BASELINE.json: 40 dB) plus margin, not a rounding-noise bound."""
    dec, _ = full_decoder
    lat = torch.from_numpy(make_latents(2, 9, 64, 47)).cuda()
    outs = {}
    try:
        for k in (0, 1):
            dec.set_option("block_kernel", k)
            off = dec.decode(lat)
            cache = dec.new_cache(4)
            idx = torch.tensor([3, 1], device="cuda")
            parts, t = [], 0
            for n in (1, 2, 1, 5):
                parts.append(dec.decode(lat[:, t:t + n], cache=cache, sample_indices=idx, use_cache=True))
                t += n
            assert torch.equal(torch.cat(parts, -1), off), k
            outs[k] = off
    finally:
        dec.set_option("block_kernel", 1)
    m = parity_metrics(outs[1].cpu().numpy(), outs[0].cpu().numpy())
    assert m["snr_db"] > 48.0 and m["rel_max"] < 1e-2, m


def test_block_ws_on_and_off_stream_equal_offline(full_decoder):
    """Stage 4 (C = 128) as one kernel (block_ws, the default) and as mixer_kernel + ffn_stream_kernel (block_ws 0): each
    reproduces its own offline decode bit for bit when streamed in uneven chunks, and the two agree to the path's tolerance
    (block_ws folds ffn_norm_w / ffn_gamma into the bf16 weights like block_tma_kernel, see the test above)."""
    dec, _ = full_decoder
    lat = torch.from_numpy(make_latents(2, 9, 64, 53)).cuda()
    outs = {}
    try:
        for k in (0, 1):
            dec.set_option("block_ws", k)
            off = dec.decode(lat)
            cache = dec.new_cache(4)
            idx = torch.tensor([0, 2], device="cuda")
            parts, t = [], 0
            for n in (2, 1, 1, 5):
                parts.append(dec.decode(lat[:, t:t + n], cache=cache, sample_indices=idx, use_cache=True))
                t += n
            assert torch.equal(torch.cat(parts, -1), off), k
            outs[k] = off
    finally:
        dec.set_option("block_ws", 1)
    m = parity_metrics(outs[1].cpu().numpy(), outs[0].cpu().numpy())
    assert m["snr_db"] > 48.0 and m["rel_max"] < 1e-2, m


def test_ffn_stream256_on_and_off_stream_equal_offline(full_decoder):
    """Stage 3's FFN half as ffn_stream_kernel<256> (default) and as two GEMMs: each setting reproduces its own offline
    decode bit for bit when streamed, and the two agree to well inside the path's tolerance (same bf16 operands; only the
    order of the fp32 operations around the hidden activation differs)."""
    dec, _ = full_decoder
    lat = torch.from_numpy(make_latents(2, 8, 64, 59)).cuda()
    outs = {}
    try:
        for k in (0, 1):
            dec.set_option("ffn_stream256", k)
            off = dec.decode(lat)
            cache = dec.new_cache(3)
            idx = torch.tensor([2, 0], device="cuda")
            parts, t = [], 0
            for n in (1, 3, 4):
                parts.append(dec.decode(lat[:, t:t + n], cache=cache, sample_indices=idx, use_cache=True))
                t += n
            assert torch.equal(torch.cat(parts, -1), off), k
            outs[k] = off
    finally:
        dec.set_option("ffn_stream256", 1)
    m = parity_metrics(outs[1].cpu().numpy(), outs[0].cpu().numpy())
    assert m["snr_db"] > 55.0 and m["rel_max"] < 5e-3, m


def test_state_zero_export_import(full_decoder):
    dec, _ = full_decoder
    lat = torch.from_numpy(make_latents(1, 4, 64, 43)).cuda()
    cache = dec.new_cache(4)
    a0 = dec.decode(lat[:, :2], cache=cache, sample_indices=[1], use_cache=True)
    blob = cache.export_state(1)
    a1 = dec.decode(lat[:, 2:], cache=cache, sample_indices=[1], use_cache=True)
    cache.import_state(3, blob)                                        # resume the same stream in another slot
    b1 = dec.decode(lat[:, 2:], cache=cache, sample_indices=[3], use_cache=True)
    assert torch.equal(a1, b1)
    cache.set_to_zero([1])
    c0 = dec.decode(lat[:, :2], cache=cache, sample_indices=[1], use_cache=True)
    assert torch.equal(c0, a0)                                          # zeroed slot == first chunk
    del a0


def test_halo_chunk_equals_whole(full_decoder):
    """Time-chunk sharding: frames [s-H, e) minus the first H*hop samples == the monolithic decode."""
    dec, _ = full_decoder
    H = dec.halo_frames
    assert H == 58
    T = H + 30
    lat = torch.from_numpy(make_latents(1, T, 64, 44)).cuda()
    whole = dec.decode(lat)
    s = H + 7
    part = dec.decode(lat[:, s - H:])[:, :, H * 3200:]
    m = parity_metrics(part.cpu().numpy(), whole[:, :, s * 3200:].cpu().numpy())
    assert m["snr_db"] > 90.0, m                    # fp32 residual stream: far-tap leakage only
    short = dec.decode(lat[:, s - 20:])[:, :, 20 * 3200:]
    m2 = parity_metrics(short.cpu().numpy(), whole[:, :, s * 3200:].cpu().numpy())
    assert m2["snr_db"] < m["snr_db"]               # a too-short halo is visibly worse


def test_workspace_chunking_is_invisible(full_decoder):
    """A small workspace forces utterance- and time-chunked passes (state carried on device): same output."""
    import vvdec
    dec, _ = full_decoder
    lat = torch.from_numpy(make_latents(3, 9, 64, 45)).cuda()
    ref = dec.decode(lat)
    saved, dec._ws = dec.max_workspace_bytes, None
    try:
        dec.max_workspace_bytes = dec.workspace_bytes(1, 9) + 4096            # one utterance per pass
        a = dec.decode(lat)
        dec._ws = None
        dec.max_workspace_bytes = dec.workspace_bytes(1, 4) + (4 << 20)       # time chunks of >= 4 frames
        b = dec.decode(lat)
    finally:
        dec.max_workspace_bytes, dec._ws = saved, None
    assert torch.equal(a, ref) and torch.equal(b, ref)


def test_errors_are_loud(full_decoder):
    import vvdec
    dec, _ = full_decoder
    with pytest.raises(ValueError):
        dec.decode(torch.zeros(1, 5, 7, device="cuda"))
    with pytest.raises(RuntimeError):
        dec.decode(torch.zeros(1, 5, 64))
    with pytest.raises(AssertionError):
        dec.decode(torch.zeros(2, 1, 64, device="cuda"), cache=dec.new_cache(4), sample_indices=[0], use_cache=True)
    with pytest.raises(ValueError):
        vvdec.B200AcousticDecoder(vvdec.Geometry(n_filters=4, vae_dim=16, depths=(2, 1, 1, 1, 1, 1, 2)), device=0)
