"""Generate tests/golden/*.npz by running the UNMODIFIED reference (build container only).

TEST INFRASTRUCTURE ONLY.  Usage (from the repo root, in the container that has /root/reference):

    python -m oracle.gen_golden            # writes tests/golden/*.npz (~2 MB)

Each fixture records the reference's fp32 CPU output of ``VibeVoiceAcousticTokenizerModel.decode``
(modular_vibevoice_tokenizer.py:1099-1108) on weights/latents that ``oracle/weights.py`` regenerates
bit-exactly from the seeds stored in the file, so the weights themselves (1.3 GB) are never committed.
Per-layer taps (forward hooks on every upsample layer and Block1D) are stored as strided subsamples
plus sum / sum-of-squares so a failing kernel localises.
"""
from __future__ import annotations

import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle.geometry import FULL, TINY, DecoderGeometry          # noqa: E402
from oracle.weights import make_weights, make_latents, make_encoder_weights, make_audio   # noqa: E402
from oracle.ref_shim import build_reference_model, import_reference   # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def tap_summary(x: np.ndarray, n: int = 256) -> np.ndarray:
    """[n strided samples..., sum, sumsq] of the [B,C,T] activation, float64."""
    f = np.asarray(x, dtype=np.float64).ravel()
    step = max(1, f.size // n)
    return np.concatenate([f[::step][:n], np.zeros(max(0, n - f[::step][:n].size)), [f.sum(), (f * f).sum()]])


def run_with_taps(model, latents_bt_c: np.ndarray, geom: DecoderGeometry):
    import torch
    taps = {}
    hooks = []
    dec = model.decoder
    # forward_features (:914-946) calls the inner S-layers directly and inlines Block1D, so hooks on the
    # nn.Sequential / Block1D containers never fire: tap the inner conv's output, and the *inputs* of
    # ffn_norm (= mixer-half output) and of the next norm / head (= block output) with pre-hooks.
    def pre(key):
        return lambda m, a: taps.__setitem__(key, a[0].detach().numpy().copy())
    prev_out_key = None
    for i in range(geom.n_stages):
        hooks.append(dec.upsample_layers[i][0].register_forward_hook(
            lambda m, a, out, key=f"up{i}": taps.__setitem__(key, out.detach().numpy().copy())))
        for j in range(geom.depths[i]):
            blk = dec.stages[i][j]
            if prev_out_key is not None and j > 0:
                hooks.append(blk.norm.register_forward_pre_hook(pre(prev_out_key)))
            hooks.append(blk.ffn_norm.register_forward_pre_hook(pre(f"decoder.stages.{i}.{j}.mixer_out")))
            prev_out_key = f"decoder.stages.{i}.{j}.out"
        # the last block of a stage feeds the next upsample layer (or the head)
        nxt = dec.upsample_layers[i + 1][0] if i + 1 < geom.n_stages else dec.head
        hooks.append(nxt.register_forward_pre_hook(pre(prev_out_key)))
    with torch.no_grad():
        y = model.decode(torch.from_numpy(latents_bt_c)).numpy()
    for h in hooks:
        h.remove()
    return y, taps


def offline_case(name, geom, init, wseed, B, T, lseed, with_taps=True):
    t0 = time.time()
    w = make_weights(geom, init, wseed)
    model = build_reference_model(geom, w)
    lat = make_latents(B, T, geom.vae_dim, lseed)
    if with_taps:
        y, taps = run_with_taps(model, lat, geom)
    else:
        import torch
        with torch.no_grad():
            y = model.decode(torch.from_numpy(lat)).numpy()
        taps = {}
    meta = dict(kind="offline", geom=geom.__dict__, init=init, wseed=wseed, B=B, T=T, lseed=lseed,
                tap_names=sorted(taps.keys()))
    arrs = {"audio": y.astype(np.float32), "meta": np.array(json.dumps(meta))}
    for k, v in taps.items():
        arrs["tap:" + k] = tap_summary(v)
    np.savez(os.path.join(GOLD, name + ".npz"), **arrs)
    print(f"{name}: out {y.shape} peak {np.abs(y).max():.4g} rms {np.sqrt((y**2).mean()):.4g}  ({time.time()-t0:.1f}s)")
    return model, w


def streaming_case(name, geom, init, wseed, schedule, lseed, n_streams, model=None):
    """Replay a call pattern like modeling_vibevoice_inference.py:540-544,:635-641 against the reference.

    schedule: list of ["decode", [sample indices], T] | ["zero", [sample indices]].
    Latents for (step s, stream i) are rows of one seeded tensor so any replayer can regenerate them.
    """
    import torch
    _, _, Cache = import_reference()
    if model is None:
        model = build_reference_model(geom, make_weights(geom, init, wseed))
    n_dec = sum(1 for s in schedule if s[0] == "decode")
    maxT = max(s[2] for s in schedule if s[0] == "decode")
    lat = make_latents(n_dec * n_streams, maxT, geom.vae_dim, lseed).reshape(n_dec, n_streams, maxT, geom.vae_dim)
    cache = Cache()
    outs = {}
    d = 0
    for s, step in enumerate(schedule):
        if step[0] == "zero":
            cache.set_to_zero(torch.tensor(step[1]))
            continue
        idx, T = step[1], step[2]
        x = torch.from_numpy(np.ascontiguousarray(lat[d, idx, :T]))              # [n, T, D]
        with torch.no_grad():
            y = model.decode(x, cache=cache, sample_indices=torch.tensor(idx), use_cache=True)
        outs[f"out:{s}"] = y.numpy().astype(np.float32)
        d += 1
    meta = dict(kind="streaming", geom=geom.__dict__, init=init, wseed=wseed, lseed=lseed,
                n_streams=n_streams, schedule=schedule)
    np.savez(os.path.join(GOLD, name + ".npz"), meta=np.array(json.dumps(meta)), **outs)
    print(f"{name}: {len(outs)} decode calls recorded")


def encoder_case(name, geom, init, wseed, B, L, aseed):
    """Reference ``VibeVoiceAcousticTokenizerModel.encode(audio).mean`` (:1081-1085) on seeded weights / audio, plus
    tap summaries of every downsample layer's output and of every stage's output (pre-hooks on the consumer)."""
    import torch
    t0 = time.time()
    w = make_encoder_weights(geom, init, wseed)
    model = build_reference_model(geom, w)
    audio = make_audio(B, L, aseed)
    taps, hooks = {}, []
    enc = model.encoder
    for i in range(geom.n_stages):
        hooks.append(enc.downsample_layers[i][0].register_forward_hook(
            lambda m, a, out, key=f"down{i}": taps.__setitem__(key, out.detach().numpy().copy())))
        nxt = enc.downsample_layers[i + 1][0] if i + 1 < geom.n_stages else enc.head
        hooks.append(nxt.register_forward_pre_hook(
            lambda m, a, key=f"encoder.stages.{i}.{geom.enc_depth(i) - 1}.out": taps.__setitem__(key, a[0].detach().numpy().copy())))
    with torch.no_grad():
        out = model.encode(torch.from_numpy(audio))
    for h in hooks:
        h.remove()
    mean = out.mean.numpy().astype(np.float32)
    meta = dict(kind="encode", geom=geom.__dict__, init=init, wseed=wseed, B=B, L=L, aseed=aseed,
                std=float(out.std), tap_names=sorted(taps.keys()))
    arrs = {"mean": mean, "meta": np.array(json.dumps(meta))}
    for k, v in taps.items():
        arrs["tap:" + k] = tap_summary(v)
    np.savez(os.path.join(GOLD, name + ".npz"), **arrs)
    print(f"{name}: mean {mean.shape} peak {np.abs(mean).max():.4g} rms {np.sqrt((mean**2).mean()):.4g}  ({time.time()-t0:.1f}s)")


def encoder_stream_case(name, geom, init, wseed, n_streams, schedule, aseed):
    """Replay streaming encode calls (modeling_vibevoice_inference.py:656-662 pattern) against the reference.
    schedule: list of ["encode", [sample indices], hops] | ["zero", [sample indices]]; call s of stream i reads the
    next `hops` * hop samples of that stream's seeded audio."""
    import torch
    _, _, Cache = import_reference()
    w = make_encoder_weights(geom, init, wseed)
    model = build_reference_model(geom, w)
    total = sum(st[2] for st in schedule if st[0] == "encode") * geom.hop
    audio = make_audio(n_streams, total, aseed)
    pos = [0] * n_streams
    cache = Cache()
    outs = {}
    for s, step in enumerate(schedule):
        if step[0] == "zero":
            cache.set_to_zero(torch.tensor(step[1]))
            continue
        idx, n = step[1], step[2] * geom.hop
        x = np.stack([audio[i, :, pos[i]:pos[i] + n] for i in idx])
        for i in idx:
            pos[i] += n
        with torch.no_grad():
            y = model.encode(torch.from_numpy(x), cache=cache, sample_indices=torch.tensor(idx), use_cache=True).mean
        outs[f"out:{s}"] = y.numpy().astype(np.float32)
    meta = dict(kind="encode_stream", geom=geom.__dict__, init=init, wseed=wseed, aseed=aseed, n_streams=n_streams,
                schedule=schedule, total=total)
    np.savez(os.path.join(GOLD, name + ".npz"), meta=np.array(json.dumps(meta)), **outs)
    print(f"{name}: {len(outs)} encode calls recorded")


def main_encoder():
    """Encoder fixtures: ragged lengths exercise the right extra padding of every strided conv
    (:127-133); an exact multiple of the hop exercises the no-padding path; default init pins the real init scale."""
    os.makedirs(GOLD, exist_ok=True)
    encoder_case("enc_full_stress_L6500", FULL, "stress", 4321, 1, 3200 * 2 + 100, 5)
    encoder_case("enc_full_mild_B2_L12800", FULL, "mild", 99, 2, 3200 * 4, 6)
    encoder_case("enc_full_default_L9999", FULL, "default", 0, 1, 9999, 7)
    encoder_case("enc_full_stress_L1", FULL, "stress", 4321, 2, 1, 8)            # shorter than every kernel
    encoder_case("enc_tiny_stress_L10117", TINY, "stress", 7, 2, 3200 * 3 + 517, 5)
    # semantic tokenizer shape (N1): same encoder class with vae_dim 128
    encoder_case("enc_sem128_stress_L7001", DecoderGeometry(vae_dim=128, depths=(2, 1, 1, 1, 1, 1, 2)), "stress", 11, 1, 7001, 9)
    # streaming encode: steady state, a late joiner (all-or-nothing get, :204-207), set_to_zero, multi-hop chunks
    sched = [["encode", [0, 1], 1], ["encode", [0, 1], 1], ["encode", [0], 2], ["zero", [1]], ["encode", [0, 1], 1],
             ["encode", [0, 1, 2], 1], ["encode", [2, 0], 3]]
    encoder_stream_case("enc_stream_stress", DecoderGeometry(depths=(2, 1, 1, 1, 1, 1, 2)), "stress", 13, 3, sched, 10)
    encoder_stream_case("enc_stream_tiny", TINY, "stress", 7, 3, sched, 10)


def main():
    if "--encoder-only" in sys.argv:
        return main_encoder()
    os.makedirs(GOLD, exist_ok=True)
    # --- full geometry -------------------------------------------------------------------
    model, _ = offline_case("full_stress_T6", FULL, "stress", 1234, 1, 6, 11)
    # streaming on the same weights: steady state, a late joiner (all-or-nothing get, :204-207),
    # set_to_zero at a segment end (:234-241), a multi-frame chunk
    sched = [["decode", [0, 1], 1], ["decode", [0, 1], 1], ["decode", [0], 1], ["decode", [0, 1], 2],
             ["zero", [1]], ["decode", [0, 1], 1], ["decode", [0, 1, 2], 1], ["decode", [2, 0], 1]]
    streaming_case("full_stress_stream", FULL, "stress", 1234, sched, 21, 3, model=model)
    del model
    offline_case("full_default_T6", FULL, "default", 0, 1, 6, 12)
    # K1 = BASELINE.json configs[0]: batch 1, 10 s (75 frames) — under stress init so every layer counts
    offline_case("full_stress_K1_T75", FULL, "stress", 1234, 1, 75, 1, with_taps=False)
    offline_case("full_mild_T8", FULL, "mild", 77, 2, 8, 13, with_taps=False)
    # --- reduced geometry (fast CPU tests; generic-C code paths) -----------------------------
    model, _ = offline_case("tiny_stress_T12", TINY, "stress", 5, 2, 12, 14)
    sched = [["decode", [0, 1, 2], 1]] * 3 + [["decode", [1], 3], ["zero", [0, 2]], ["decode", [0, 1, 2], 2],
             ["decode", [2, 1], 1], ["decode", [0, 1, 2, 3], 1], ["decode", [3], 7]]
    streaming_case("tiny_stress_stream", TINY, "stress", 5, sched, 22, 4, model=model)
    # T == vae_dim layout ambiguity (:1102): [B, T, D] with T == D is read as [B, D, T]
    offline_case("tiny_stress_T16_ambiguous", TINY, "stress", 5, 1, 16, 15, with_taps=False)
    main_encoder()


if __name__ == "__main__":
    main()
