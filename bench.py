#!/usr/bin/env python
"""bench.py — decoded audio-seconds per second of the acoustic-tokenizer decode path on N B200s.

    python bench.py --gpus N --steps K --warmup W            (N > 1: launched under torch.distributed.run)
    python bench.py --impl reference ...                      CPU arm: the numpy port of the reference path

A step = one pass of the hot path over one batch of synthetic latents: BASELINE.json configs[1],
batch 64 x 60 s (450 latent frames) per GPU, bf16 latents in / bf16 waveform out.  Utterances are sharded
across ranks with no collective on the data path (weak scaling: every rank decodes its own 64).
One JSON line on rank 0; see DESIGN.md "Measurement" for every key.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "vibevoice-rocm_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

FRAME_RATE = 7.5
METRIC = "decoded audio-sec/sec"
UNIT = "audio-s/s"
K2_B, K2_T = 64, 450
CPU_SAMPLE_B, CPU_SAMPLE_T = 1, 75          # configs[0]: the reference's own CPU-runnable case (10 s)


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(path):
        with open(path) as f:
            p = json.load(f)
        return dict(hbm_gbs=float(p["hbm_gbs"]), tf_burst=float(p["bf16_tflops"]),
                    tf_sustained=float(p.get("bf16_tflops_sustained", p["bf16_tflops"])), source="measured")
    return dict(hbm_gbs=6650.0, tf_burst=1590.0, tf_sustained=1400.0, source="fallback")


# --------------------------------------------------------------------------------------------------
# CPU arm: the reference's own CPU path = its ATen call sequence (oracle/torch_ref.py: F.conv1d / conv_transpose1d /
# linear / gelu, fp32, pinned to the reference's outputs by tests/test_oracle_torch_ref.py), on all host threads.
# /root/reference cannot travel to the GPU box, so the modules are restated call for call: kind = "port".
# --------------------------------------------------------------------------------------------------
def workload_config(B, T):
    return {"workload": f"acoustic decoder bf16, batch {B} x {T / FRAME_RATE:.0f} s ({T} latent frames) synthetic "
                        f"latents per GPU, random-init (stress) weights",
            "l2": "inputs >> L2 (each pass streams >700 MB of activations per utterance group)",
            "sharding": "utterances per rank, no data-path collective"}


def cpu_reference_audio_s_per_s(sd, steps: int, warmup: int):
    """-> (audio-s/s, seconds per decode, threads used): configs[0] (batch 1 x 75 frames = 10 s) per step."""
    import torch
    from oracle import torch_ref
    from oracle.geometry import FULL
    threads = torch_ref.use_all_host_threads()
    w = torch_ref.to_weights(sd)
    gen = torch.Generator().manual_seed(1)
    lat = torch.randn(CPU_SAMPLE_B, CPU_SAMPLE_T, FULL.vae_dim, generator=gen).to(torch.bfloat16).float()
    for _ in range(warmup):
        torch_ref.decode(w, FULL, lat)
    ts = []
    for _ in range(steps):
        t0 = time.perf_counter()
        torch_ref.decode(w, FULL, lat)
        ts.append(time.perf_counter() - t0)
    sec = sum(ts) / len(ts)
    return CPU_SAMPLE_B * CPU_SAMPLE_T / FRAME_RATE / sec, sec, threads


CPU_SAMPLE = (f"batch {CPU_SAMPLE_B} x {CPU_SAMPLE_T} frames (10 s = BASELINE.json configs[0]) per step; the reference's "
              f"ATen call sequence (torch CPU fp32: conv1d / conv_transpose1d / linear / erf-GELU), all host threads")


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    from vvdec.synthetic import random_state_dict
    sd = random_state_dict()
    v, sec, threads = cpu_reference_audio_s_per_s(sd, max(1, args.steps), max(0, args.warmup))
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.batch, args.frames),
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "sample": CPU_SAMPLE,
                         "torch": torch.__version__},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------------
# parity samples at the benchmark's own sizes: the CPU reference (oracle/torch_ref.py, fp32) on ONE utterance / row /
# window of the tensors the timed legs just produced.  Tolerance = BASELINE.json: SNR >= 40 dB, max-abs <= 2e-2 of peak.
# --------------------------------------------------------------------------------------------------
def _metrics(test, ref):
    import torch
    t, r = test.double().flatten().cpu(), ref.double().flatten().cpu()
    d = t - r
    peak = float(r.abs().max()) or 1.0
    den = float((d * d).sum())
    snr = float("inf") if den == 0 else float(10 * torch.log10((r * r).sum().clamp_min(1e-300) / den))
    return {"snr_db": snr, "rel_max": float(d.abs().max()) / peak}


def _verdict(m):
    m["ok"] = bool(m["snr_db"] >= 40.0 and m["rel_max"] <= 2e-2)
    return m


def parity_decode_sample(sd, geom, lat_cpu_f32, y_dev, what):
    """lat_cpu_f32 [1, T, D] (the bf16-representable latents of one utterance), y_dev the path's output for it."""
    from oracle import torch_ref
    torch_ref.use_all_host_threads()
    t0 = time.perf_counter()
    ref = torch_ref.decode(torch_ref.to_weights(sd), geom, lat_cpu_f32)
    m = _verdict(_metrics(y_dev.float().cpu(), ref))
    m.update({"what": what, "cpu_seconds": time.perf_counter() - t0})
    return m


# --------------------------------------------------------------------------------------------------
# clocks
# --------------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    REASONS = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
               0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting"}

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz, self._halt = index, [], set(), None, threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        while not self._halt.is_set():
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                r = self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                for bit, name in self.REASONS.items():
                    if r & bit and name != "gpu_idle":
                        self.reasons.add(name)
            except Exception:
                pass
            self._halt.wait(0.05)

    def finish(self):
        self._halt.set()
        self.join(timeout=2)
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


# --------------------------------------------------------------------------------------------------
# roofline bookkeeping: algorithmic work per latent frame of each kernel (DESIGN.md "Kernels")
# --------------------------------------------------------------------------------------------------
def kernel_work(geom, kind: str, stage: int):
    """-> (bound, algorithmic bytes per frame, algorithmic flop per frame, bytes this design moves per frame)."""
    C = geom.stage_channels(stage)
    M = 1
    for r in geom.ratios[:stage]:
        M *= r
    cm = C * M
    if kind == "mixer":        # 4*C*M bytes (bf16 in + out), ~20*C*M flop; we move fp32 x, fp32 y, bf16 a
        return "hbm", 4 * cm, 20 * cm, 10 * cm
    if kind in ("ffn1", "ffn2"):   # FFN half = 4*C*M bytes and 16*C^2*M flop, split evenly over its two launches
        bound = "tensor" if C >= 128 else "hbm"
        moved = (2 * cm + 8 * cm) if kind == "ffn1" else (8 * cm + 4 * cm + 4 * cm)
        return bound, 2 * cm, 8 * C * cm, moved
    if kind == "ffn_fused":    # whole FFN half in one launch: a (bf16) + resid (fp32) in, out (fp32)
        return ("tensor" if C >= 128 else "hbm"), 4 * cm, 16 * C * cm, 10 * cm
    if kind == "ffn_pair":     # both GEMMs in L2-resident row chunks: the hidden tensor never reaches HBM
        return "tensor", 4 * cm, 16 * C * cm, 10 * cm
    if kind == "block_fused":  # whole Block1D in one launch (C <= 128): SURVEY's mixer half + FFN half = 8*C*M bytes;
        # this design moves exactly that (fp32 x in, fp32 x out).  (20 + 16 C) / 8 flop per byte: 132 at C = 64 (below the
        # ~212 flop/B ridge of 1382 TFLOP/s over 6.5 TB/s: HBM-bound), 258 at C = 128 (above it: tensor-bound)
        return ("tensor" if C >= 128 else "hbm"), 8 * cm, 20 * cm + 16 * C * cm, 8 * cm
    if kind == "upsample":     # 2*(Cin*Min + Cout*Mout) bytes, 4*Cin*Cout*r*Min flop
        cin, r = geom.stage_channels(stage - 1), geom.ratios[stage - 1]
        min_ = M // r
        bound = "tensor" if cin >= 256 else "hbm"
        return bound, 2 * (cin * min_ + cm), 4 * cin * C * r * min_, 2 * cin * min_ + 4 * cm
    if kind == "stem":
        return "hbm", 2 * geom.vae_dim + 2 * cm, 2 * 7 * geom.vae_dim * C, 2 * geom.vae_dim + 4 * cm
    if kind == "head":
        return "hbm", 2 * cm + 4 * M, 2 * 7 * cm, 4 * cm + 2 * M
    return "hbm", 0, 0, 0


def streaming_leg(dec, geom, dev, sd, n_streams=256, steps=300, warmup=30):
    """BASELINE.json configs[3]: frame-by-frame decode of `n_streams` concurrent streams with the device-resident
    conv state; per-step latency from CUDA events, eager launches and one CUDA graph per step."""
    import torch
    cache = dec.new_cache(n_streams)
    idx = torch.arange(n_streams, device=dev, dtype=torch.int32)
    gen = torch.Generator().manual_seed(4)
    lat = torch.randn(n_streams, 1, geom.vae_dim, generator=gen).to(torch.bfloat16).to(dev)

    def step():
        return dec.decode(lat, cache=cache, sample_indices=idx, use_cache=True)

    def timed(fn, n):
        ts = []
        for _ in range(n):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            e1.synchronize()
            ts.append(e0.elapsed_time(e1))
        ts.sort()
        return ts[len(ts) // 2], ts[min(len(ts) - 1, int(len(ts) * 0.99))]

    for _ in range(warmup):
        step()
    torch.cuda.synchronize()
    dec.take_launch_count()
    p50, p99 = timed(step, steps)
    launches = dec.take_launch_count() // steps
    out = {"streams": n_streams, "frames_per_step": 1, "eager_p50_ms": p50, "eager_p99_ms": p99,
           "launches_per_step": int(launches), "realtime_budget_ms": 1e3 / FRAME_RATE}
    try:
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(3):
                step()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            y = step()
        for _ in range(warmup):
            graph.replay()
        torch.cuda.synchronize()
        g50, g99 = timed(graph.replay, steps)
        out.update({"graph_p50_ms": g50, "graph_p99_ms": g99, "audio_s_per_s_graph": n_streams / FRAME_RATE / (g50 / 1e3)})
        del y
    except Exception as e:                                     # report, do not hide
        out["graph_error"] = f"{type(e).__name__}: {e}"[:200]
    # the semantic tokenizer's streaming encode of the 3200 samples just decoded (same TokenizerEncoder,
    # vae_dim 128; modeling_vibevoice_inference.py:656-662), audio handed over on the device
    try:
        import vvdec
        from vvdec.synthetic import random_encoder_state_dict
        sgeom = vvdec.Geometry(vae_dim=128)
        senc = vvdec.B200AcousticEncoder(sgeom, device=dev.index, fix_std=0.0, std_dist_type="none")
        senc.load_state_dict(random_encoder_state_dict(sgeom, seed=77))
        ecache = senc.new_cache(n_streams)
        chunk = step()                                         # [n_streams, 1, 3200] bf16, stays on the device

        def estep():
            return senc.encode(chunk, cache=ecache, sample_indices=idx, use_cache=True).mean
        for _ in range(warmup):
            estep()
        torch.cuda.synchronize()
        senc.take_launch_count()
        e50, e99 = timed(estep, steps)
        out["semantic_encode"] = {"eager_p50_ms": e50, "eager_p99_ms": e99,
                                  "launches_per_step": int(senc.take_launch_count() // steps)}

        def both():
            return senc.encode(step(), cache=ecache, sample_indices=idx, use_cache=True).mean
        b50, b99 = timed(both, steps)
        out["decode_plus_semantic_encode"] = {"eager_p50_ms": b50, "eager_p99_ms": b99}
        del senc, ecache
    except Exception as e:
        out["semantic_encode"] = {"error": f"{type(e).__name__}: {e}"[:200]}
    # the PCM16 step + streamer hand-off of the chunk just decoded (main.py:937-951, streamer.py:50-55)
    try:
        import time
        import numpy as np
        import vvdec
        chunk = step()
        for _ in range(5):
            vvdec.pcm16(chunk)
        torch.cuda.synchronize()
        k50, k99 = timed(lambda: vvdec.pcm16(chunk), steps)
        streamer = vvdec.B200PcmStreamer(batch_size=n_streams)
        host_idx = list(range(n_streams))

        def wall(fn, n):
            ts = []
            for _ in range(n):
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                fn()
                ts.append((time.perf_counter() - t0) * 1e3)
            ts.sort()
            return ts[len(ts) // 2]

        def put_ours():
            streamer.put(chunk, host_idx)
            for q in streamer.audio_queues:                    # drain, as the consumer thread would
                q.get_nowait()

        def put_reference_pattern():                           # streamer.py:50-55 + main.py:937-951, per sample
            for i in range(n_streams):
                a = chunk[i].detach().cpu().float().numpy().astype(np.float32).squeeze()
                if np.max(np.abs(a)) > 1.0:
                    a = a / np.max(np.abs(a))
                (a * 32767).astype(np.int16)
        wall(put_ours, 5)
        out["pcm16"] = {"kernel_p50_ms": k50, "kernel_p99_ms": k99, "launches_per_step": 1,
                        "bytes_per_step": int(chunk.numel() * (chunk.element_size() + 2)),
                        "streamer_put_wall_p50_ms": wall(put_ours, 50),
                        "reference_pattern_wall_p50_ms": wall(put_reference_pattern, 10),
                        "d2h_bytes_per_step": int(chunk.numel() * 2)}
    except Exception as e:
        out["pcm16"] = {"error": f"{type(e).__name__}: {e}"[:200]}
    # SURVEY 8d K4 schedule through the product's frame step: random active subsets, set_to_zero events
    try:
        out["schedule"] = streaming_schedule_leg(dec, geom, dev, sd, n_streams=n_streams)
    except Exception as e:
        out["schedule"] = {"error": f"{type(e).__name__}: {e}"[:300]}
    return out


def sampler_leg(dev, n_streams=256, steps=20, iters=20, parity=True):
    """the diffusion-head sampler that produces the latents the decoder consumes -- sample_speech_tokens
    (modeling_vibevoice_inference.py:695-708) for `n_streams` streams, `steps` DPM-Solver++ steps of the 1.5B model's head
    (hidden 1536, 4 layers) on 2 x n_streams rows.  Per-call device time, eager and as one CUDA graph; two streams are
    checked against the numpy oracle (fp32)."""
    import numpy as np
    import torch
    import vvdec
    from oracle.sampler_oracle import OracleSampler, SamplerGeometry, make_sampler_weights
    og = SamplerGeometry()
    w = make_sampler_weights(og, 2026)
    smp = vvdec.B200DiffusionSampler(vvdec.SamplerGeometry(inference_steps=steps), device=dev.index)
    smp.load_state_dict({k: torch.from_numpy(v) for k, v in w.items()})
    gen = torch.Generator().manual_seed(6)
    cond = torch.randn(n_streams, og.hidden, generator=gen).to(torch.bfloat16)
    neg = torch.randn(n_streams, og.hidden, generator=gen).to(torch.bfloat16)
    noise = torch.randn(n_streams, og.latent, generator=gen)
    cd, nd, zd = cond.to(dev), neg.to(dev), noise.to(dev)

    def call():
        return smp.sample_speech_tokens(cd, nd, cfg_scale=3.0, noise=zd)

    def timed(fn, k):
        ts = []
        for _ in range(k):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record(); e1.synchronize()
            ts.append(e0.elapsed_time(e1))
        ts.sort()
        return ts[len(ts) // 2]
    for _ in range(3):
        y = call()
    torch.cuda.synchronize()
    smp.take_launch_count()
    eager = timed(call, iters)
    launches = smp.take_launch_count() // iters
    flop = 2.0 * steps * 2 * n_streams * ((3 * og.layers + 2) * og.hidden * og.hidden + og.layers * 3 * og.hidden * og.ffn
                                          + 2 * og.hidden * og.latent)
    out = {"workload": f"sample_speech_tokens, {n_streams} streams x {steps} steps, head hidden {og.hidden} x {og.layers} layers "
                       f"({2 * n_streams} rows per head pass), bf16 conditions",
           "eager_p50_ms": eager, "launches_per_call": int(launches), "tflops_eager": flop / (eager * 1e-3) / 1e12}
    try:
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            call()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            yg = call()
        for _ in range(3):
            graph.replay()
        torch.cuda.synchronize()
        gms = timed(graph.replay, iters)
        out.update({"graph_p50_ms": gms, "tflops_graph": flop / (gms * 1e-3) / 1e12,
                    "graph_equals_eager": bool(torch.equal(yg, y))})
    except Exception as e:
        out["graph_error"] = f"{type(e).__name__}: {e}"[:200]
    if parity:
        k = 2
        t0 = time.perf_counter()
        ref = OracleSampler(w, og).sample(cond[:k].float().numpy(), neg[:k].float().numpy(), noise[:k].numpy(), cfg_scale=3.0, steps=steps)
        sub = smp.sample_speech_tokens(cd[:k], nd[:k], cfg_scale=3.0, noise=zd[:k])
        m = _verdict(_metrics(sub.float().cpu(), torch.from_numpy(ref)))
        m.update({"what": f"{k} streams vs the numpy oracle (fp32)", "cpu_seconds": time.perf_counter() - t0})
        out["parity"] = m
    del smp
    return out


def streaming_schedule_leg(dec, geom, dev, sd, n_streams=256, steps=320, p_active=0.8, zero_every=300, track=(3, 200)):
    """SURVEY 8d's K4 schedule: every step a random subset of the streams (p = 0.8) decodes one frame, every stream ends a
    speech segment (set_to_zero on both caches) about once per `zero_every` steps -- driven through vvdec.FrameStep
    (decode -> semantic streaming encode -> PCM16 as one CUDA graph per stream-count bucket, indices on the device, no
    host sync inside the loop).  Per-step device time from CUDA events read after the loop.  Two tracked streams are then
    checked against the CPU reference: their longest segment's chunks (streaming) vs the reference's OFFLINE decode of the
    same latents (the reference's streaming form equals its offline form), and the semantic means vs the
    reference's encode of the audio that was handed over."""
    import numpy as np
    import torch
    import vvdec
    from vvdec.synthetic import random_encoder_state_dict
    sgeom = vvdec.Geometry(vae_dim=128)
    senc = vvdec.B200AcousticEncoder(sgeom, device=dev.index, fix_std=None, std_dist_type="none")
    sd_sem = random_encoder_state_dict(sgeom, seed=77)
    senc.load_state_dict(sd_sem)
    fs = vvdec.FrameStep(dec, senc, pcm=True, n_slots=n_streams, bucket=32)
    rng = np.random.Generator(np.random.Philox(key=[4, 4]))
    gen = torch.Generator().manual_seed(4)
    lat_all = torch.randn(steps, n_streams, 1, geom.vae_dim, generator=gen).to(torch.bfloat16)
    lat_dev = lat_all.to(dev)
    act = [np.nonzero(rng.random(n_streams) < p_active)[0] for _ in range(steps)]
    zero = [np.nonzero(rng.random(n_streams) < 1.0 / zero_every)[0] for _ in range(steps)]
    idx_dev = [torch.from_numpy(a.astype(np.int64)).to(dev) for a in act]
    zero_dev = [torch.from_numpy(z.astype(np.int64)).to(dev) if z.size else None for z in zero]
    for nb in sorted({-(-len(a) // 32) * 32 for a in act}):         # build every bucket's graph before timing
        fs(lat_dev[0, :nb], torch.arange(nb, device=dev))
    fs.set_to_zero(torch.arange(n_streams, device=dev))
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    seg = {k: {"lat": [], "audio": [], "sem": []} for k in track}
    best = {k: None for k in track}

    def close(k):
        if best[k] is None or len(seg[k]["lat"]) > len(best[k]["lat"]):
            best[k] = seg[k]
        seg[k] = {"lat": [], "audio": [], "sem": []}
    torch.cuda.set_sync_debug_mode("error")                         # a host synchronisation inside the loop raises
    try:
        t_wall = time.perf_counter()
        for t in range(steps):
            ev[t][0].record()
            o = fs(lat_dev[t].index_select(0, idx_dev[t]), idx_dev[t])
            ev[t][1].record()
            for k in track:
                pos = np.nonzero(act[t] == k)[0]
                if pos.size:
                    seg[k]["lat"].append(lat_all[t, k]); seg[k]["audio"].append(o.audio[int(pos[0])]); seg[k]["sem"].append(o.semantic[int(pos[0])])
            if zero_dev[t] is not None:
                fs.set_to_zero(zero_dev[t])
                for k in track:
                    if k in zero[t]:
                        close(k)
    finally:
        torch.cuda.set_sync_debug_mode("default")
    torch.cuda.synchronize()
    wall_ms = (time.perf_counter() - t_wall) * 1e3 / steps
    fs.check()
    for k in track:
        close(k)
    ts = sorted(a.elapsed_time(b) for a, b in ev)
    res = {"what": f"{steps} steps, {n_streams} streams, each active with p = {p_active} per step, a set_to_zero per stream "
                   f"about every {zero_every} steps; vvdec.FrameStep = decode + semantic encode + PCM16 in one CUDA graph",
           "active_mean": float(np.mean([len(a) for a in act])), "zero_events": int(sum(z.size for z in zero)),
           "step_p50_ms": ts[len(ts) // 2], "step_p99_ms": ts[min(len(ts) - 1, int(len(ts) * 0.99))],
           "host_wall_ms_per_step": wall_ms, "graphs": len(fs._plans), "graph_launches": fs.graph_launches}
    # parity of the tracked streams (CPU reference, fp32)
    from oracle import torch_ref
    torch_ref.use_all_host_threads()
    wd, ws = torch_ref.to_weights(sd), torch_ref.to_weights(sd_sem)
    checks = []
    for k in track:
        b = best[k]
        n = min(len(b["lat"]), 40)
        if n < 4:
            continue
        lat = torch.stack(b["lat"][:n], 0).reshape(1, n, geom.vae_dim).float()
        audio = torch.cat([a.reshape(-1) for a in b["audio"][:n]]).float().cpu()
        sem = torch.cat([x.reshape(1, -1) for x in b["sem"][:n]], 0).float().cpu()
        t0 = time.perf_counter()
        ref_a = torch_ref.decode(wd, geom, lat).reshape(-1)
        ref_s = torch_ref.encode(ws, sgeom, audio.reshape(1, 1, -1))[0]
        m = {"stream": int(k), "frames": n, "decode": _verdict(_metrics(audio, ref_a)), "semantic": _verdict(_metrics(sem, ref_s)),
             "cpu_seconds": time.perf_counter() - t0}
        checks.append(m)
    res["parity"] = checks
    res["parity_ok"] = bool(checks) and all(c["decode"]["ok"] and c["semantic"]["ok"] for c in checks)
    del senc, fs
    return res


def torch_eager_leg(sd, geom, dev, dec, B=8, T=450, steps=3):
    """Context number, not an arm: the reference's call sequence (oracle/torch_ref.py: the ATen calls the reference module
    makes -- cuDNN depthwise / transposed conv, cuBLAS linear, F.gelu) in bf16 eager on the same GPU and weights, on a
    bounded sample.  Lives in bench.py only; the product path never uses it."""
    import torch
    from oracle import torch_ref
    w = torch_ref.to_weights(sd, dev, torch.bfloat16)
    gen = torch.Generator().manual_seed(9)
    lat = torch.randn(B, T, geom.vae_dim, generator=gen).to(torch.bfloat16).to(dev)
    with torch.no_grad():
        y = torch_ref.decode(w, geom, lat[:1, :8])    # agreement check on a small piece (both bf16 paths vs each other)
        ours = dec.decode(lat[:1, :8])
        err = (y.float() - ours.float())
        snr = float(10 * torch.log10(ours.float().pow(2).sum() / err.pow(2).sum().clamp_min(1e-30)))
        del y, ours, err
        torch_ref.decode(w, geom, lat)                # warm-up (cuDNN heuristics, allocator)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            torch_ref.decode(w, geom, lat)
        e1.record()
        torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    del w
    torch.cuda.empty_cache()
    return {"what": "stock PyTorch ops (cuDNN/cuBLAS), bf16 eager, same GPU and weights; context only",
            "sample": f"batch {B} x {T} frames", "ms_per_step": ms, "value": B * T / FRAME_RATE / (ms / 1e3), "unit": UNIT,
            "snr_vs_b200_path_db": snr, "torch": torch.__version__}


def longform_leg(dec, geom, dev, world, rank, dist, steps, sd=None, frames=40_500, parity=True):
    """BASELINE.json configs[2]: one 90-minute utterance (40,500 latent frames) through vvdec.decode_sharded -- time chunks
    across the ranks, each rank recomputing its causal halo (58 frames), no collective on the data path, one batched
    point-to-point gather of the finished waveform slices into rank 0's tensor (strong scaling: the utterance is fixed,
    N varies).  All ranks call this.  Rank 0 then checks sampled windows of the gathered waveform against the CPU
    reference (one window straddles a shard boundary when there is one) and, at N > 1, times the same utterance on its
    GPU alone so that the line carries the strong-scaling speed-up measured in one run."""
    import torch
    import vvdec
    from vvdec.sharded import plan_for
    halo, hop, D = dec.halo_frames, geom.hop, geom.vae_dim
    gen = torch.Generator().manual_seed(3)
    lat_all = torch.randn(1, frames, D, generator=gen)
    seg = frames // 4
    for k, mu in enumerate((-0.5, 0.0, 0.5, 1.0)):                    # four "speakers": four segments, shifted means
        lat_all[:, k * seg:(k + 1) * seg if k < 3 else frames] += mu
    lat_all = lat_all.to(torch.bfloat16)
    lat_dev = lat_all.to(dev)
    shards = plan_for([frames], world, halo)

    def sync():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    final = vvdec.decode_sharded(dec, lat_dev, shards=shards)
    sync()
    t_dec = t_all = 0.0
    for _ in range(steps):
        sync()
        tm = {}
        e0 = torch.cuda.Event(enable_timing=True)
        e0.record()
        final = vvdec.decode_sharded(dec, lat_dev, shards=shards, timings=tm)
        sync()
        t_dec += e0.elapsed_time(tm["decode_done"]); t_all += e0.elapsed_time(tm["gather_done"])
    t = torch.tensor([t_dec / steps, t_all / steps], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_dec, ms_all = float(t[0]), float(t[1])
    out = {"workload": f"one {frames / FRAME_RATE / 60:.0f}-minute utterance ({frames} frames) through vvdec.decode_sharded: "
                       f"time-chunk shards with a {halo}-frame recomputed halo, waveform slices sent point-to-point into "
                       f"rank 0's tensor",
           "scaling": "strong", "n_gpus": world, "shards": len(shards), "halo_frames": halo,
           "frames_owned_rank0": sum(s.frames for s in shards if s.rank == 0),
           "recomputed_halo_frames_total": sum(s.halo for s in shards),
           "ms_decode": ms_dec, "ms_decode_plus_gather": ms_all, "ms_gather": ms_all - ms_dec,
           "gather_bytes": (sum(s.frames for s in shards if s.rank != 0) * hop * 2 if world > 1 else 0),
           "value": frames / FRAME_RATE / (ms_all / 1e3), "unit": UNIT}
    if rank != 0:
        return out
    out["finite"] = bool(torch.isfinite(final.float()).all())
    if world > 1:                                                     # the same utterance on rank 0's GPU alone
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        solo = dec.decode(lat_dev)
        torch.cuda.synchronize()
        e0.record()
        solo = dec.decode(lat_dev)
        e1.record()
        torch.cuda.synchronize()
        ms1 = e0.elapsed_time(e1)
        out["single_gpu_ms_same_run"] = ms1
        out["speedup_vs_single_gpu"] = ms1 / ms_all
        out["gathered_vs_single_gpu"] = _metrics(final, solo)          # halo recompute vs one causal pass: rounding only
        del solo
    if parity and sd is not None:
        bounds = sorted(s.start for s in shards if s.start > 0)
        wins = [(0, 8), ((bounds[len(bounds) // 2] - 4) if bounds else frames // 2, 8), (frames - 8, 8)]
        from oracle import torch_ref
        torch_ref.use_all_host_threads()
        wts = torch_ref.to_weights(sd)
        res = []
        for w0, n in wins:
            lo = max(0, w0 - halo)
            t0 = time.perf_counter()
            ref = torch_ref.decode(wts, geom, lat_all[:, lo:w0 + n].float())[:, :, (w0 - lo) * hop:]
            m = _verdict(_metrics(final[:, :, w0 * hop:(w0 + n) * hop].float().cpu(), ref))
            m.update({"frames": [w0, w0 + n], "straddles_shard_boundary": any(w0 < b < w0 + n for b in bounds),
                      "cpu_seconds": time.perf_counter() - t0})
            res.append(m)
        out["parity_windows"] = res
        out["parity_ok"] = all(m["ok"] for m in res)
    return out


def enc_kernel_work(geom, kind: str, stage: int):
    """Encoder counterpart of kernel_work: per SAMPLE-ROW of the stage (not per frame) -> (bound, algorithmic bytes,
    flop, bytes moved) per row.  Stage i has C = n_filters * 2^i channels."""
    n = len(geom.depths)
    C = geom.n_filters * 2 ** stage
    if kind == "mixer":
        return "hbm", 4 * C, 20 * C, 10 * C
    if kind in ("ffn1", "ffn2"):
        moved = (2 * C + 8 * C) if kind == "ffn1" else (8 * C + 4 * C + 4 * C)
        return ("tensor" if C >= 128 else "hbm"), 2 * C, 8 * C * C, moved
    if kind in ("ffn_fused", "ffn_pair"):
        return ("tensor" if C >= 128 else "hbm"), 4 * C, 16 * C * C, 10 * C
    if kind == "block_fused":
        return "hbm", 8 * C, 20 * C + 16 * C * C, 8 * C
    if kind == "downsample":   # per OUTPUT row: reads r rows of C/2 bf16, writes C fp32 (algorithmic: bf16 out)
        r = geom.ratios[n - 1 - stage]
        cin = C // 2
        return ("tensor" if cin >= 256 else "hbm"), 2 * r * cin + 2 * C, 2 * 2 * r * cin * C, 2 * r * cin + 4 * C
    if kind == "enc_stem":
        return "hbm", 2 + 2 * C, 2 * 7 * C, 4 + 4 * C
    if kind == "enc_head":
        return "hbm", 2 * C + 2 * geom.vae_dim, 2 * 7 * C * geom.vae_dim, 2 * C + 4 * geom.vae_dim
    return "hbm", 0, 0, 0


def encoder_leg(geom, dev, pk, steps, warmup, B=32, L=720_000, kernels_out="", cpu_baseline=True):
    """BASELINE.json configs[4]: acoustic tokenizer ENCODER on voice prompts, batch 32 x 30 s, through
    B200AcousticEncoder.encode (the C ABI's vv_encode).  Device-resident throughput, an end-to-end figure with pinned
    host buffers, and the per-kernel table of one profiled pass."""
    import torch
    import vvdec
    from vvdec.synthetic import random_encoder_state_dict
    enc = vvdec.B200AcousticEncoder(geom, device=dev.index)
    sd_cpu = random_encoder_state_dict(geom)
    enc.load_state_dict(sd_cpu)
    gen = torch.Generator().manual_seed(5)
    a_host = (torch.randn(B, 1, L, generator=gen) * 0.1).to(torch.bfloat16).pin_memory()
    audio = a_host.to(dev)
    T = enc.frames(L)
    m_host = torch.empty((B, T, geom.vae_dim), dtype=torch.bfloat16).pin_memory()
    for _ in range(max(2, warmup)):
        enc.encode(audio)
    torch.cuda.synchronize()
    enc.take_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        enc.encode(audio)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    launches = enc.take_launch_count() // steps

    def e2e_step():
        m_host.copy_(enc.encode(a_host.to(dev, non_blocking=True)).mean, non_blocking=True)
    e2e_step()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(steps):
        e2e_step()
    e1.record()
    torch.cuda.synchronize()
    ms_e2e = e0.elapsed_time(e1) / steps
    audio_s = B * L / 24000.0
    enc.set_profiling(True)
    enc.encode(audio)
    prof = enc.read_profile()
    enc.set_profiling(False)
    total = sum(r["ms"] for r in prof) or 1.0
    n = len(geom.depths)
    rows_of = [L]
    for i in range(1, n):
        r = geom.ratios[n - 1 - i]
        rows_of.append(-(-rows_of[-1] // r))
    table = []
    for r in prof:
        bound, abytes, aflop, moved = enc_kernel_work(geom, r["kind"], r["stage"])
        nl = max(1, r["launches"])
        nblk = geom.depths[n - 1 - r["stage"]] if r["kind"] in ("mixer", "ffn1", "ffn2", "ffn_fused", "block_fused", "ffn_pair") else 1
        passes = max(1, nl // nblk)
        units = B * rows_of[r["stage"]] / passes
        msl = r["ms"] / nl
        table.append(dict(kind=r["kind"], stage=r["stage"], launches=nl, ms_total=r["ms"], share=r["ms"] / total,
                          ms_per_launch=msl, rows_per_launch=units, bound=bound,
                          gbs_algorithmic=abytes * units / (msl * 1e-3) / 1e9 if msl > 0 else 0.0,
                          gbs_moved=moved * units / (msl * 1e-3) / 1e9 if msl > 0 else 0.0,
                          tflops=aflop * units / (msl * 1e-3) / 1e12 if msl > 0 else 0.0))
    table.sort(key=lambda r: -r["ms_total"])
    top = table[0]
    if kernels_out:
        with open(kernels_out, "w") as f:
            json.dump({"config": {"B": B, "L": L, "ms_per_step": ms}, "kernels": table}, f, indent=1)
    row = 11 % B                                               # one K5 row kept for the parity sample below
    mean_row = enc.encode(audio).mean[row:row + 1].float().cpu()
    del enc
    cpu, parity = None, None
    if cpu_baseline:
        # the reference's ATen call sequence (oracle/torch_ref.py, fp32, all host threads) on ONE row of this batch: the
        # reported CPU figure for the encoder AND the K5-sized parity sample (tolerance as for the decoder)
        from oracle import torch_ref
        threads = torch_ref.use_all_host_threads()
        wts = torch_ref.to_weights(sd_cpu)
        a_row = a_host[row:row + 1].float()
        torch_ref.encode(wts, geom, a_row[:, :, :32_000])
        t0 = time.perf_counter()
        ref = torch_ref.encode(wts, geom, a_row)
        sec = time.perf_counter() - t0
        cpu = {"value": (L / 24000.0) / sec, "unit": "encoded audio-s/s", "cores": threads, "kind": "port",
               "sample": f"batch 1 x {L / 24000:.0f} s ({L} samples = one row of this batch), the reference's ATen call "
                         f"sequence in fp32, {sec:.2f} s"}
        parity = _verdict(_metrics(mean_row, ref))
        parity["what"] = f"row {row} of the batch ({L} samples -> {T} frames) vs the CPU reference"
    return {"cpu_baseline": cpu, "parity": parity,
            "workload": f"acoustic tokenizer encoder, batch {B} x {L / 24000:.0f} s ({L} samples -> {T} frames), bf16 audio in",
            "value": audio_s / (ms / 1e3), "unit": "encoded audio-s/s", "ms_per_step": ms, "gpu_launches": int(launches),
            "e2e": {"value": audio_s / (ms_e2e / 1e3), "ms_per_step": ms_e2e, "h2d_bytes_per_step": a_host.numel() * 2,
                    "d2h_bytes_per_step": m_host.numel() * 2},
            "top_kernel": {"kernel": f"{top['kind']}@stage{top['stage']}", "share": top["share"], "bound": top["bound"],
                           "ms_per_launch": top["ms_per_launch"],
                           "achieved": top["tflops"] if top["bound"] == "tensor" else top["gbs_algorithmic"],
                           "peak": pk["tf_sustained"] if top["bound"] == "tensor" else pk["hbm_gbs"],
                           "unit": "TFLOP/s" if top["bound"] == "tensor" else "GB/s"}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=K2_B)
    ap.add_argument("--frames", type=int, default=K2_T)
    ap.add_argument("--workspace-gb", type=float, default=24.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-streaming", action="store_true")
    ap.add_argument("--no-encoder", action="store_true")
    ap.add_argument("--no-longform", action="store_true")
    ap.add_argument("--no-torch-eager", action="store_true")
    ap.add_argument("--encoder-kernels-out", default="", help="per-kernel table of the encoder leg (JSON)")
    ap.add_argument("--kernels-out", default="", help="write the per-kernel timing table (JSON) here")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if args.impl == "reference":
        return run_reference_arm(args)

    import torch
    import torch.distributed as dist
    import vvdec
    from vvdec.synthetic import random_state_dict

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    B, T = args.batch, args.frames
    geom = vvdec.Geometry()
    sd = random_state_dict(geom)
    dec = vvdec.B200AcousticDecoder(geom, device=local, max_workspace_bytes=int(args.workspace_gb * (1 << 30)))
    dec.load_state_dict(sd)

    gen = torch.Generator().manual_seed(2 + rank)
    lat_host = torch.randn(B, T, geom.vae_dim, generator=gen).to(torch.bfloat16).pin_memory()
    lat = lat_host.to(dev)
    out_host = torch.empty((B, 1, T * geom.hop), dtype=torch.bfloat16).pin_memory()
    audio_s = B * T / FRAME_RATE

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms: float) -> float:
        if world == 1:
            return ms
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- device-resident throughput --------------------------------------------------------------
    for _ in range(args.warmup):
        y = dec.decode(lat)
    barrier()
    dec.take_launch_count()
    sampler = ClockSampler(local)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        y = dec.decode(lat)               # 738 MB of fresh activations per pass >> 126 MB L2: no flush needed
    e1.record()
    barrier()
    clocks = sampler.finish()
    launches = dec.take_launch_count()
    ms_step = max_over_ranks(e0.elapsed_time(e1) / args.steps)
    value = world * audio_s / (ms_step / 1e3)

    # ---- end to end through the public decode() with host buffers --------------------------------
    def e2e_step():
        x = lat_host.to(dev, non_blocking=True)
        out_host.copy_(dec.decode(x), non_blocking=True)
    for _ in range(2):
        e2e_step()
    barrier()
    e0.record()
    for _ in range(args.steps):
        e2e_step()
    e1.record()
    barrier()
    ms_e2e = max_over_ranks(e0.elapsed_time(e1) / args.steps)
    e2e = {"value": world * audio_s / (ms_e2e / 1e3), "unit": UNIT, "ms_per_step": ms_e2e,
           "h2d_bytes_per_step": lat_host.numel() * 2, "d2h_bytes_per_step": out_host.numel() * 2}

    # ---- long-form decode sharded in time (BASELINE.json configs[2]); every rank takes part -------------------
    longform = None
    if not args.no_longform:
        y = None
        try:
            longform = longform_leg(dec, geom, dev, world, rank, dist, max(2, min(5, args.steps // 2)), sd=sd,
                                    parity=not args.no_cpu_baseline)
        except Exception as e:                                  # report, do not hide
            longform = {"error": f"{type(e).__name__}: {e}"[:300]}
            if world > 1:
                raise

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- per-kernel timing, live, CUDA events on the decode stream (one extra step) --------------
    pk = peaks()
    dec.set_profiling(True)
    dec.decode(lat)
    prof = dec.read_profile()
    dec.set_profiling(False)
    total_ms = sum(r["ms"] for r in prof) or 1.0
    table = []
    frames = B * T
    for r in prof:
        bound, abytes, aflop, moved = kernel_work(geom, r["kind"], r["stage"])
        n = max(1, r["launches"])
        nblk = geom.depths[r["stage"]] if r["kind"] in ("mixer", "ffn1", "ffn2", "ffn_fused", "block_fused", "ffn_pair") else 1
        passes = n // nblk                                   # decode passes the workspace split the batch into
        units = frames / passes                              # latent frames one launch processes
        ms_launch = r["ms"] / n
        row = dict(kind=r["kind"], stage=r["stage"], launches=n, ms_total=r["ms"], share=r["ms"] / total_ms,
                   ms_per_launch=ms_launch, frames_per_launch=units, bound=bound,
                   gbs_algorithmic=abytes * units / (ms_launch * 1e-3) / 1e9 if ms_launch > 0 else 0.0,
                   gbs_moved=moved * units / (ms_launch * 1e-3) / 1e9 if ms_launch > 0 else 0.0,
                   tflops=aflop * units / (ms_launch * 1e-3) / 1e12 if ms_launch > 0 else 0.0)
        table.append(row)
    table.sort(key=lambda r: -r["ms_total"])
    top = table[0]
    if top["bound"] == "tensor":
        roof = {"bound": "tensor", "achieved": top["tflops"], "peak": pk["tf_sustained"], "unit": "TFLOP/s",
                "frac": top["tflops"] / pk["tf_sustained"], "hbm_gbs_algorithmic": top["gbs_algorithmic"],
                "hbm_frac": top["gbs_algorithmic"] / pk["hbm_gbs"]}
    else:
        roof = {"bound": "hbm", "achieved": top["gbs_algorithmic"], "peak": pk["hbm_gbs"], "unit": "GB/s",
                "frac": top["gbs_algorithmic"] / pk["hbm_gbs"], "moved_gbs": top["gbs_moved"]}
    kname = f"{top['kind']}@stage{top['stage']}"
    traffic, traffic_src = None, None
    try:                                                      # DRAM bytes per launch from the committed ncu capture
        with open(os.path.join(ROOT, "profiles", "traffic_r02.json")) as f:
            tr = json.load(f).get(kname)
        if tr:
            traffic, traffic_src = tr["dram_bytes_per_frame"] * top["frames_per_launch"], tr["source"]
    except OSError:
        pass
    roof.update({"kernel": kname, "share_of_step": top["share"], "traffic": traffic, "traffic_source": traffic_src,
                 "algorithmic_bytes_per_launch": kernel_work(geom, top["kind"], top["stage"])[1] * top["frames_per_launch"],
                 "peak_source": pk["source"], "ms_per_launch": top["ms_per_launch"]})
    if args.kernels_out:
        with open(args.kernels_out, "w") as f:
            json.dump({"config": {"B": B, "T": T, "ms_per_step": ms_step, "value": value}, "kernels": table}, f, indent=1)

    # ---- streaming per-frame latency (second half of BASELINE.json's metric) ---------------------------
    streaming = None
    if not args.no_streaming:
        y = None
        streaming = streaming_leg(dec, geom, dev, sd)

    # ---- diffusion-head sampler ---------------------------------------------------------------
    sampler = None
    if not args.no_streaming:
        try:
            sampler = sampler_leg(dev, parity=not args.no_cpu_baseline)
        except Exception as e:                                  # report, do not hide
            sampler = {"error": f"{type(e).__name__}: {e}"[:300]}
            torch.cuda.empty_cache()

    # ---- encoder leg (BASELINE.json configs[4]) ---------------------------------------------------------
    encoder = None
    if not args.no_encoder:
        y = None
        dec._ws = None                                          # release the decode workspace
        torch.cuda.empty_cache()
        try:
            encoder = encoder_leg(geom, dev, pk, args.steps, args.warmup, kernels_out=args.encoder_kernels_out,
                                  cpu_baseline=not args.no_cpu_baseline)
        except Exception as e:                                  # report, do not hide
            encoder = {"error": f"{type(e).__name__}: {e}"[:300]}

    # ---- stock-PyTorch eager decode on the same GPU (context for the CPU baseline; bounded sample) ----------------
    eager = None
    if not args.no_torch_eager:
        try:
            dec._ws = None
            torch.cuda.empty_cache()
            eager = torch_eager_leg(sd, geom, dev, dec)
        except Exception as e:                                  # report, do not hide
            eager = {"error": f"{type(e).__name__}: {e}"[:300]}
            torch.cuda.empty_cache()

    # ---- CPU baseline (reported, not the target): bounded sample on the host cores -----------------
    cpu, parity = None, None
    if not args.no_cpu_baseline:
        v, sec, threads = cpu_reference_audio_s_per_s(sd, 3, 1)
        cpu = {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
               "sample": CPU_SAMPLE + f"; {sec:.2f} s per decode", "torch": torch.__version__}
        # ---- parity at the headline's own size: one utterance of the K2 batch against the CPU reference ------------------
        u = 17 % B
        y = dec.decode(lat)                                     # the timed call, once more (same routing, same tiles)
        parity = {"k2": parity_decode_sample(sd, geom, lat_host[u:u + 1].float(), y[u:u + 1],
                                             f"utterance {u} of the batch {B} x {T} frames ({T / FRAME_RATE:.0f} s), bf16 output "
                                             f"vs the CPU reference (fp32)"),
                  "k3": ({"ok": longform.get("parity_ok"), "windows": longform.get("parity_windows")} if longform else None),
                  "k4": ({"ok": streaming.get("schedule", {}).get("parity_ok"), "streams": streaming.get("schedule", {}).get("parity")}
                         if streaming else None),
                  "k5": (encoder.get("parity") if encoder else None),
                  "tolerance": "SNR >= 40 dB and max|d| <= 2e-2 of the reference peak (BASELINE.json north_star)"}
        del y

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
        "data": "synthetic",
        "config": workload_config(B, T),
        "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roof, "cpu_baseline": cpu,
        "parity": parity,
        "streaming": streaming, "sampler": sampler, "encoder": encoder, "longform": longform,
        "torch_eager_gpu": eager,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
