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
# CPU arm: the oracle port, timed on the host cores (checker code used as the reported CPU baseline)
# --------------------------------------------------------------------------------------------------
def cpu_port_audio_s_per_s(sd, steps: int, warmup: int):
    import numpy as np
    from oracle.geometry import FULL
    from oracle.vv_oracle import OracleDecoder
    from oracle.weights import make_latents
    dec = OracleDecoder({k: v.numpy() for k, v in sd.items()}, FULL)
    lat = make_latents(CPU_SAMPLE_B, CPU_SAMPLE_T, FULL.vae_dim, 1)
    for _ in range(warmup):
        dec.decode(lat)
    ts = []
    for _ in range(steps):
        t0 = time.perf_counter()
        dec.decode(lat)
        ts.append(time.perf_counter() - t0)
    sec = sum(ts) / len(ts)
    return CPU_SAMPLE_B * CPU_SAMPLE_T / FRAME_RATE / sec, sec


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    from vvdec.synthetic import random_state_dict
    torch.set_num_threads(os.cpu_count() or 1)
    sd = random_state_dict()
    v, sec = cpu_port_audio_s_per_s(sd, max(1, args.steps), max(0, args.warmup))
    sample = f"batch {CPU_SAMPLE_B} x {CPU_SAMPLE_T} frames (10 s) per step, numpy fp32 port of the reference path"
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"acoustic decoder, batch {K2_B} x 60 s ({K2_T} latent frames) per GPU; CPU arm "
                               f"times a bounded sample: {sample}"},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": os.cpu_count(), "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


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
    if kind == "block_fused":  # whole Block1D in one launch (C <= 64): SURVEY's mixer half + FFN half = 8*C*M bytes;
        return "hbm", 8 * cm, 20 * cm + 16 * C * cm, 8 * cm   # this design moves exactly that (fp32 x in, fp32 x out)
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


def streaming_leg(dec, geom, dev, n_streams=256, steps=300, warmup=30):
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
    return out


def torch_eager_leg(sd, geom, dev, dec, B=8, T=450, steps=3):
    """Context number, not an arm: the same decode written with stock PyTorch ops (cuDNN depthwise / transposed conv,
    cuBLAS linear, F.gelu — the calls the reference module makes, modular_vibevoice_tokenizer.py:914-951) in bf16 eager on
    the same GPU and weights, on a bounded sample.  Lives in bench.py only; the product path never uses it."""
    import torch
    import torch.nn.functional as F
    w = {k: v.to(dev, torch.bfloat16) for k, v in sd.items()}
    n = len(geom.depths)

    def rms(x, wt):                                   # ConvRMSNorm :77-91 (fp32 statistics, cast, then * weight)
        xt = x.transpose(1, 2)
        xf = xt.float()
        xn = (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + geom.eps)).to(x.dtype) * wt
        return xn.transpose(1, 2)

    def decode(lat):                                  # [B, T, D] -> [B, 1, T*hop]
        x = lat.transpose(1, 2)
        x = F.conv1d(F.pad(x, (6, 0)), w["decoder.upsample_layers.0.0.conv.conv.weight"], w["decoder.upsample_layers.0.0.conv.conv.bias"])
        for i in range(n):
            if i > 0:
                r = geom.ratios[i - 1]
                p = f"decoder.upsample_layers.{i}.0.convtr.convtr."
                y = F.conv_transpose1d(x, w[p + "weight"], w[p + "bias"], stride=r)
                x = y[..., :y.shape[-1] - r]          # causal trim of k - r samples on the right (:563-571)
            C = x.shape[1]
            for j in range(geom.depths[i]):
                p = f"decoder.stages.{i}.{j}."
                h = rms(x, w[p + "norm.weight"])
                h = F.conv1d(F.pad(h, (6, 0)), w[p + "mixer.conv.conv.conv.weight"], w[p + "mixer.conv.conv.conv.bias"], groups=C)
                x = x + h * w[p + "gamma"][None, :, None]
                h = rms(x, w[p + "ffn_norm.weight"]).transpose(1, 2)
                h = F.linear(F.gelu(F.linear(h, w[p + "ffn.linear1.weight"], w[p + "ffn.linear1.bias"])),
                             w[p + "ffn.linear2.weight"], w[p + "ffn.linear2.bias"]).transpose(1, 2)
                x = x + h * w[p + "ffn_gamma"][None, :, None]
        return F.conv1d(F.pad(x, (6, 0)), w["decoder.head.conv.conv.weight"], w["decoder.head.conv.conv.bias"])

    gen = torch.Generator().manual_seed(9)
    lat = torch.randn(B, T, geom.vae_dim, generator=gen).to(torch.bfloat16).to(dev)
    with torch.no_grad():
        y = decode(lat[:1, :8])                       # agreement check on a small piece (both bf16 paths vs each other)
        ours = dec.decode(lat[:1, :8])
        err = (y.float() - ours.float())
        snr = float(10 * torch.log10(ours.float().pow(2).sum() / err.pow(2).sum().clamp_min(1e-30)))
        del y, ours, err
        decode(lat)                                   # warm-up (cuDNN heuristics, allocator)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            decode(lat)
        e1.record()
        torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    del w
    torch.cuda.empty_cache()
    return {"what": "stock PyTorch ops (cuDNN/cuBLAS), bf16 eager, same GPU and weights; context only",
            "sample": f"batch {B} x {T} frames", "ms_per_step": ms, "value": B * T / FRAME_RATE / (ms / 1e3), "unit": UNIT,
            "snr_vs_b200_path_db": snr, "torch": torch.__version__}


def longform_leg(dec, geom, dev, world, rank, dist, steps, frames=40_500):
    """BASELINE.json configs[2]: one 90-minute utterance (40,500 latent frames) split into time chunks across the ranks,
    each rank recomputing its causal halo (58 frames); no collective on the data path, one NCCL gather of the finished
    waveform slices to rank 0 at the end (strong scaling: the utterance is fixed, N varies).  All ranks call this."""
    import torch
    from vvdec.shard import plan_shards
    halo, hop, D = dec.halo_frames, geom.hop, geom.vae_dim
    shards = plan_shards([frames], world, halo)
    mine = sorted((s for s in shards if s.rank == rank), key=lambda s: s.start)
    gen = torch.Generator().manual_seed(3)
    lat_all = torch.randn(1, frames, D, generator=gen)
    seg = frames // 4
    for k, mu in enumerate((-0.5, 0.0, 0.5, 1.0)):                    # four "speakers": four segments, shifted means
        lat_all[:, k * seg:(k + 1) * seg if k < 3 else frames] += mu
    lats = [lat_all[:, s.start - s.halo:s.end].to(torch.bfloat16).to(dev) for s in mine]
    owned = sum(s.frames for s in mine)
    per_rank = [sum(s.frames for s in shards if s.rank == r) for r in range(world)]
    cap = max(per_rank) * hop
    buf = torch.zeros(cap, dtype=torch.bfloat16, device=dev)
    gl = [torch.empty(cap, dtype=torch.bfloat16, device=dev) for _ in range(world)] if (rank == 0 and world > 1) else None
    final = torch.empty(frames * hop, dtype=torch.bfloat16, device=dev) if rank == 0 else None

    def step():
        off = 0
        for s, x in zip(mine, lats):
            y = dec.decode(x)                                         # [1, 1, (halo + owned) * hop]
            n = s.frames * hop
            buf[off:off + n].copy_(y.view(-1)[s.halo * hop:])
            off += n

    def gather():
        if world > 1:
            dist.gather(buf, gl, dst=0)
        if rank == 0:
            offs = [0] * world
            for s in sorted(shards, key=lambda s: (s.rank, s.start)):
                src = gl[s.rank] if world > 1 else buf
                n = s.frames * hop
                final[s.start * hop:s.start * hop + n].copy_(src[offs[s.rank]:offs[s.rank] + n])
                offs[s.rank] += n

    def sync():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
    step(); gather(); sync()
    e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    t_dec = t_all = 0.0
    for _ in range(steps):
        sync()
        e0.record(); step(); e1.record(); gather(); e2.record()
        sync()
        t_dec += e0.elapsed_time(e1); t_all += e0.elapsed_time(e2)
    t = torch.tensor([t_dec / steps, t_all / steps], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_dec, ms_all = float(t[0]), float(t[1])
    ok = bool(torch.isfinite(final.float()).all()) if rank == 0 else True
    return {"workload": f"one {frames / FRAME_RATE / 60:.0f}-minute utterance ({frames} frames), time-chunk shards with a "
                        f"{halo}-frame recomputed halo, waveform gathered to rank 0",
            "scaling": "strong", "n_gpus": world, "shards": len(shards), "halo_frames": halo,
            "frames_owned_rank0": owned, "recomputed_halo_frames_total": sum(s.halo for s in shards),
            "ms_decode": ms_dec, "ms_decode_plus_gather": ms_all, "gather_bytes": (frames * hop * 2 if world > 1 else 0),
            "value": frames / FRAME_RATE / (ms_all / 1e3), "unit": UNIT, "finite": ok}


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
    del enc
    cpu = None
    if cpu_baseline:
        # reported CPU figure for the encoder: the numpy port on a bounded sample (batch 1 x 2 s), all host threads
        from oracle.geometry import FULL
        from oracle.vv_oracle import OracleEncoder
        from oracle.weights import make_audio
        oe = OracleEncoder({k: v.numpy() for k, v in sd_cpu.items()}, FULL)
        a = make_audio(1, 48_000, 5)
        oe.encode(a[:, :, :6400])
        t0 = time.perf_counter()
        oe.encode(a)
        sec = time.perf_counter() - t0
        cpu = {"value": 2.0 / sec, "unit": "encoded audio-s/s", "cores": os.cpu_count(), "kind": "port",
               "sample": f"batch 1 x 2 s (48000 samples), numpy fp32 port, {sec:.2f} s"}
    return {"cpu_baseline": cpu,
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
            longform = longform_leg(dec, geom, dev, world, rank, dist, max(2, args.steps // 2))
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
                "frac": top["tflops"] / pk["tf_sustained"]}
    else:
        roof = {"bound": "hbm", "achieved": top["gbs_algorithmic"], "peak": pk["hbm_gbs"], "unit": "GB/s",
                "frac": top["gbs_algorithmic"] / pk["hbm_gbs"], "moved_gbs": top["gbs_moved"]}
    kname = f"{top['kind']}@stage{top['stage']}"
    traffic, traffic_src = None, None
    try:                                                      # DRAM bytes per launch from the committed ncu capture
        with open(os.path.join(ROOT, "profiles", "traffic_r01.json")) as f:
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
        streaming = streaming_leg(dec, geom, dev)

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
    cpu = None
    if not args.no_cpu_baseline:
        torch.set_num_threads(os.cpu_count() or 1)
        v, sec = cpu_port_audio_s_per_s(sd, 2, 1)
        cpu = {"value": v, "unit": UNIT, "cores": os.cpu_count(), "kind": "port",
               "sample": f"batch {CPU_SAMPLE_B} x {CPU_SAMPLE_T} frames (10 s), numpy fp32 port, {sec:.2f} s per decode"}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
        "data": "synthetic",
        "config": {"workload": f"acoustic decoder bf16, batch {B} x {T / FRAME_RATE:.0f} s ({T} latent frames) synthetic "
                               f"latents per GPU, random-init (stress) weights",
                   "l2": "inputs >> L2 (each pass streams >700 MB of activations per utterance group)",
                   "sharding": "utterances per rank, no data-path collective"},
        "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roof, "cpu_baseline": cpu,
        "streaming": streaming, "encoder": encoder, "longform": longform,
        "torch_eager_gpu": eager,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
