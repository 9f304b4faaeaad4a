#!/usr/bin/env python
"""K4 under a profiler: `--steps` frame-by-frame decode steps of `--streams` concurrent streams (random weights, device-
resident state), nothing else.  For `ncu --metrics gpu__time_duration.sum` launch lists of ONE streaming step.
    python tools/stream_step.py --streams 256 --steps 8"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "vibevoice-rocm_b200")):
    sys.path.insert(0, p)
import torch          # noqa: E402
import vvdec          # noqa: E402
from vvdec.synthetic import random_state_dict   # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--streams", type=int, default=256)
ap.add_argument("--steps", type=int, default=8)
ap.add_argument("--profile", action="store_true", help="per-kernel CUDA-event times of the step (eager launches, warm caches)")
a = ap.parse_args()
g = vvdec.Geometry()
dec = vvdec.B200AcousticDecoder(g, device=0)
dec.load_state_dict(random_state_dict(g))
cache = dec.new_cache(a.streams)
idx = torch.arange(a.streams, device="cuda", dtype=torch.int32)
lat = torch.randn(a.streams, 1, g.vae_dim, device="cuda").to(torch.bfloat16)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for i in range(a.steps):
    if i == a.steps - 1:
        e0.record()
    y = dec.decode(lat, cache=cache, sample_indices=idx, use_cache=True)
e1.record()
torch.cuda.synchronize()
print(f"last step {e0.elapsed_time(e1):.3f} ms, launches per step {dec.take_launch_count() // a.steps}, out {tuple(y.shape)}")
if a.profile:
    dec.set_profiling(True)
    n = 20
    for _ in range(n):
        dec.decode(lat, cache=cache, sample_indices=idx, use_cache=True)
    rows = dec.read_profile()
    dec.set_profiling(False)
    tot = sum(r["ms"] for r in rows) / n
    print(f"per step (event-timed, sum {tot * 1e3:.0f} us):")
    for r in sorted(rows, key=lambda r: (r["stage"], r["kind"])):
        print(f"  stage {r['stage']} {r['kind']:12s} {r['launches'] // n:3d} launches  {r['ms'] / n * 1e3:7.1f} us  "
              f"({r['ms'] / r['launches'] * 1e3:6.1f} us each)")
