#!/usr/bin/env python
"""N4 alone: time vvdec.B200DiffusionSampler.sample_speech_tokens (256 streams x 20 steps) eager and as one CUDA graph.
    python tools/sampler_step.py [--streams 256] [--steps 20] [--opt key=value ...]"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "vibevoice-rocm_b200")):
    sys.path.insert(0, p)
import torch          # noqa: E402
import bench          # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--streams", type=int, default=256)
ap.add_argument("--steps", type=int, default=20)
a = ap.parse_args()
out = bench.sampler_leg(torch.device("cuda", 0), n_streams=a.streams, steps=a.steps, parity=False)
print({k: (round(v, 3) if isinstance(v, float) else v) for k, v in out.items() if k != "workload"})
