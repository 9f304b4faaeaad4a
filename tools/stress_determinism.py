#!/usr/bin/env python
"""Race hunt without a sanitizer (compute-sanitizer is closed on this pool): random shapes through the whole decode / encode
paths, every call made twice and compared BIT FOR BIT (a data race in the mbarrier / TMEM / TMA protocols shows up as
run-to-run differences), streamed decode in random chunkings compared bit for bit with the offline decode of the same
latents, all outputs finite.  Random weights (vvdec.synthetic), full geometry.
    python tools/stress_determinism.py [--cases 24] [--seed 1]"""
import argparse
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "vibevoice-rocm_b200")):
    sys.path.insert(0, p)
import numpy as np      # noqa: E402
import torch            # noqa: E402
import vvdec            # noqa: E402
from vvdec.synthetic import random_state_dict, random_encoder_state_dict   # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--cases", type=int, default=24)
ap.add_argument("--seed", type=int, default=1)
a = ap.parse_args()
rng = np.random.default_rng(a.seed)
g = vvdec.Geometry()
dec = vvdec.B200AcousticDecoder(g, device=0)
dec.load_state_dict(random_state_dict(g))
enc = vvdec.B200AcousticEncoder(g, device=0)
enc.load_state_dict(random_encoder_state_dict(g))
bad = 0
t0 = time.time()
for case in range(a.cases):
    B = int(rng.integers(1, 13))
    T = int(rng.choice([1, 2, 3, 5, 8, 13, 21, 40, 75, 150, 225, 450]))
    lat = torch.from_numpy(rng.standard_normal((B, T, g.vae_dim)).astype(np.float32)).cuda().to(torch.bfloat16)
    y1 = dec.decode(lat)
    y2 = dec.decode(lat)
    ok = bool(torch.isfinite(y1.float()).all()) and torch.equal(y1, y2)
    # streamed, random chunking, random slots
    cache = dec.new_cache(16)
    slots = torch.from_numpy(rng.permutation(16)[:B].astype(np.int32)).cuda()
    parts, t = [], 0
    while t < T:
        n = int(min(T - t, rng.integers(1, 8)))
        parts.append(dec.decode(lat[:, t:t + n], cache=cache, sample_indices=slots, use_cache=True))
        t += n
    ys = torch.cat(parts, -1)
    ok_s = torch.equal(ys, y1)
    # encode the decoded audio (ragged length), twice
    L = int(y1.shape[-1] - rng.integers(0, 3199))
    au = y1[..., :max(L, 1)].float().contiguous()
    e1 = enc.encode(au).mean
    e2 = enc.encode(au).mean
    ok_e = bool(torch.isfinite(e1.float()).all()) and torch.equal(e1, e2)
    print(f"case {case:2d} B={B} T={T:3d} L={au.shape[-1]:7d} repeat={'ok' if ok else 'DIFF'} stream=={'ok' if ok_s else 'DIFF'} "
          f"encode={'ok' if ok_e else 'DIFF'}", flush=True)
    bad += (not ok) + (not ok_s) + (not ok_e)
print(f"{a.cases} cases, {bad} failures, {time.time() - t0:.1f} s")
sys.exit(1 if bad else 0)
