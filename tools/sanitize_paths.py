"""compute-sanitizer target (race detection): smoke-sized offline decode, streaming decode (in-place state
update: the only in-place hazard on the path), offline + streaming encode and PCM16 — every kernel class of the product
path once, on the full channel ladder with one block per stage.  Run as
    compute-sanitizer --tool memcheck|racecheck python tools/sanitize_paths.py [--small]
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "vibevoice-rocm_b200")):
    sys.path.insert(0, p)
import torch  # noqa: E402
import vvdec  # noqa: E402
from vvdec.synthetic import random_encoder_state_dict, random_state_dict  # noqa: E402

small = "--small" in sys.argv
g = vvdec.Geometry(depths=(1, 1, 1, 1, 1, 1, 1))
dec = vvdec.B200AcousticDecoder(g, device=0).load_state_dict(random_state_dict(g))
enc = vvdec.B200AcousticEncoder(g, device=0).load_state_dict(random_encoder_state_dict(g))
gen = torch.Generator().manual_seed(1)
lat = torch.randn(2, 2 if small else 3, 64, generator=gen).cuda()
y = dec.decode(lat)
cache = dec.new_cache(4)
idx = torch.tensor([0, 2], device="cuda")
ys = [dec.decode(lat[:, t:t + 1], cache=cache, sample_indices=idx, use_cache=True) for t in range(lat.shape[1])]
torch.cuda.synchronize()
print("stream==offline", bool(torch.equal(torch.cat(ys, -1), y)))
m = enc.encode(y).mean
ecache = enc.new_cache(4)
ms = [enc.encode(c, cache=ecache, sample_indices=idx, use_cache=True).mean for c in ys]
p = vvdec.pcm16(y)
torch.cuda.synchronize()
print("done", tuple(y.shape), tuple(m.shape), tuple(p.shape), float(y.float().abs().max()))
