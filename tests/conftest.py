"""pytest wiring: markers, import paths, golden-fixture helpers."""
import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "vibevoice-rocm_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "slow: longer CPU test")


def load_golden(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    meta = json.loads(str(z["meta"]))
    return z, meta


def geom_from_meta(meta):
    from oracle.geometry import DecoderGeometry
    g = dict(meta["geom"])
    g["ratios"] = tuple(g["ratios"])
    g["depths"] = tuple(g["depths"])
    return DecoderGeometry(**g)


@pytest.fixture(scope="session")
def weight_cache():
    """(geom, init, seed) -> weights dict; full-geometry weights take seconds to draw."""
    from oracle.weights import make_weights
    cache = {}

    def get(geom, init, seed):
        key = (geom, init, seed)
        if key not in cache:
            if len(cache) >= 2:                       # 1.3 GB each for the full geometry
                cache.pop(next(iter(cache)))
            cache[key] = make_weights(geom, init, seed)
        return cache[key]
    return get


@pytest.fixture(scope="session")
def full_decoder(weight_cache):
    """(B200AcousticDecoder with the golden fixtures' stress weights, the weight dict).  GPU only."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    import vvdec
    from oracle.geometry import FULL
    w = weight_cache(FULL, "stress", 1234)
    dec = vvdec.B200AcousticDecoder(vvdec.Geometry(), device=0)
    dec.load_state_dict({k: torch.from_numpy(v) for k, v in w.items()})
    return dec, w
