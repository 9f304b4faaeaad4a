"""Generate tests/golden/pcm16_*.npz with the reference's OWN convert_to_16_bit_wav (build container only).

TEST INFRASTRUCTURE ONLY.  Usage:  python -m oracle.gen_golden_pcm16

main.py (and demo/gradio_demo.py) import gradio at module level, which this image does not have, so the function's
source is lifted out of /root/reference/main.py with ``ast`` and executed unmodified with numpy and torch in scope —
the reference's code runs, nothing of it is copied into the repo.  Inputs are regenerated from seeds (oracle/pcm16.py
make_wave); only the int16 outputs are stored.
"""
from __future__ import annotations

import ast
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
from oracle.pcm16 import make_wave          # noqa: E402

REF_FILE = "/root/reference/main.py"
GOLD = os.path.join(ROOT, "tests", "golden")

# (name, B, L, seed, scale, bf16): streamed chunks (3200), a ragged length, quiet rows only (no normalisation), a long row
CASES = [
    ("pcm16_chunks_B6_L3200", 6, 3200, 11, 2.0, False),
    ("pcm16_chunks_bf16_B5_L3200", 5, 3200, 12, 3.0, True),
    ("pcm16_ragged_B3_L4097", 3, 4097, 13, 1.5, False),
    ("pcm16_quiet_B4_L800", 4, 800, 14, 0.2, False),
    ("pcm16_long_B2_L40003", 2, 40003, 15, 6.9, False),
]


def reference_function():
    tree = ast.parse(open(REF_FILE).read())
    fn = next(n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "convert_to_16_bit_wav")
    mod = ast.Module(body=[fn], type_ignores=[])
    ns = {"np": np, "torch": torch}
    exec(compile(mod, REF_FILE, "exec"), ns)
    return ns["convert_to_16_bit_wav"], (fn.lineno, fn.end_lineno)


def main():
    ref, lines = reference_function()
    os.makedirs(GOLD, exist_ok=True)
    for name, B, L, seed, scale, bf16 in CASES:
        x = make_wave(B, L, seed, scale, bf16)
        out = np.stack([ref(x[b].copy()) for b in range(B)])          # one chunk per sample, as main.py:937-951 does
        assert out.dtype == np.int16
        meta = {"B": B, "L": L, "seed": seed, "scale": scale, "bf16": bf16,
                "reference": f"main.py:{lines[0]}-{lines[1]} convert_to_16_bit_wav", "numpy": np.__version__}
        np.savez_compressed(os.path.join(GOLD, name + ".npz"), pcm=out, meta=json.dumps(meta))
        print(name, out.shape, "peaks", np.abs(x).max(axis=1))


if __name__ == "__main__":
    main()
