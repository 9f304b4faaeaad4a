#!/usr/bin/env python
"""Run ONE kernel class of the decode path at a K2-pass shape, alone (for ncu captures and quick A/B timing).

    python tools/kbench.py --kind ffn --stage 6 --frames 2048 --iters 5

Weights are random (vvdec.synthetic); inputs are random and larger than L2 at the default sizes.
Prints the CUDA-event time per call and the bytes/s this design moves."""
import argparse
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "vibevoice-rocm_b200")):
    sys.path.insert(0, p)
import torch          # noqa: E402
import vvdec          # noqa: E402
from vvdec.synthetic import random_state_dict   # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--kind", default="ffn", choices=["mixer", "ffn", "upsample", "head", "block"])
    ap.add_argument("--stage", type=int, default=6)
    ap.add_argument("--frames", type=int, default=2048)
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--gemm-path", type=int, default=0)
    ap.add_argument("--inplace", action="store_true", help="ffn: out == resid (C = 256: TMA reduce-add mode)")
    a = ap.parse_args()
    g = vvdec.Geometry()
    # only the stage under test needs real depth; keep the load fast
    dec = vvdec.B200AcousticDecoder(g, device=0, gemm_path=a.gemm_path)
    dec.load_state_dict(random_state_dict(g))
    lib, h = dec._lib, dec._h
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    Cc = g.stage_channels(a.stage)
    M = 1
    for r in g.ratios[:a.stage]:
        M *= r
    rows = a.frames * M
    P = lambda t: C.c_void_p(t.data_ptr())
    if a.kind == "mixer":
        x = torch.randn(1, rows, Cc, device="cuda")
        y = torch.empty_like(x)
        ab = torch.empty(1, rows, Cc, device="cuda", dtype=torch.bfloat16)
        call = lambda: lib.vv_kernel_mixer(h, a.stage, 0, P(x), P(y), P(ab), 1, rows, st)
        moved = rows * Cc * 10
    elif a.kind == "ffn":
        ab = torch.randn(rows, Cc, device="cuda").to(torch.bfloat16)
        res = torch.randn(rows, Cc, device="cuda")
        out = torch.empty_like(res)
        hid = torch.empty(rows, 4 * Cc, device="cuda", dtype=torch.bfloat16)
        if a.inplace:
            out = res
        call = lambda: lib.vv_kernel_ffn(h, a.stage, 0, P(ab), P(res), P(out), P(hid), 1, rows, st)
        moved = rows * Cc * 10
    elif a.kind == "block":
        x = torch.randn(1, rows, Cc, device="cuda")
        out = torch.empty_like(x)
        call = lambda: lib.vv_kernel_block(h, a.stage, 0, P(x), P(out), 1, rows, st)
        moved = rows * Cc * 8
    elif a.kind == "upsample":
        cin = g.stage_channels(a.stage - 1)
        r = g.ratios[a.stage - 1]
        rin = rows // r
        xb = torch.randn(1, rin + 1, cin, device="cuda").to(torch.bfloat16)
        y = torch.empty(1, rows, Cc, device="cuda")
        call = lambda: lib.vv_kernel_upsample(h, a.stage, P(xb), P(y), 1, rin, st)
        moved = rin * cin * 2 + rows * Cc * 4
    else:
        x = torch.randn(1, rows, 32, device="cuda")
        out = torch.empty(1, rows, device="cuda")
        call = lambda: lib.vv_kernel_head(h, P(x), P(out), 0, 1, rows, st)
        moved = rows * 32 * 4 + rows * 4
    for _ in range(2):
        assert call() == 0, lib.vv_last_error()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.iters):
        assert call() == 0
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.iters
    print(f"{a.kind} stage {a.stage} C={Cc} rows={rows}: {ms:.3f} ms/call, {moved / ms / 1e6:.0f} GB/s moved")


if __name__ == "__main__":
    main()
