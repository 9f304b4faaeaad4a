"""Timeline of block_ws_kernel / block_tma_kernel (build with -DVV_BW_TRACE, VVDEC_LIB=that build): clock64 stamps of CTA 0,
tiles 8..23.   python tools/micro/bw_trace.py [stage]"""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (ROOT, os.path.join(ROOT, "vibevoice-rocm_b200")):
    sys.path.insert(0, p)
import torch, vvdec
from vvdec.synthetic import random_state_dict
g = vvdec.Geometry()
dec = vvdec.B200AcousticDecoder(g, device=0)
dec.load_state_dict(random_state_dict(g))
lib, h = dec._lib, dec._h
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
stage = int(sys.argv[1]) if len(sys.argv) > 1 else 4          # 4: block_ws_kernel; 5 / 6: block_tma_kernel
Cc = g.stage_channels(stage)
rows = 5760 * (102400 // Cc)
P = lambda t: C.c_void_p(t.data_ptr())
x = torch.randn(1, rows, Cc, device="cuda"); out = torch.empty_like(x)
for _ in range(3):
    assert lib.vv_kernel_block(h, stage, 0, P(x), P(out), 1, rows, st) == 0
torch.cuda.synchronize()
buf = (C.c_longlong * 512)()
raw = C.CDLL(os.environ["VVDEC_LIB"])
assert raw.vv_debug_read_bw_trace(buf, 512) == 0
names = ["load", "R:x_full", "R:st_rdy", "S2:start", "S2:y_full", "R:y_seen", "R:a_full", "R:d2init", "G1:start", "G1:last",
         "G2:commit", "E2:d2full", "E2:done", "XY:y_rdy", "XY:st_read"]
t0 = min(v for v in buf if v > 0)
e1 = ["c0:wait", "c0:d1full", "c0:ld0", "c0:st0", "c0:ld1", "c0:st1", "c0:hfull", "", "c2:wait", "c2:d1full"]
if stage != 4:
    e1 = ["wait", "d1full", "gelu0 done", "h_empty", "", "", "hfull"]
for ti in range(16):
    row = [buf[ti * 32 + s] - t0 if buf[ti * 32 + s] else -1 for s in range(15)]
    print(f"it={8 + ti:3d} " + " ".join(f"{n}={v:6d}" for n, v in zip(names, row)))
    row = [buf[ti * 32 + 16 + s] - t0 if buf[ti * 32 + 16 + s] else -1 for s in range(10)]
    print("   E1   " + " ".join(f"{n}={v:6d}" for n, v in zip(e1, row) if n))
