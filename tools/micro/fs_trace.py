"""Timeline of ffn_stream_kernel (build with -DVV_FS_TRACE, VVDEC_LIB=that build): clock64 stamps of CTA 0."""
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
stage = int(sys.argv[1]) if len(sys.argv) > 1 else 4            # 4: C = 128 (set VV_BLOCK_WS irrelevant: vv_kernel_ffn), 3: C = 256
Cc = g.stage_channels(stage)
rows = 5760 * (102400 // Cc) if stage >= 4 else 5760 * 200
P = lambda t: C.c_void_p(t.data_ptr())
ab = torch.randn(rows, Cc, device="cuda").to(torch.bfloat16)
res = torch.randn(rows, Cc, device="cuda"); out = torch.empty_like(res)
hid = torch.empty(16, device="cuda")
for _ in range(3):
    assert lib.vv_kernel_ffn(h, stage, 0, P(ab), P(res), P(out), P(hid), 1, rows, st) == 0
torch.cuda.synchronize()
buf = (C.c_longlong * 512)()
raw = C.CDLL(os.environ["VVDEC_LIB"])
assert raw.vv_debug_read_trace(buf, 512) == 0
names = ["G1 w:d1e", "G1 w:w1", "G1 issued", "G2 w:w2", "G2 w:h", "G2 issued", "E1 start", "E1 d1full", "E1 ld done",
         "E1 gelu done", "E1 h_empty", "E1 arrived"]
t0 = min(v for v in buf if v > 0)
for gi in range(32):
    row = [buf[gi * 16 + s] - t0 if buf[gi * 16 + s] else -1 for s in range(12)]
    print(f"g={256 + gi:4d} " + " ".join(f"{n}={v:6d}" for n, v in zip(names, row)))
