"""Per-role clock64 timeline of CTA 0 of the fused LN + 1x1 + depthwise kernel (needs a -DB200IR_TIMELINE build).
Usage: B200IR_LIB=tools/ab/lib_tl.so python tools/fused_timeline.py [first_row n_rows]"""
import ctypes
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import b200ir  # noqa: E402
from b200ir import ops  # noqa: E402
L = b200ir._lib
NAMES = {1: "P.xempty", 10: "M.top", 11: "M.xnorm", 12: "M.rowempty", 13: "M.commit", 20: "L.top", 21: "L.xfull",
         22: "L.done", 23: "L.stats", 24: "L.reduced", 25: "L.normalized", 26: "L.fenced", 14: "M.acc_issued", 30: "C.top", 31: "C.rowfull", 32: "C.done"}
C, Hd, HW, B = 96, 255, 256, 16
g = torch.Generator().manual_seed(0)
x = (torch.rand(B, C, HW, HW, generator=g) - 0.5).bfloat16().cuda()
w = ((torch.rand(2 * Hd, C, generator=g) - 0.5) / C ** 0.5).cuda()
lw, lb = (torch.rand(C, generator=g) + 0.5).cuda(), (torch.rand(C, generator=g) - 0.5).cuda()
dw = ((torch.rand(2 * Hd, 1, 3, 3, generator=g) - 0.5) * 0.5).cuda()
wf = ops.pack_fused_proj_dw(w, None, L.LN_WITHBIAS, lw, lb, Hd, True)
out = torch.empty((B, Hd, HW, HW), device="cuda", dtype=torch.bfloat16)
lib = ctypes.CDLL(L.LIB_PATH)
lib.b200ir_debug_fused_timeline.argtypes = [ctypes.c_void_p, ctypes.c_int]
for _ in range(3):
    ops.ln_proj_dw(x, wf, dw, None, L.LN_WITHBIAS, Hd, True, out=out)
lib.b200ir_debug_fused_timeline(None, 0)
ops.ln_proj_dw(x, wf, dw, None, L.LN_WITHBIAS, Hd, True, out=out)
buf = (ctypes.c_ulonglong * (2 * 16384))()
n = lib.b200ir_debug_fused_timeline(buf, 16384)
ev = sorted((buf[2 * i + 1], buf[2 * i] >> 32, buf[2 * i] & 0xffffffff) for i in range(n) if buf[2 * i + 1])
base = ev[0][0]
print(len(ev), "events, span", ev[-1][0] - base, "cycles")
r0, nr = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (100, 6)
for clk, tag, ident in ev:
    row = ident - (2 if tag >= 30 else 0)      # conv events carry the newest input row (centre + 1) -> output row id
    if r0 <= row < r0 + nr:
        print(f"{clk - base:9d}  row {ident:4d}  {NAMES.get(tag, tag)}")
# per-role mean period
import collections
per = collections.defaultdict(list)
for clk, tag, ident in ev:
    per[tag].append(clk)
for tag, v in sorted(per.items()):
    if len(v) > 50:
        d = [b - a for a, b in zip(v[40:-1], v[41:])]
        print(NAMES.get(tag, tag), "mean period", sum(d) / len(d), "n", len(v))
