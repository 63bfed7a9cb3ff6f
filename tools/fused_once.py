"""One shape of the fused kernel, a few launches (ncu target).  python tools/fused_once.py [C Hd HW B]"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import b200ir
from b200ir import ops
L = b200ir._lib
C, Hd, HW, B = (int(v) for v in sys.argv[1:5]) if len(sys.argv) >= 5 else (96, 255, 256, 16)
g = torch.Generator().manual_seed(0)
x = (torch.rand(B, C, HW, HW, generator=g) - 0.5).bfloat16().cuda()
w = ((torch.rand(2 * Hd, C, generator=g) - 0.5) / C ** 0.5).cuda()
lw, lb = (torch.rand(C, generator=g) + 0.5).cuda(), (torch.rand(C, generator=g) - 0.5).cuda()
dw = ((torch.rand(2 * Hd, 1, 3, 3, generator=g) - 0.5) * 0.5).cuda()
wf = ops.pack_fused_proj_dw(w, None, L.LN_WITHBIAS, lw, lb, Hd, True)
out = torch.empty((B, Hd, HW, HW), device="cuda", dtype=torch.bfloat16)
for _ in range(4):
    ops.ln_proj_dw(x, wf, dw, None, L.LN_WITHBIAS, Hd, True, out=out)
torch.cuda.synchronize()
print("ok", float(out.float().abs().mean()))
