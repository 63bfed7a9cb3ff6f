"""Single launches of backward-only kernels at cfg4 shapes (for ncu --set full)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import b200ir
from b200ir import ops
L = b200ir._lib
bf = torch.bfloat16
B, C, H, W = 8, 48, 128, 128
x = (torch.rand(B, C, H, W, device="cuda") - 0.5).to(bf)
g = (torch.rand(B, C, H, W, device="cuda") - 0.5).to(bf)
r = (torch.rand(B, C, H, W, device="cuda") - 0.5).to(bf)
lw = torch.rand(C, device="cuda") + 0.5
for _ in range(2):
    ops.ln_bwd(x, g, L.LN_WITHBIAS, lw, r)
torch.cuda.synchronize()
