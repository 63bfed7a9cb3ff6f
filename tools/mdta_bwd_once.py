"""A few launches of the attention-core backward at one cfg4 level (for ncu --set full -k regex:mdta_bwd_chain).
python tools/mdta_bwd_once.py C heads H"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import b200ir
from b200ir import autograd_ops as A
C, heads, H = (int(a) for a in sys.argv[1:4]) if len(sys.argv) > 3 else (96, 1, 128)
B = 8
g = torch.Generator().manual_seed(0)
qkv = (torch.randn(B, 3 * C, H, H, generator=g) * 0.5).cuda().bfloat16().requires_grad_(True)
temp = torch.ones(heads, 1, 1, device="cuda", requires_grad=True)
w = (torch.randn(C, C, 1, 1, generator=g) / C ** 0.5).cuda().requires_grad_(True)
res = torch.randn(B, C, H, H, generator=g).cuda().bfloat16()
dy = torch.randn(B, C, H, H, generator=g).cuda().bfloat16()
for _ in range(3):
    y = A.mdta_core(qkv, temp, w, None, heads, res)
    y.backward(dy)
torch.cuda.synchronize()
print("ok")
