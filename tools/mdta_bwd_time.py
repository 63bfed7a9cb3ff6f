"""Warm CUDA-event times of the two halves of the attention-core backward at the cfg4 level shapes (B8).
python tools/mdta_bwd_time.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import b200ir
from b200ir import ops
B = 8
for C, heads, H in ((48, 1, 128), (96, 1, 128), (96, 2, 64), (192, 4, 32), (384, 8, 16)):
    g = torch.Generator().manual_seed(0)
    qkv = (torch.randn(B, 3 * C, H, H, generator=g) * 0.5).cuda().bfloat16()
    partials = ops.mdta_gram(qkv, heads)
    w = (torch.randn(C, C, generator=g) / C ** 0.5).cuda()
    temp = torch.ones(heads, device="cuda")
    _, attn = ops.mdta_softmax_fold(partials, temp, w, want_attn=True)
    dM = torch.randn(B, C, C, generator=g).cuda()
    res = {}
    for name, fn in (("chain", lambda: ops.mdta_bwd_chain(partials, attn, temp, w, dM)),
                     ("fold", lambda: ops.mdta_bwd_fold(attn, w, dM, heads))):
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(50):
            fn()
        e1.record()
        torch.cuda.synchronize()
        res[name] = e0.elapsed_time(e1) / 50 * 1e3
    print("C=%3d heads=%d HW=%5d nslab=%2d : chain(+dtemp fold) %6.1f us   fold(+dW fold) %6.1f us" %
          (C, heads, H * H, partials.shape[2], res["chain"], res["fold"]))
