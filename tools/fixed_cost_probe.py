"""Fixed cost per launch of the main kernels: a tiny problem (one or two tiles), 64 back-to-back launches replayed from a
CUDA graph.  python tools/fixed_cost_probe.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import b200ir
from b200ir import ops
L = b200ir._lib
dev = "cuda"
bf = torch.bfloat16


def graph_time(fn, n=64, reps=20):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for _ in range(3):
            fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(n):
            fn()
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps / n * 1e3


def rnd(*s, dtype=bf):
    return (torch.rand(*s, device=dev) - 0.5).to(dtype)


for (B, C, H) in ((1, 48, 16), (8, 48, 128), (8, 384, 16), (8, 192, 32)):
    x = rnd(B, C, H, H)
    p = ops.pack_conv1x1(rnd(3 * C, C, dtype=torch.float32), None, L.LN_WITHBIAS, torch.ones(C, device=dev), torch.zeros(C, device=dev))
    out = torch.empty(B, 3 * C, H, H, device=dev, dtype=bf)
    t1 = graph_time(lambda: ops.gemm_tc(x, p, 3 * C, L.LN_WITHBIAS, out=out))
    dy = rnd(B, 3 * C, H, H)
    t2 = graph_time(lambda: ops.conv_wgrad(dy, x, 1))
    w = rnd(3 * C, 1, 3, 3, dtype=torch.float32)
    t3 = graph_time(lambda: ops.dwconv3x3(dy, w, None, out=out))
    parts = torch.rand(64, 3 * C * C, device=dev)
    o = torch.empty(3 * C * C, device=dev)
    t4 = graph_time(lambda: ops.reduce_partials(parts, 64, 3 * C * C, out=o))
    t5 = graph_time(lambda: ops.ln_bwd(x, x, L.LN_WITHBIAS, torch.ones(C, device=dev), None, torch.zeros(C, device=dev)))
    print(f"B{B} C{C} {H}x{H}: gemm_tc {t1:6.2f} us   conv_wgrad(+fold) {t2:6.2f} us   dwconv3x3 {t3:6.2f} us   reduce_partials {t4:6.2f} us   ln_bwd(+fold) {t5:6.2f} us")
