"""Micro-benchmarks of the backward-only kernels at the cfg4 (B8, 128x128) level shapes."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import b200ir
from b200ir import ops
L = b200ir._lib
bf = torch.bfloat16


def timeit(fn, iters=20):
    """us per call, replayed from a CUDA graph of `iters` calls (no host launch overhead in the number)."""
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    st = torch.cuda.Stream()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.stream(st):
        fn()
        with torch.cuda.graph(g, stream=st):
            for _ in range(iters):
                fn()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (3 * iters) * 1e3


for (B, C, H, W) in [(8, 48, 128, 128), (8, 96, 128, 128), (8, 96, 64, 64), (8, 192, 32, 32), (8, 384, 16, 16)]:
    x = (torch.rand(B, C, H, W, device="cuda") - 0.5).to(bf)
    g = (torch.rand(B, C, H, W, device="cuda") - 0.5).to(bf)
    r = (torch.rand(B, C, H, W, device="cuda") - 0.5).to(bf)
    lw = torch.rand(C, device="cuda") + 0.5
    lb = torch.rand(C, device="cuda") - 0.5
    mb = B * C * H * W * 2 / 1e6
    t1 = timeit(lambda: ops.ln_bwd(x, g, L.LN_WITHBIAS, lw, r))
    t2 = timeit(lambda: ops.ln_fwd(x, L.LN_WITHBIAS, lw, lb))
    print(f"B{B} C{C} {H}x{W}: ln_bwd(+reduce) {t1:7.1f} us ({4*mb/t1:5.2f} TB/s)   ln_fwd {t2:7.1f} us ({2*mb/t2:5.2f} TB/s)")
