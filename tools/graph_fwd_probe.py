"""Eager vs CUDA-graph replay of the cfg2 forward (is the eager forward losing time between kernels?)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import b200ir
from oracle import restormer_oracle as O

net = b200ir.Restormer().cuda().eval().set_compute_dtype(torch.bfloat16)
net.load_state_dict(O.synth_state_dict({}, seed=0), strict=True)
xs = [O.synth_image(16, 256, 256, seed=i).cuda() for i in range(12)]


def timeit(fn, n=10):
    for i in range(3):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


with torch.no_grad():
    print("eager   %.3f ms" % timeit(lambda i: net(xs[i % 12])))
    static_in = xs[0].clone()
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        for _ in range(2):
            net(static_in)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=st):
            static_out = net(static_in)
    torch.cuda.synchronize()

    def replay(i):
        static_in.copy_(xs[i % 12])
        g.replay()
    print("graphed %.3f ms" % timeit(replay))
    ref = net(xs[3])
    replay(3)
    torch.cuda.synchronize()
    print("max diff", (ref - static_out).abs().max().item())
