"""A/B: the cfg2 forward (B16 256x256 bf16) as ONE chain vs. the batch cut into S independent sub-batches on S streams inside
the same CUDA graph (S parallel branches: the latency-bound deep-level kernels of one branch overlap those of the others).
python tools/split_fwd_probe.py [S ...]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import b200ir
from b200ir import ops
from oracle import restormer_oracle as O

B, H = 16, 256
net = b200ir.Restormer().cuda().eval().set_compute_dtype(torch.bfloat16)
net.load_state_dict(O.synth_state_dict({}, seed=0), strict=True)
xs = [O.synth_image(B, H, H, seed=100 + i).cuda() for i in range(3)]


def timeit(fn, n=20):
    for i in range(5):
        fn(xs[i % 3])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n):
        fn(xs[i % 3])
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


class Split:
    def __init__(self, net, example, S):
        self.x = example.clone()
        self.S = S
        self.streams = [torch.cuda.Stream() for _ in range(S)]
        with torch.no_grad():
            for _ in range(2):
                self._run()
            torch.cuda.synchronize()
            self.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph):
                self.y = self._run()

    def _run(self):
        main = torch.cuda.current_stream()
        y = torch.empty_like(self.x)
        n = self.x.shape[0] // self.S
        for i, s in enumerate(self.streams):
            s.wait_stream(main)
            with torch.cuda.stream(s):
                yi = net(self.x[i * n:(i + 1) * n])
                ops.copy_cast(yi, y[i * n:(i + 1) * n])
        for s in self.streams:
            main.wait_stream(s)
        return y

    def __call__(self, x):
        self.x.copy_(x)
        self.graph.replay()
        return self.y


with torch.no_grad():
    ref = net(xs[0]).clone()
    g1 = b200ir.GraphedForward(net, xs[0])
    print("one chain      : %.2f ms" % timeit(g1))
    for S in [int(a) for a in sys.argv[1:]] or [2, 4]:
        sp = Split(net, xs[0], S)
        t = timeit(sp)
        d = (sp(xs[0]).float() - ref.float()).abs().max().item()
        print("%d sub-batches  : %.2f ms   max |diff| vs one chain %.3g" % (S, t, d))
