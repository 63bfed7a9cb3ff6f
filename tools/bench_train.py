"""Training-step timing (cfg4: Restormer, L1 + AdamW, 128x128 patches, batch 8 per GPU) with a per-kernel breakdown.
usage: python tools/bench_train.py [--batch 8] [--size 128] [--dtype bf16|fp32] [--steps 5] [--breakdown]"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200ir  # noqa: E402
from b200ir import ops  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--size", type=int, default=128)
    ap.add_argument("--dtype", default="bf16")
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--breakdown", action="store_true")
    ap.add_argument("--graph", action="store_true")
    ap.add_argument("--no-side", action="store_true", help="single-stream backward (A/B of the side-stream weight gradients)")
    ap.add_argument("--no-fork", action="store_true", help="attention backward on one stream (A/B)")
    a = ap.parse_args()
    if a.no_side:
        b200ir.training.SIDE_WGRAD = False
    if a.no_fork:
        b200ir.autograd_ops.FORK_ATTN_BWD = False
    torch.manual_seed(0)
    net = b200ir.Restormer().cuda()
    net.set_compute_dtype(torch.bfloat16 if a.dtype == "bf16" else torch.float32)
    opt = b200ir.AdamW(net.parameters(), lr=2e-4)
    g = torch.Generator().manual_seed(1234)
    clean = torch.rand(a.batch, 3, a.size, a.size, generator=g)
    deg = (clean + 25 / 255 * torch.randn(clean.shape, generator=g)).clamp(0, 1).cuda()
    clean = clean.cuda()
    step = b200ir.GraphedTrainStep(net, opt, deg, clean) if a.graph else (lambda d, c: b200ir.train_step(net, opt, d, c))
    for _ in range(2):
        loss = step(deg, clean)
    torch.cuda.synchronize()
    l0 = ops.LAUNCHES
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        loss = step(deg, clean)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.steps
    mp = a.batch * a.size * a.size / 1e6
    print(f"train step: {ms:.2f} ms  {mp / ms * 1e3:.2f} MP/s  loss {loss.item():.5f}  launches/step "
          f"{(ops.LAUNCHES - l0) / a.steps:.0f}  peak mem {torch.cuda.max_memory_allocated() / 2**30:.2f} GiB")
    if a.breakdown:
        ops.timing_start()
        b200ir.train_step(net, opt, deg, clean)
        torch.cuda.synchronize()
        t = ops.timing_stop()
        agg = {}
        for k, (n, mean) in t.items():
            e = agg.setdefault(k[0], [0, 0.0])
            e[0] += n
            e[1] += n * mean
        tot = sum(v[1] for v in agg.values())
        for k, (n, s) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            print(f"  {k:28s} {n:5d} launches {s:9.3f} ms  {100 * s / tot:5.1f}%")
        print(f"  sum of kernels {tot:.2f} ms")
        big = sorted(t.items(), key=lambda kv: -kv[1][0] * kv[1][1])[:25]
        for k, (n, mean) in big:
            print(f"    {str(k):80s} x{n:3d} {mean * 1e3:9.1f} us")


if __name__ == "__main__":
    main()
