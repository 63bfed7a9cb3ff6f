"""Per-kernel micro-benchmarks at the BASELINE configs[1] shapes (B=16, 256x256): CUDA-event time, algorithmic
bytes and fraction of the measured HBM peak.  Usage: python tools/bench_ops.py [names...] [--iters N] [--once]"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import b200ir  # noqa: E402
from b200ir import ops  # noqa: E402
import importlib  # noqa: E402

A = importlib.import_module("b200ir.autograd_ops")
L = b200ir._lib
PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(
    os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
dev = "cuda"
bf = torch.bfloat16


def rnd(*shape, dtype=bf):
    return (torch.rand(*shape, device=dev) - 0.5).to(dtype)


def level(name):
    # (B, C, heads, H, W)
    return {"l1a": (16, 48, 1, 256, 256), "l1b": (16, 96, 1, 256, 256), "l2": (16, 96, 2, 128, 128),
            "l3": (16, 192, 4, 64, 64), "l4": (16, 384, 8, 32, 32),
            # training shapes of cfg4 (B8, 128x128 patches): the four U-Net levels
            "t1": (8, 48, 1, 128, 128), "t2": (8, 96, 2, 64, 64), "t3": (8, 192, 4, 32, 32), "t4": (8, 384, 8, 16, 16)}[name]


VARIANT = int(__import__('os').environ.get('BENCH_VARIANT', '0'))   # L.GEMM_AUTO / RESIDENT / STREAM / RESIDENT_CH (A/B timing)


def make(opname, lv):
    B, C, h, H, W = level(lv)
    Hd = int(C * 2.66)
    px = B * H * W
    if opname == "k1":
        x = [rnd(B, C, H, W) for _ in range(3)]
        p = ops.pack_conv1x1(rnd(3 * C, C, dtype=torch.float32), None, L.LN_WITHBIAS, torch.ones(C, device=dev), torch.zeros(C, device=dev))
        out = torch.empty(B, 3 * C, H, W, device=dev, dtype=bf)
        return (lambda i: ops.gemm_tc(x[i % 3], p, 3 * C, L.LN_WITHBIAS, out=out, variant=VARIANT)), px * (C + 3 * C) * 2
    if opname == "k4":
        x = [rnd(B, C, H, W) for _ in range(3)]
        p = ops.pack_conv1x1(rnd(2 * Hd, C, dtype=torch.float32), None, L.LN_WITHBIAS, torch.ones(C, device=dev), torch.zeros(C, device=dev))
        out = torch.empty(B, 2 * Hd, H, W, device=dev, dtype=bf)
        return (lambda i: ops.gemm_tc(x[i % 3], p, 2 * Hd, L.LN_WITHBIAS, out=out, variant=VARIANT)), px * (C + 2 * Hd) * 2
    if opname == "k6":
        x = [rnd(B, Hd, H, W) for _ in range(2)]
        r = [rnd(B, C, H, W) for _ in range(2)]
        p = ops.pack_conv1x1(rnd(C, Hd, dtype=torch.float32))
        out = torch.empty(B, C, H, W, device=dev, dtype=bf)
        return (lambda i: ops.gemm_tc(x[i % 2], p, C, L.LN_NONE, r[i % 2], out=out, variant=VARIANT)), px * (Hd + 2 * C) * 2
    if opname == "k3b":
        qkv = [rnd(B, 3 * C, H, W) for _ in range(2)]
        r = [rnd(B, C, H, W) for _ in range(2)]
        p = ops.pack_conv1x1(rnd(B, C, C, dtype=torch.float32), None, per_sample=True)
        out = torch.empty(B, C, H, W, device=dev, dtype=bf)
        return (lambda i: ops.gemm_tc(qkv[i % 2][:, 2 * C:], p, C, L.LN_NONE, r[i % 2], out=out, variant=VARIANT)), px * 3 * C * 2
    if opname == "k6n":     # K6 without the residual (A/B: what the epilogue's residual path costs)
        x = [rnd(B, Hd, H, W) for _ in range(2)]
        p = ops.pack_conv1x1(rnd(C, Hd, dtype=torch.float32))
        out = torch.empty(B, C, H, W, device=dev, dtype=bf)
        return (lambda i: ops.gemm_tc(x[i % 2], p, C, L.LN_NONE, None, out=out, variant=VARIANT)), px * (Hd + C) * 2
    if opname == "k3bn":
        qkv = [rnd(B, 3 * C, H, W) for _ in range(2)]
        p = ops.pack_conv1x1(rnd(B, C, C, dtype=torch.float32), None, per_sample=True)
        out = torch.empty(B, C, H, W, device=dev, dtype=bf)
        return (lambda i: ops.gemm_tc(qkv[i % 2][:, 2 * C:], p, C, L.LN_NONE, None, out=out, variant=VARIANT)), px * 2 * C * 2
    if opname == "dw":
        x = [rnd(B, 3 * C, H, W) for _ in range(2)]
        w = rnd(3 * C, 1, 3, 3, dtype=torch.float32)
        out = torch.empty(B, 3 * C, H, W, device=dev, dtype=bf)
        return (lambda i: ops.dwconv3x3(x[i % 2], w, None, out=out)), px * 6 * C * 2
    if opname == "gate":
        x = [rnd(B, 2 * Hd, H, W) for _ in range(2)]
        w = rnd(2 * Hd, 1, 3, 3, dtype=torch.float32)
        out = torch.empty(B, Hd, H, W, device=dev, dtype=bf)
        return (lambda i: ops.dwconv3x3_gelu_gate(x[i % 2], w, None, out=out)), px * 3 * Hd * 2
    if opname == "gram":
        x = [rnd(B, 3 * C, H, W) for _ in range(2)]
        return (lambda i: ops.mdta_gram(x[i % 2], h)), px * 2 * C * 2
    if opname == "fold":
        x = rnd(B, 3 * C, H, W)
        part = ops.mdta_gram(x, h)
        t = torch.ones(h, device=dev)
        wo = rnd(C, C, dtype=torch.float32)
        return (lambda i: ops.mdta_softmax_fold(part, t, wo)), part.numel() * 4 + B * C * C * 4
    raise KeyError(opname)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("names", nargs="*", default=["k1", "k3b", "k4", "k6", "dw", "gate", "gram", "fold"])
    ap.add_argument("--levels", default="l1b")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--once", action="store_true", help="single launch per op (for ncu)")
    a = ap.parse_args()
    for lv in a.levels.split(","):
        for n in a.names:
            fn, nbytes = make(n, lv)
            if a.once:
                fn(0)
                torch.cuda.synchronize()
                continue
            for i in range(3):
                fn(i)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(a.iters):
                fn(i)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / a.iters
            gbs = nbytes / ms / 1e6
            print(f"{lv} {n:5s} {ms*1e3:9.1f} us  {nbytes/1e6:9.1f} MB  {gbs:8.0f} GB/s  {gbs/PEAK*100:5.1f}% of measured HBM")
            del fn
            torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
