"""Does the GDFN intermediate stay in L2 if project_in and the dwconv gate run per batch CHUNK over one reused buffer?
Times K4 (LN+project_in GEMM) + K5 (dwconv+GELU gate) for the whole batch vs chunked over the batch."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import b200ir
from b200ir import ops
L = b200ir._lib


def run(B, C, H, W, chunks):
    Hd = int(C * 2.66)
    g = torch.Generator(device="cuda").manual_seed(0)
    x = (torch.rand(B, C, H, W, device="cuda", generator=g) - 0.5).bfloat16()
    w4 = (torch.rand(2 * Hd, C, device="cuda", generator=g) - 0.5) / C ** 0.5
    lw = torch.ones(C, device="cuda"); lb = torch.zeros(C, device="cuda")
    wd = (torch.rand(2 * Hd, 1, 3, 3, device="cuda", generator=g) - 0.5)
    packed = ops.pack_conv1x1(w4, None, L.LN_WITHBIAS, lw, lb)
    out = torch.empty(B, Hd, H, W, device="cuda", dtype=torch.bfloat16)
    flush = torch.empty(256 << 20, device="cuda", dtype=torch.uint8)
    res = {}
    for nc in chunks:
        cb = B // nc
        t = torch.empty(cb, 2 * Hd, H, W, device="cuda", dtype=torch.bfloat16)
        def step():
            for i in range(nc):
                ops.gemm_tc(x[i * cb:(i + 1) * cb], packed, 2 * Hd, L.LN_WITHBIAS, out=t)
                ops.dwconv3x3_gelu_gate(t, wd, None, out=out[i * cb:(i + 1) * cb])
        for _ in range(3):
            step()
        ts = []
        for _ in range(5):
            flush.zero_()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); step(); e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        res[nc] = (min(ts), out.float().abs().sum().item())
    return res


if __name__ == "__main__":
    for (B, C, H, W) in [(16, 96, 256, 256), (16, 96, 128, 128), (16, 192, 64, 64), (16, 48, 256, 256)]:
        r = run(B, C, H, W, [1, 2, 4, 8, 16])
        print((B, C, H, W), {k: f"{v[0]*1e3:.0f}us" for k, v in r.items()}, "t bytes/chunk MB:",
              {k: round(B // k * 2 * int(C * 2.66) * H * W * 2 / 1e6, 1) for k in r})
        cs = {round(v[1], 0) for v in r.values()}
        assert len(cs) == 1, cs
