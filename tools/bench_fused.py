"""Time the fused LN + 1x1 + depthwise (+gate) kernel against the two-kernel path it replaces, per U-Net level of cfg2
(B16, 256x256 input).  CUDA events, inputs rotate over > L2-sized sets.   python tools/bench_fused.py [gate|plain|all]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import b200ir  # noqa: E402
from b200ir import ops  # noqa: E402

L = b200ir._lib


def timeit(fn, n=20, warm=3):
    for _ in range(warm):
        fn(0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3   # us


def main():
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    B = 16
    levels = [(96, 255, 256), (48, 127, 256), (96, 255, 128), (192, 510, 64)]
    if os.environ.get("FUSED_LEVELS"):
        levels = levels[:int(os.environ["FUSED_LEVELS"])]
    g = torch.Generator().manual_seed(0)
    rows = []
    for C, Hd, HW in levels:
        for gate in ([True, False] if which == "all" else [which == "gate"]):
            Cout = Hd if gate else 3 * C
            Ctot = 2 * Hd if gate else 3 * C
            nset = max(2, int(300e6 / (B * C * HW * HW * 2)) + 1)
            xs = [(torch.rand(B, C, HW, HW, generator=g) - 0.5).bfloat16().cuda() for _ in range(nset)]
            w = ((torch.rand(Ctot, C, generator=g) - 0.5) / C ** 0.5).cuda()
            lw, lb = (torch.rand(C, generator=g) + 0.5).cuda(), (torch.rand(C, generator=g) - 0.5).cuda()
            dw = ((torch.rand(Ctot, 1, 3, 3, generator=g) - 0.5) * 0.5).cuda()
            if not ops.ln_proj_dw_usable(xs[0], Cout, gate):
                continue
            wf = ops.pack_fused_proj_dw(w, None, L.LN_WITHBIAS, lw, lb, Cout, gate)
            packed = ops.pack_conv1x1(w, None, L.LN_WITHBIAS, lw, lb)
            out = torch.empty((B, Cout, HW, HW), device="cuda", dtype=torch.bfloat16)
            tbuf = torch.empty((B, Ctot, HW, HW), device="cuda", dtype=torch.bfloat16)
            t_f = timeit(lambda i: ops.ln_proj_dw(xs[i % nset], wf, dw, None, L.LN_WITHBIAS, Cout, gate, out=out))
            t_g = timeit(lambda i: ops.gemm_tc(xs[i % nset], packed, Ctot, L.LN_WITHBIAS, out=tbuf))
            if gate:
                t_d = timeit(lambda i: ops.dwconv3x3_gelu_gate(tbuf, dw, None, out=out))
            else:
                t_d = timeit(lambda i: ops.dwconv3x3(tbuf, dw, None, out=out))
            px = B * HW * HW
            alg = (C + Cout) * 2 * px
            rows.append(dict(C=C, Cout=Cout, HW=HW, gate=gate, fused_us=t_f, gemm_us=t_g, dw_us=t_d,
                             speedup=(t_g + t_d) / t_f, fused_gbs=alg / t_f / 1e3))
            print(json.dumps(rows[-1]))
    return rows


if __name__ == "__main__":
    main()
