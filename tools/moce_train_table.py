"""Per-kernel CUDA-event table of one eager MoCE-IR training step (B8, 128x128, bf16).  python tools/moce_train_table.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import b200ir
from b200ir import ops
from oracle.gen_golden_cases import MOCEIR_CFG3, synth_like
net = b200ir.MoCEIR(**MOCEIR_CFG3)
net.load_state_dict(synth_like(net.state_dict(), 81), strict=True)
net = net.cuda().train().set_compute_dtype(torch.bfloat16)
opt = b200ir.AdamW(net.parameters(), lr=2e-4)
c = torch.rand(8, 3, 128, 128, device="cuda")
d = (c + 0.1 * torch.randn_like(c)).clamp(0, 1)
for _ in range(3):
    b200ir.train_step(net, opt, d, c, balance_loss_weight=0.01)
torch.cuda.synchronize()
ops.timing_start()
b200ir.train_step(net, opt, d, c, balance_loss_weight=0.01)
torch.cuda.synchronize()
t = ops.timing_stop()
agg = {}
for k, (n, m) in t.items():
    e = agg.setdefault(k[0], [0, 0.0]); e[0] += n; e[1] += n * m
tot = sum(v[1] for v in agg.values())
print("sum of libb200ir kernels %.2f ms in %d launches" % (tot, sum(v[0] for v in agg.values())))
for k, (n, s) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:16]:
    print(f"  {k:28s} {n:5d} launches {s:9.3f} ms")
big = sorted(t.items(), key=lambda kv: -kv[1][0] * kv[1][1])[:12]
for k, (n, m) in big:
    print(f"  {n:3d} x {m * 1e3:8.1f} us  {k}")
