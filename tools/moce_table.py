"""Per-kernel CUDA-event table of one MoCEIR forward (BASELINE configs[2]).  python tools/moce_table.py [B] [HW]"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import b200ir
from b200ir import ops
from oracle import restormer_oracle as O
from oracle.gen_golden_cases import MOCEIR_CFG3, synth_like
B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
HW = int(sys.argv[2]) if len(sys.argv) > 2 else 256
net = b200ir.MoCEIR(**MOCEIR_CFG3)
net.load_state_dict(synth_like(net.state_dict(), 81), strict=True)
net = net.cuda().eval().set_compute_dtype(torch.bfloat16)
x = O.synth_image(B, HW, HW, seed=5).cuda()
with torch.no_grad():
    for _ in range(2):
        net(x)
    torch.cuda.synchronize()
    ops.timing_start()
    net(x)
    torch.cuda.synchronize()
    t = ops.timing_stop()
rows = sorted(((n * m, n, m, k) for k, (n, m) in t.items()), reverse=True)
tot = sum(r[0] for r in rows)
print("total ms (sum of kernels)", tot)
for tm, n, m, k in rows[:40]:
    print(f"{tm:8.3f} ms  {n:3d} x {m * 1e3:8.1f} us  {k}")
