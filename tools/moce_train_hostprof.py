"""cProfile of the host side of eager MoCE-IR training steps (B8, 128x128, bf16): where the ~48 us per launch go.
python tools/moce_train_hostprof.py [steps]"""
import cProfile, os, pstats, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import b200ir
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
n = int(sys.argv[1]) if len(sys.argv) > 1 else 3
t0 = time.perf_counter()
for _ in range(n):
    b200ir.train_step(net, opt, d, c, balance_loss_weight=0.01)
torch.cuda.synchronize()
print("unprofiled: %.1f ms / step" % ((time.perf_counter() - t0) / n * 1e3))
pr = cProfile.Profile()
pr.enable()
for _ in range(n):
    b200ir.train_step(net, opt, d, c, balance_loss_weight=0.01)
torch.cuda.synchronize()
pr.disable()
st = pstats.Stats(pr)
st.sort_stats("tottime").print_stats(45)
