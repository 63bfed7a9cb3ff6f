"""cProfile of one eager MoCE-IR training step (where does the host time per launch go).  python tools/profile_moce_train.py"""
import cProfile, os, pstats, sys
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
pr = cProfile.Profile()
pr.enable()
for _ in range(3):
    b200ir.train_step(net, opt, d, c, balance_loss_weight=0.01)
torch.cuda.synchronize()
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(28)
