"""bf16 training-step gradients vs the fp32 oracle: relative error per parameter tensor (worst first)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import b200ir
from oracle import restormer_oracle as O
NET_CFG = dict(dim=16, num_blocks=[1, 2, 1, 1], num_refinement_blocks=1, heads=[1, 2, 4, 8])   # tests/test_backward_gpu.py
cfg = dict(O.DEFAULT_CFG); cfg.update(NET_CFG)
sd = O.synth_state_dict(NET_CFG, seed=5)
clean = O.synth_image(2, 32, 32, seed=21)
noise = torch.randn(clean.shape, generator=torch.Generator().manual_seed(22)) * (25.0 / 255.0)
degraded = (clean + noise).clamp(0, 1)
losses, grads, new_sd = O.train_step_reference(sd, NET_CFG, degraded, clean, steps=3)
net = b200ir.Restormer(**cfg).cuda(); net.load_state_dict(sd, strict=True); net.set_compute_dtype(torch.bfloat16)
opt = b200ir.AdamW(net.parameters(), lr=2e-4)
got = [b200ir.train_step(net, opt, degraded.cuda(), clean.cuda()).item() for _ in range(3)]
params = dict(net.named_parameters())
rel = lambda a, b: ((a.detach().cpu().double() - b.double()).norm() / b.double().norm().clamp_min(1e-30)).item()
rs = sorted(((rel(params[n].grad, g), n) for n, g in grads.items()), reverse=True)
print("losses", got, losses)
print("bad(>0.1):", sum(r > 0.1 for r, _ in rs), "of", len(rs))
for r, n in rs[:14]:
    print(f"  {r:.3f} {n}")
