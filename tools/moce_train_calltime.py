"""Host time spent INSIDE each libb200ir entry point during one eager MoCE-IR training step (a launch that blocks shows up
here): python tools/moce_train_calltime.py"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import b200ir
from b200ir import ops
from oracle.gen_golden_cases import MOCEIR_CFG3, synth_like


class Proxy:
    def __init__(self, lib):
        self._lib, self.t, self.slow = lib, {}, []

    def __getattr__(self, name):
        f = getattr(self._lib, name)
        if not name.startswith("b200ir_"):
            return f

        def g(*a):
            t0 = time.perf_counter()
            r = f(*a)
            dt = time.perf_counter() - t0
            e = self.t.setdefault(name, [0, 0.0])
            e[0] += 1; e[1] += dt
            if dt > 2e-4:
                self.slow.append((dt, name, time.perf_counter() - self.t0))
            return r
        return g


net = b200ir.MoCEIR(**MOCEIR_CFG3)
net.load_state_dict(synth_like(net.state_dict(), 81), strict=True)
net = net.cuda().train().set_compute_dtype(torch.bfloat16)
opt = b200ir.AdamW(net.parameters(), lr=2e-4)
c = torch.rand(8, 3, 128, 128, device="cuda")
d = (c + 0.1 * torch.randn_like(c)).clamp(0, 1)
for _ in range(3):
    b200ir.train_step(net, opt, d, c, balance_loss_weight=0.01)
torch.cuda.synchronize()
p = Proxy(ops.lib)
ops.lib = p
p.t0 = time.perf_counter()
b200ir.train_step(net, opt, d, c, balance_loss_weight=0.01)
t1 = time.perf_counter() - p.t0
torch.cuda.synchronize()
t2 = time.perf_counter() - p.t0
print("host returned after %.1f ms, GPU done after %.1f ms" % (t1 * 1e3, t2 * 1e3))
tot = sum(v[1] for v in p.t.values())
print("inside libb200ir calls: %.1f ms in %d calls" % (tot * 1e3, sum(v[0] for v in p.t.values())))
for k, (n, s) in sorted(p.t.items(), key=lambda kv: -kv[1][1])[:12]:
    print("  %-40s %5d calls %8.2f ms  (%.1f us each)" % (k, n, s * 1e3, s / n * 1e6))
print("calls above 200 us:")
for dt, name, at in p.slow[:40]:
    print("  at %7.1f ms  %-36s %8.1f us" % (at * 1e3, name, dt * 1e6))
