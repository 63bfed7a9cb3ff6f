"""A/B of the side-stream weight gradients: eager / graphed x side / single stream must agree bitwise."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import b200ir
from oracle import restormer_oracle as O
NET_CFG = dict(dim=16, num_blocks=[1, 2, 1, 1], num_refinement_blocks=1, heads=[1, 2, 4, 8])
cfg = dict(O.DEFAULT_CFG); cfg.update(NET_CFG)
sd = O.synth_state_dict(NET_CFG, seed=5)
clean = O.synth_image(2, 32, 32, seed=21).cuda()
deg = (clean + 0.1 * torch.randn(clean.shape, generator=torch.Generator().manual_seed(3)).cuda()).clamp(0, 1)
def make():
    net = b200ir.Restormer(**cfg).cuda(); net.load_state_dict(sd, strict=True); net.set_compute_dtype(torch.bfloat16)
    return net, b200ir.AdamW(net.parameters(), lr=2e-4)
from b200ir import autograd_ops as A
res = {}
for side in (False, "lnfold", "gatefold", True):
    for graph in (False, True):
        if side is not True and side is not False and not graph:
            continue
        b200ir.training.SIDE_WGRAD = bool(side)
        A.SIDE_SITES = {"conv1x1", "dw3x3", "conv3x3", "lnfold", "gatefold"} if side in (True, False) else {side}
        net, opt = make()
        step = b200ir.GraphedTrainStep(net, opt, deg, clean) if graph else (lambda d, c: b200ir.train_step(net, opt, d, c))
        ls = [step(deg, clean).item() for _ in range(3)]
        torch.cuda.synchronize()
        res[(side, graph)] = (ls, opt.flat_p.clone(), opt.flat_g.clone(), dict(net.named_parameters()), opt)
        print("side", side, "graph", graph, ls)
ref = res[(False, False)]
for k, v in res.items():
    print(k, "losses equal", v[0] == ref[0], "params equal", torch.equal(v[1], ref[1]), "grads equal", torch.equal(v[2], ref[2]))
    if not torch.equal(v[2], ref[2]):
        # which parameters differ
        off = 0
        bad = []
        for (n, p) in v[3].items():
            a = v[2][off:off + p.numel()]; b = ref[2][off:off + p.numel()]
            if not torch.equal(a, b): bad.append((n, float((a - b).abs().max()), float(b.abs().max())))
            off += p.numel()
        print("   differing:", bad[:12], len(bad))
