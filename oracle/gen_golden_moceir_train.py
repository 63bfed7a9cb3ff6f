"""Generate tests/golden/moceir_train_*.npz from the UNMODIFIED reference ``moce_ir.MoCEIR`` in TRAINING mode (build
container only): SparseDispatcher routing (moce_ir.py:71-141, 659-667), importance + load auxiliary losses (:760-800),
``total_loss`` (:1205-1230) and the training objective of MoCE-IR-main/src/train.py:59-72,
``L1(restored, clean) + balance_loss_weight * net.total_loss`` with balance_loss_weight = 0.01
(MoCE-IR-main/src/options.py).  Seeded synthetic weights / inputs, ``torch.manual_seed`` immediately before the forward
(router noise, moce_ir.py:741).  Stored: output, total_loss, loss, the experts picked and EVERY parameter gradient (per-tensor scale + fp16 mantissas).
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")
fv, fvnn = types.ModuleType("fvcore"), types.ModuleType("fvcore.nn")
fvnn.FlopCountAnalysis, fvnn.flop_count_table = object, (lambda *a, **k: "")
fv.nn = fvnn
sys.modules["fvcore"], sys.modules["fvcore.nn"] = fv, fvnn

from oracle import restormer_oracle as O  # noqa: E402
from oracle.gen_golden_cases import MOCEIR_TRAIN_CASES, synth_like  # noqa: E402

import moce_ir as MR  # noqa: E402  (the reference)

GOLD = os.path.join(ROOT, "tests", "golden")
BALANCE_W = 0.01


def main():
    for name, cfg, b, h, w, wseed, iseed, fseed in MOCEIR_TRAIN_CASES:
        torch.manual_seed(0)
        net = MR.MoCEIR(**cfg).train()
        sd = synth_like(net.state_dict(), wseed)
        net.load_state_dict(sd, strict=True)
        picks = []
        for m in net.modules():
            if isinstance(m, MR.RoutingFunction):
                m.register_forward_hook(lambda mod, inp, out: picks.append(out[1].detach().clone()))
        clean = O.synth_image(b, h, w, seed=iseed)
        noise = torch.randn(clean.shape, generator=torch.Generator().manual_seed(iseed + 1)) * (25.0 / 255.0)
        x = (clean + noise).clamp(0, 1)
        torch.manual_seed(fseed)
        y = net(x)
        bal = net.total_loss
        loss = torch.nn.functional.l1_loss(y, clean) + BALANCE_W * bal
        loss.backward()
        rec = {"y": y.detach().numpy(), "total_loss": np.array(float(bal)), "loss": np.array(float(loss)),
               "picks": torch.stack([p.reshape(-1) for p in picks]).numpy(), "balance_w": np.array(BALANCE_W)}
        ng = 0
        for k, p in net.named_parameters():
            if p.grad is not None:
                g = p.grad.numpy()
                sc = float(np.abs(g).max()) or 1.0
                rec["g." + k] = (g / sc).astype(np.float16)      # gradients scaled to [-1, 1] per tensor, fp16 (fixture size)
                rec["s." + k] = np.array(sc, dtype=np.float64)
                ng += 1
        np.savez_compressed(os.path.join(GOLD, name + ".npz"), **rec)
        print(name, tuple(y.shape), "loss", float(loss), "total_loss", float(bal), "picks", rec["picks"].tolist(), "grads", ng,
              "of", sum(1 for _ in net.parameters()))


if __name__ == "__main__":
    main()
