"""Generate tests/golden/moceir_*.npz from the UNMODIFIED reference ``moce_ir.MoCEIR`` (build container only).

The whole all-in-one model of BASELINE configs[2] at small sizes: seeded synthetic weights (``synth_like`` over the
reference's own state_dict keys / shapes), seeded input, ``torch.manual_seed`` immediately before the eval-mode forward
(the router adds noise in eval mode too, moce_ir.py:741).  Stored: the output and, per DecoderBlock, the
expert indices the reference picked (so a mismatch can be told apart from a numerical one).  ``fvcore`` is stubbed.
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
from oracle.gen_golden_cases import MOCEIR_CASES, synth_like  # noqa: E402

import moce_ir as MR  # noqa: E402  (the reference)

GOLD = os.path.join(ROOT, "tests", "golden")


def main():
    for name, cfg, b, h, w, wseed, iseed, fseed in MOCEIR_CASES:
        torch.manual_seed(0)
        net = MR.MoCEIR(**cfg).eval()
        sd = synth_like(net.state_dict(), wseed)
        net.load_state_dict(sd, strict=True)
        picks = []
        for m in net.modules():
            if isinstance(m, MR.RoutingFunction):
                m.register_forward_hook(lambda mod, inp, out: picks.append(out[1].detach().clone()))
        x = O.synth_image(b, h, w, seed=iseed)
        torch.manual_seed(fseed)
        with torch.no_grad():
            y = net(x)
        rec = {"y": y.numpy(), "picks": torch.stack([p.reshape(-1) for p in picks]).numpy(),
               "nparams": np.array(sum(v.numel() for v in sd.values())),
               "keys": np.array(sorted(sd)), "shapes": np.array([str(tuple(sd[k].shape)) for k in sorted(sd)]),
               "complexity": np.stack([sd[k].numpy() for k in sorted(sd) if k.endswith("complexity")])}
        np.savez_compressed(os.path.join(GOLD, name + ".npz"), **rec)
        print(name, tuple(y.shape), "experts picked per decoder block:", rec["picks"].tolist(), "state tensors", len(sd))


if __name__ == "__main__":
    main()
