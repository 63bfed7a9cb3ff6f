"""Generate tests/golden/gated_mdta_*.npz from the UNMODIFIED reference experiment module (build container only).

Pins the gated MDTA variant: ``GatedMDTA`` of
/root/reference/experiment/baseline_restormer_new_train/20251116_use_headwise_gate_model/MoCE-IR-main/src/net/moce_ir.py
(:84-171) -- outputs and gradients (inputs + every parameter) of the reference's own autograd.  ``fvcore`` (FLOP
counting only) is stubbed.  Only seeds and OUTPUTS are stored.
"""
import importlib.util
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
fv, fvnn = types.ModuleType("fvcore"), types.ModuleType("fvcore.nn")
fvnn.FlopCountAnalysis, fvnn.flop_count_table = object, (lambda *a, **k: "")
fv.nn = fvnn
sys.modules["fvcore"], sys.modules["fvcore.nn"] = fv, fvnn

from oracle import restormer_oracle as O  # noqa: E402
from oracle.gen_golden_cases import GATED_CASES  # noqa: E402

REF = ("/root/reference/experiment/baseline_restormer_new_train/20251116_use_headwise_gate_model/MoCE-IR-main/src/net/"
       "moce_ir.py")
spec = importlib.util.spec_from_file_location("gated_ref", REF)
GR = importlib.util.module_from_spec(spec)
spec.loader.exec_module(GR)

GOLD = os.path.join(ROOT, "tests", "golden")


def main():
    for name, dim, heads, bias, gate_type, b, h, w, seed in GATED_CASES:
        mod = GR.GatedMDTA(dim, heads, bias, gate_type=gate_type)
        sd = O.synth_gated_attention_params(dim, heads, bias, gate_type, seed)
        assert set(sd) == set(mod.state_dict()), sorted(set(sd) ^ set(mod.state_dict()))
        mod.load_state_dict(sd, strict=True)
        x = (O.synth_image(b, h, w, seed=seed + 100, c=dim) * 2 - 1).requires_grad_(True)
        gy = torch.randn(b, dim, h, w, generator=torch.Generator().manual_seed(seed + 200))
        y = mod(x)
        y.backward(gy)
        rec = {"y": y.detach().numpy(), "dx": x.grad.numpy()}
        for n, q in mod.named_parameters():
            rec["d_" + n] = q.grad.numpy()
        np.savez_compressed(os.path.join(GOLD, name + ".npz"), **rec)
        print(name, tuple(y.shape), len(rec))


if __name__ == "__main__":
    main()
