"""Generate tests/golden/moce_encgrp_*.npz and moce_cross_*.npz from the UNMODIFIED reference moce_ir.py (build container only).

Pins the MoCE-IR flavour of the block (``EncoderBlock`` / ``EncoderResidualGroup``, moce_ir.py:805-834, 903-932: same
math as Restormer's TransformerBlock under the names ``norms.0``, ``norms.1``, ``mixer``, ``ffn``; bias=True as the
MoCE-IR encoder uses it, moce_ir.py:1149).  ``fvcore`` (imported at moce_ir.py:16 for FLOP counting only) is not
installed: a stub module is injected before the import.  Only seeds and OUTPUTS are stored.
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
from oracle.gen_golden_cases import CROSS_CASES, MOCE_CASES, moce_block_params  # noqa: E402

import moce_ir as MR  # noqa: E402  (the reference)

GOLD = os.path.join(ROOT, "tests", "golden")


def main():
    for name, dim, heads, f, bias, ln, nblk, b, h, w, seed in MOCE_CASES:
        grp = MR.EncoderResidualGroup(dim=dim, num_heads=heads, num_blocks=nblk, ffn_expansion=f, LayerNorm_type=ln,
                                      bias=bias)
        sd = moce_block_params(dim, heads, f, bias, ln, nblk, seed)
        assert set(sd) == set(grp.state_dict()), sorted(set(sd) ^ set(grp.state_dict()))
        grp.load_state_dict(sd, strict=True)
        x = (O.synth_image(b, h, w, seed=seed + 100, c=dim) * 2 - 1).requires_grad_(True)
        gy = torch.randn(b, dim, h, w, generator=torch.Generator().manual_seed(seed + 200))
        y = grp(x)
        y.backward(gy)
        out = {"y": y.detach().numpy(), "dx": x.grad.numpy()}
        for n, q in grp.named_parameters():
            out["d_" + n] = q.grad.numpy()
        np.savez_compressed(os.path.join(GOLD, name + ".npz"), **out)
        print(name, tuple(y.shape), len(out))
    # CrossAttention (decoder mixer): outputs of the reference module for seeded parameters and two seeded inputs
    for name, dim, heads, bias, b, h, w, seed in CROSS_CASES:
        mod = MR.CrossAttention(dim, heads, bias)
        sd = O.synth_cross_attention_params(dim, heads, bias, seed)
        assert set(sd) == set(mod.state_dict()), sorted(set(sd) ^ set(mod.state_dict()))
        mod.load_state_dict(sd, strict=True)
        x = (O.synth_image(b, h, w, seed=seed + 100, c=dim) * 2 - 1).requires_grad_(True)
        y = (O.synth_image(b, h, w, seed=seed + 300, c=dim) * 2 - 1).requires_grad_(True)
        gy = torch.randn(b, dim, h, w, generator=torch.Generator().manual_seed(seed + 200))
        out = mod(x, y)
        out.backward(gy)
        rec = {"y": out.detach().numpy(), "dx": x.grad.numpy(), "dy": y.grad.numpy()}
        for n, q in mod.named_parameters():
            rec["d_" + n] = q.grad.numpy()
        np.savez_compressed(os.path.join(GOLD, name + ".npz"), **rec)
        print(name, tuple(out.shape), len(rec))


if __name__ == "__main__":
    main()
