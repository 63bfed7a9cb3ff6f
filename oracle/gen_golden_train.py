"""Generate tests/golden/train_*.npz from the UNMODIFIED reference (build container only; needs /root/reference).

Pins the backward / training-step oracle (``block_grads``, ``train_step_reference`` in restormer_oracle.py) against the
reference itself: ``Restormer.TransformerBlock`` / ``Restormer.Restormer`` are imported by path, loaded with the seeded
synthetic weights, and differentiated by torch autograd exactly as ``loss.backward()`` does inside the reference's
Lightning ``training_step`` (MoCE-IR-main/src/train.py:59-72), followed by ``optim.AdamW(lr=2e-4)`` (:83).
Only seeds and OUTPUTS (gradients, losses, updated parameters) are stored.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")

from oracle import restormer_oracle as O  # noqa: E402
from oracle.gen_golden_cases import TRAIN_BLOCK_CASES, TRAIN_NET_CASE  # noqa: E402

import Restormer as R  # noqa: E402  (the reference)

GOLD = os.path.join(ROOT, "tests", "golden")


def main():
    for name, dim, heads, f, bias, ln, b, h, w, seed in TRAIN_BLOCK_CASES:
        p = O.synth_block_params(dim, heads, f, bias, ln, seed)
        blk = R.TransformerBlock(dim, heads, f, bias, ln)
        blk.load_state_dict(p, strict=True)
        x = (O.synth_image(b, h, w, seed=seed + 100, c=dim) * 2 - 1).requires_grad_(True)
        gy = torch.randn(b, dim, h, w, generator=torch.Generator().manual_seed(seed + 200))
        blk(x).backward(gy)
        out = {"dx": x.grad.numpy()}
        for n, q in blk.named_parameters():
            out["d_" + n] = q.grad.numpy()
        np.savez_compressed(os.path.join(GOLD, name + ".npz"), **out)
        print(name, len(out))

    name, cfg, b, h, w, wseed, iseed, steps = TRAIN_NET_CASE
    full = dict(O.DEFAULT_CFG)
    full.update(cfg)
    net = R.Restormer(**full)
    net.load_state_dict(O.synth_state_dict(cfg, seed=wseed), strict=True)
    clean = O.synth_image(b, h, w, seed=iseed)
    noise = torch.randn(clean.shape, generator=torch.Generator().manual_seed(iseed + 1)) * (25.0 / 255.0)
    degraded = (clean + noise).clamp(0, 1)
    opt = torch.optim.AdamW(net.parameters(), lr=2e-4)          # MoCE-IR-main/src/train.py:83
    loss_fn = torch.nn.L1Loss()                                 # :51
    losses = []
    for _ in range(steps):
        opt.zero_grad()
        loss = loss_fn(net(degraded), clean)                    # :62-65
        loss.backward()
        opt.step()
        losses.append(loss.item())
    out = {"losses": np.array(losses, dtype=np.float64)}
    for n, q in net.named_parameters():
        out["g_" + n] = q.grad.numpy()                          # gradients of the last step
        out["p_" + n] = q.detach().numpy()                      # parameters after the last step
    np.savez_compressed(os.path.join(GOLD, name + ".npz"), **out)
    print(name, losses)


if __name__ == "__main__":
    main()
