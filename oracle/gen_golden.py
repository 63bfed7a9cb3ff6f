"""Generate tests/golden/*.npz from the UNMODIFIED reference (run in the build container only).

    python oracle/gen_golden.py            # needs /root/reference (read-only mount)

The reference modules are imported by path (never copied), given the seeded synthetic
weights from ``oracle.restormer_oracle.synth_*`` through ``load_state_dict(strict=True)``
(which also proves that the names/shapes in ``restormer_param_shapes`` are the
reference's), and run on CPU in fp32.  Only seeds, configs and OUTPUTS are stored; inputs
and weights are regenerated from the seeds by the tests.  `/root/reference` does not exist
on the GPU box, so nothing else may import it.
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

import Restormer as R  # noqa: E402  (the reference, /root/reference/Restormer.py)

GOLD = os.path.join(ROOT, "tests", "golden")

from oracle.gen_golden_cases import BLOCK_CASES, NET_CASES  # noqa: E402


def main():
    os.makedirs(GOLD, exist_ok=True)
    torch.set_grad_enabled(False)
    torch.manual_seed(0)

    for name, dim, heads, f, bias, ln, b, h, w, seed in BLOCK_CASES:
        p = O.synth_block_params(dim, heads, f, bias, ln, seed)
        blk = R.TransformerBlock(dim, heads, f, bias, ln).eval()
        blk.load_state_dict(p, strict=True)
        x = O.synth_image(b, h, w, seed=seed + 100, c=dim) * 2 - 1
        out = {
            "y_block": blk(x),
            "y_norm1": blk.norm1(x),
            "y_attn": blk.attn(blk.norm1(x)),
            "y_ffn": blk.ffn(blk.norm2(x)),
        }
        np.savez_compressed(os.path.join(GOLD, name + ".npz"),
                            meta=np.array([dim, heads, int(bias), b, h, w, seed]), f=np.array([f]),
                            ln=np.array([ln]), **{k: v.numpy() for k, v in out.items()})
        print(name, {k: tuple(v.shape) for k, v in out.items()})

    # resizing modules + patch embed
    for name, mod, c in [("down_c48", R.Downsample(48), 48), ("up_c96", R.Upsample(96), 96),
                         ("embed_3_48", R.OverlapPatchEmbed(3, 48), 3)]:
        gen = torch.Generator().manual_seed(21)
        sd = {k: O.synth_tensor(tuple(v.shape), "conv", gen) for k, v in mod.state_dict().items()}
        mod.load_state_dict(sd, strict=True)
        x = O.synth_image(2, 8, 12, seed=22, c=c)
        np.savez_compressed(os.path.join(GOLD, name + ".npz"), y=mod.eval()(x).numpy())
        print(name)

    for name, cfg, b, h, w, wseed, iseed in NET_CASES:
        full = dict(O.DEFAULT_CFG)
        full.update(cfg)
        net = R.Restormer(**full).eval()
        sd = O.synth_state_dict(cfg, seed=wseed)
        assert list(sd.keys()) == list(net.state_dict().keys()), "state_dict order/name mismatch"
        net.load_state_dict(sd, strict=True)
        x = O.synth_image(b, h, w, seed=iseed, c=full["inp_channels"])
        y = net(x)
        np.savez_compressed(os.path.join(GOLD, name + ".npz"), y=y.numpy(),
                            meta=np.array([b, h, w, wseed, iseed]))
        print(name, tuple(y.shape), float(y.abs().mean()))


if __name__ == "__main__":
    main()
