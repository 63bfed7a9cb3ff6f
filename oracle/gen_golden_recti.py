"""Generate tests/golden/recti_*.npz and adair_cca_*.npz from the UNMODIFIED reference modules (build container only):
``experiment/RectiFormer.py`` GatedSpatialFFN (:231-266) and ElementwiseGatedAttention (:270-325, with and without the
per-sample dynamic temperature) and ``AdaIR-main/net/model.py`` Chanel_Cross_Attention (:177-216).  SURVEY.md §8 f2 / f3."""
import importlib.util
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
from oracle import restormer_oracle as O  # noqa: E402
from oracle.gen_golden_cases import RECTI_CASES, ADAIR_CCA_CASES, FREMODULE_CASES, fre_sample, synth_like  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def load(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def main():
    RF = load("/root/reference/experiment/RectiFormer.py", "ref_rectiformer")
    for name, kind, dim, heads, f, bias, b, h, w, seed in RECTI_CASES:
        x = O.synth_image(b, h, w, seed=seed + 100, c=dim) * 2 - 1
        if kind == "ffn":
            mod = RF.GatedSpatialFFN(dim, f, bias).eval()
        else:
            mod = RF.ElementwiseGatedAttention(dim, heads, bias).eval()
        mod.load_state_dict(synth_like(mod.state_dict(), seed), strict=True)
        with torch.no_grad():
            if kind == "attn_dyn":
                dt = (torch.rand(b, generator=torch.Generator().manual_seed(seed + 7)) * 2).view(b, 1, 1, 1)
                y = mod(x, dt)
            else:
                y = mod(x)
        # the same module under autograd (round 2: the variants train): input, parameter and dynamic-temperature gradients
        rec = {}
        xg = x.clone().requires_grad_(True)
        gy = torch.randn(y.shape, generator=torch.Generator().manual_seed(seed + 9))
        if kind == "attn_dyn":
            dtg = dt.clone().requires_grad_(True)
            mod(xg, dtg).backward(gy)
            rec["ddt"] = dtg.grad.numpy()
        else:
            mod(xg).backward(gy)
        rec["dx"] = xg.grad.numpy()
        for k, q in mod.named_parameters():
            rec["g." + k] = q.grad.numpy()
        np.savez_compressed(os.path.join(GOLD, name + ".npz"), y=y.numpy(), keys=np.array(sorted(mod.state_dict())), **rec)
        print(name, tuple(y.shape), "grads", len(rec))
    AD = load("/root/reference/AdaIR-main/net/model.py", "ref_adair_model")
    for name, dim, heads, bias, b, h, w, seed in ADAIR_CCA_CASES:
        mod = AD.Chanel_Cross_Attention(dim, heads, bias).eval()
        mod.load_state_dict(synth_like(mod.state_dict(), seed), strict=True)
        x = O.synth_image(b, h, w, seed=seed + 100, c=dim) * 2 - 1
        y2 = O.synth_image(b, h, w, seed=seed + 300, c=dim) * 2 - 1
        with torch.no_grad():
            y = mod(x, y2)
        np.savez_compressed(os.path.join(GOLD, name + ".npz"), y=y.numpy(), keys=np.array(sorted(mod.state_dict())))
        print(name, tuple(y.shape))
    for name, dim, heads, bias, b, (hi, wi), (h, w), seed in FREMODULE_CASES:
        mod = AD.FreModule(dim, heads, bias).eval()
        sd = synth_like(mod.state_dict(), seed)
        sd["para1"] = 0.5 + torch.rand(dim, 1, 1, generator=torch.Generator().manual_seed(seed + 1))   # (init is zeros)
        mod.load_state_dict(sd, strict=True)
        img = O.synth_image(b, hi, wi, seed=seed + 100)
        y2 = O.synth_image(b, h, w, seed=seed + 300, c=dim) * 2 - 1
        with torch.no_grad():
            y = mod(img, y2)
            hf, lf = mod.fft(torch.nn.functional.interpolate(img, (h, w), mode="bilinear"))
        ys, yc = fre_sample(y)
        hs, hc = fre_sample(hf)
        ls, lc = fre_sample(lf)
        np.savez_compressed(os.path.join(GOLD, name + ".npz"), y=ys, y_sum=yc, high=hs, high_sum=hc, low=ls, low_sum=lc,
                            keys=np.array(sorted(mod.state_dict())))
        print(name, tuple(y.shape), "low band max", lf.abs().max().item())


if __name__ == "__main__":
    main()
