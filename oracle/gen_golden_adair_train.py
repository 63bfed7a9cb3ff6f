"""Generate tests/golden/adair_fre_train_*.npz from the UNMODIFIED reference ``AdaIR-main/net/model.py`` FreModule
(:295-372) under torch autograd on the CPU (build container only): what AdaIR-main/train.py differentiates.  Seeded
synthetic weights / inputs; loss = mean(out * r) with a seeded random r.  Stored: out (strided sample + checksums), loss,
the spectrum-mask half sizes per sample, d loss / d feature (strided sample + checksums) and every parameter gradient
(fp32; parameters the forward never touches -- conv, score_gen, rate_conv -- have none, as in the reference).
SURVEY.md §8 f3."""
import importlib.util
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
from oracle import restormer_oracle as O  # noqa: E402
from oracle.gen_golden_cases import ADAIR_NET_CASES, FREMODULE_TRAIN_CASES, fre_sample, grad_sample, synth_like  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def load(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def fre_train_inputs(dim, b, hi, wi, h, w, seed):
    img = O.synth_image(b, hi, wi, seed=seed + 100)
    y2 = O.synth_image(b, h, w, seed=seed + 300, c=dim) * 2 - 1
    r = torch.randn(b, dim, h, w, generator=torch.Generator().manual_seed(seed + 500))
    return img, y2, r


def fre_train_state(mod, dim, seed, rate_scale):
    sd = synth_like(mod.state_dict(), seed)
    sd["rate_conv.2.weight"] = sd["rate_conv.2.weight"] * rate_scale
    sd["para1"] = 0.5 + torch.rand(dim, 1, 1, generator=torch.Generator().manual_seed(seed + 1))   # (init is zeros)
    return sd


def adair_net_state(net, seed, rate_scale):
    sd = synth_like(net.state_dict(), seed)
    g = torch.Generator().manual_seed(seed + 1)
    for k in sd:
        if k.endswith("para1"):                                   # (init is zeros: the frequency branch would be switched off)
            sd[k] = 0.5 + torch.rand(sd[k].shape, generator=g)
    rs = torch.tensor(rate_scale if isinstance(rate_scale, tuple) else (rate_scale, rate_scale), dtype=torch.float32)
    sd["fre3.rate_conv.2.weight"] = sd["fre3.rate_conv.2.weight"] * rs.view(2, 1, 1, 1)
    return sd


def adair_net_inputs(b, h, w, seed):
    clean = O.synth_image(b, h, w, seed=seed + 100)
    noise = torch.randn(clean.shape, generator=torch.Generator().manual_seed(seed + 101)) * (25.0 / 255.0)
    return (clean + noise).clamp(0, 1), clean


def nets(AD):
    """The whole AdaIR network: restored image, and for the training cases L1 loss + every parameter gradient."""
    for name, cfg, b, h, w, seed, train, rs in ADAIR_NET_CASES:
        net = AD.AdaIR(**cfg).train()
        net.load_state_dict(adair_net_state(net, seed, rs), strict=True)
        x, clean = adair_net_inputs(b, h, w, seed)
        rec = {"keys": np.array(list(net.state_dict()))}
        if train:
            y = net(x)
            loss = torch.nn.functional.l1_loss(y, clean)
            loss.backward()
            rec["y"], rec["loss"] = y.detach().numpy(), np.array(float(loss.detach()))
            for k, p in net.named_parameters():
                if p.grad is not None:
                    g = grad_sample(p.grad.numpy())
                    sc = float(np.abs(g).max()) or 1.0
                    rec["g." + k] = (g / sc).astype(np.float16)
                    rec["s." + k] = np.array(sc, dtype=np.float64)
        else:
            with torch.no_grad():
                y = net(x)
                f3 = torch.nn.functional.interpolate(x, (h // 2, w // 2), mode="bilinear")
                thr = net.fre3.rate_conv(torch.nn.functional.adaptive_avg_pool2d(net.fre3.conv1(f3), 1)).sigmoid()[:, :, 0, 0]
            rec["y"], rec["y_sum"] = fre_sample(y)
            rec["thr3"] = thr.numpy()
        np.savez_compressed(os.path.join(GOLD, name + ".npz"), **rec)
        print(name, tuple(y.shape), "loss" if train else "fre3 thr", float(loss.detach()) if train else thr.tolist(),
              "PSNR in/out vs clean", float(-10 * torch.log10(((x - clean) ** 2).mean())),
              float(-10 * torch.log10(((y.detach() - clean) ** 2).mean())))


def main():
    AD = load("/root/reference/AdaIR-main/net/model.py", "ref_adair_model")
    nets(AD)
    for name, dim, heads, bias, b, (hi, wi), (h, w), seed, rs in FREMODULE_TRAIN_CASES:
        mod = AD.FreModule(dim, heads, bias).train()
        mod.load_state_dict(fre_train_state(mod, dim, seed, rs), strict=True)
        img, y2, r = fre_train_inputs(dim, b, hi, wi, h, w, seed)
        y2.requires_grad_(True)
        with torch.no_grad():      # the mask half sizes, for the record (model.py:341-346)
            xc = mod.conv1(torch.nn.functional.interpolate(img, (h, w), mode="bilinear"))
            thr = mod.rate_conv(torch.nn.functional.adaptive_avg_pool2d(xc, 1)).sigmoid()[:, :, 0, 0]
            counts = torch.stack([(h // 128 * thr[:, 0]).int(), (w // 128 * thr[:, 1]).int()], 1)
        out = mod(img, y2)
        loss = (out * r).mean()
        loss.backward()
        ys, yc = fre_sample(out)
        gs, gc = fre_sample(y2.grad)
        rec = {"y": ys, "y_sum": yc, "loss": np.array(float(loss.detach())), "counts": counts.numpy(), "thr": thr.numpy(),
               "dy": gs, "dy_sum": gc, "keys": np.array(sorted(mod.state_dict()))}
        ng = 0
        for k, p in mod.named_parameters():
            if p.grad is not None:
                rec["g." + k] = p.grad.numpy()
                ng += 1
        np.savez_compressed(os.path.join(GOLD, name + ".npz"), **rec)
        print(name, tuple(out.shape), "loss", float(loss), "counts", counts.tolist(), "thr", thr.tolist(), "grads", ng, "of",
              sum(1 for _ in mod.parameters()), "no grad:", [k for k, p in mod.named_parameters() if p.grad is None])


if __name__ == "__main__":
    main()
