"""SURVEY.md §8 f2 / f3 variants against outputs of the UNMODIFIED reference modules (oracle/gen_golden_recti.py):
RectiFormer's split-gate FFN and element-wise gated attention with per-sample dynamic temperature
(experiment/RectiFormer.py:231-325) and AdaIR's Chanel_Cross_Attention (AdaIR-main/net/model.py:177-216)."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from oracle import restormer_oracle as O
from oracle.gen_golden_cases import ADAIR_CCA_CASES, RECTI_CASES, synth_like

pytestmark = pytest.mark.gpu


def _rel(a, b):
    return ((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30)).item()


@pytest.mark.parametrize("case", RECTI_CASES, ids=[c[0] for c in RECTI_CASES])
def test_rectiformer_variants_vs_reference_golden(case):
    import b200ir
    name, kind, dim, heads, f, bias, b, h, w, seed = case
    gold = np.load(os.path.join(GOLDEN, name + ".npz"))
    mod = b200ir.GatedSpatialFFN(dim, f, bias) if kind == "ffn" else b200ir.ElementwiseGatedAttention(dim, heads, bias)
    assert sorted(mod.state_dict()) == [str(k) for k in gold["keys"]]
    mod.load_state_dict(synth_like(mod.state_dict(), seed), strict=True)
    mod = mod.cuda().eval()
    x = (O.synth_image(b, h, w, seed=seed + 100, c=dim) * 2 - 1).cuda()
    ref = torch.from_numpy(gold["y"])
    with torch.no_grad():
        if kind == "attn_dyn":
            dt = (torch.rand(b, generator=torch.Generator().manual_seed(seed + 7)) * 2).view(b, 1, 1, 1).cuda()
            y32 = mod(x, dt).cpu()
            y16 = mod(x.bfloat16(), dt).float().cpu()
        else:
            y32 = mod(x).cpu()
            y16 = mod(x.bfloat16()).float().cpu()
    assert (y32 - ref).abs().max() <= 1e-4 * max(1.0, ref.abs().max().item()), float((y32 - ref).abs().max())
    assert _rel(y16, ref) < 2e-2, _rel(y16, ref)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("case", RECTI_CASES, ids=[c[0] for c in RECTI_CASES])
def test_rectiformer_variants_backward_vs_reference_golden(case, dtype):
    """The same modules under autograd: input, parameter and dynamic-temperature gradients against the reference module's
    own autograd (sigmoid-gate backward of the split-gate FFN, per-sample temperature in the MDTA backward)."""
    import b200ir
    name, kind, dim, heads, f, bias, b, h, w, seed = case
    gold = np.load(os.path.join(GOLDEN, name + ".npz"))
    mod = b200ir.GatedSpatialFFN(dim, f, bias) if kind == "ffn" else b200ir.ElementwiseGatedAttention(dim, heads, bias)
    mod.load_state_dict(synth_like(mod.state_dict(), seed), strict=True)
    mod = mod.cuda().train()
    x = (O.synth_image(b, h, w, seed=seed + 100, c=dim) * 2 - 1).cuda().to(dtype).requires_grad_(True)
    gy = torch.randn(gold["y"].shape, generator=torch.Generator().manual_seed(seed + 9)).cuda().to(dtype)
    if kind == "attn_dyn":
        dt = (torch.rand(b, generator=torch.Generator().manual_seed(seed + 7)) * 2).view(b, 1, 1, 1).cuda().requires_grad_(True)
        mod(x, dt).backward(gy)
        got = {"ddt": dt.grad}
    else:
        mod(x).backward(gy)
        got = {}
    got["dx"] = x.grad
    for k, q in mod.named_parameters():
        got["g." + k] = q.grad
    tol = 2e-4 if dtype == torch.float32 else 4e-2
    for k, v in got.items():
        assert v is not None, k
        ref = torch.from_numpy(gold[k])
        r = _rel(v.float().cpu().reshape(ref.shape), ref)
        # the temperature gradient is a cancelling sum (rows of dS sum to zero): looser under bf16 activations
        lim = tol if dtype == torch.float32 or "temperature" not in k and k != "ddt" else 0.25
        assert r <= lim, (k, r)


@pytest.mark.parametrize("case", ADAIR_CCA_CASES, ids=[c[0] for c in ADAIR_CCA_CASES])
def test_adair_chanel_cross_attention_vs_reference_golden(case):
    """b200ir.CrossAttention with a 3x3 kv depthwise conv == AdaIR's Chanel_Cross_Attention (same state_dict keys)."""
    import b200ir
    name, dim, heads, bias, b, h, w, seed = case
    gold = np.load(os.path.join(GOLDEN, name + ".npz"))
    mod = b200ir.CrossAttention(dim, heads, bias)
    mod.kv_dwconv = torch.nn.Conv2d(dim * 2, dim * 2, kernel_size=3, stride=1, padding=1, groups=dim * 2, bias=bias)
    assert sorted(mod.state_dict()) == [str(k) for k in gold["keys"]]
    mod.load_state_dict(synth_like(mod.state_dict(), seed), strict=True)
    mod = mod.cuda().eval()
    x = (O.synth_image(b, h, w, seed=seed + 100, c=dim) * 2 - 1).cuda()
    y2 = (O.synth_image(b, h, w, seed=seed + 300, c=dim) * 2 - 1).cuda()
    ref = torch.from_numpy(gold["y"])
    with torch.no_grad():
        y32 = mod(x, y2).cpu()
        y16 = mod(x.bfloat16(), y2.bfloat16()).float().cpu()
    assert (y32 - ref).abs().max() <= 1e-4 * max(1.0, ref.abs().max().item())
    assert _rel(y16, ref) < 2e-2


def test_accelerate_adopts_adair_num_head_cross_attention():
    """dropin.accelerate() takes a module with Chanel_Cross_Attention's structure (`num_head`, 3x3 kv dwconv)."""
    import b200ir
    from torch_models import RefAdaIRFusion
    torch.manual_seed(2)
    net = RefAdaIRFusion(32, 2)
    x = O.synth_image(2, 16, 24, seed=6, c=32) * 2 - 1
    with torch.no_grad():
        ref = net(x)
    net = net.cuda()
    net, stats = b200ir.accelerate(net, verify=True)
    assert stats == {"cross_attention": 1}, stats
    assert isinstance(net.cca, b200ir.CrossAttention) and net.cca.num_heads == 2
    with torch.no_grad():
        got = net(x.cuda()).cpu()
    assert (got - ref).abs().max() <= 2e-4 * max(1.0, ref.abs().max().item())


def test_accelerate_adopts_rectiformer_variants():
    """dropin.accelerate() swaps modules with the structure of RectiFormer's GatedSpatialFFN / ElementwiseGatedAttention
    (the stand-ins in tests/torch_models.py are bit-identical to experiment/RectiFormer.py:231-325, checked in the build
    container), keeps the dynamic-temperature argument, and the adopted block trains."""
    import b200ir
    from torch_models import RefRectiBlock
    torch.manual_seed(4)
    net = RefRectiBlock(32, 2, 2.0, True)
    x = O.synth_image(2, 16, 24, seed=8, c=32) * 2 - 1
    dt = (torch.rand(2, generator=torch.Generator().manual_seed(5)) * 2).view(2, 1, 1, 1)
    with torch.no_grad():
        ref, ref_dt = net(x), net(x, dt)
    xr = x.clone().requires_grad_(True)
    net(xr, dt).square().mean().backward()
    gref = {k: q.grad.clone() for k, q in net.named_parameters()}
    net.zero_grad()
    net = net.cuda()
    net, stats = b200ir.accelerate(net, verify=True)
    assert stats.get("elementwise_gated_attention") == 1 and stats.get("split_gate_ffn") == 1, stats
    assert isinstance(net.attn, b200ir.ElementwiseGatedAttention) and isinstance(net.ffn, b200ir.GatedSpatialFFN)
    with torch.no_grad():
        got, got_dt = net(x.cuda()).cpu(), net(x.cuda(), dt.cuda()).cpu()
    assert (got - ref).abs().max() <= 2e-4 * max(1.0, ref.abs().max().item())
    assert (got_dt - ref_dt).abs().max() <= 2e-4 * max(1.0, ref_dt.abs().max().item())
    xg = x.cuda().requires_grad_(True)
    net(xg, dt.cuda()).square().mean().backward()
    assert _rel(xg.grad.cpu(), xr.grad) <= 1e-3
    for k, q in net.named_parameters():
        assert q.grad is not None and _rel(q.grad.cpu(), gref[k]) <= 2e-3, (k, _rel(q.grad.cpu(), gref[k]))


def test_adair_fremodule_vs_reference_golden():
    """b200ir.FreModule (bilinear resize, conv3x3, few-bin low-pass instead of fft2/mask/ifft2, three channel cross
    attentions, spatial / channel gates, per-channel affine) against AdaIR-main/net/model.py FreModule outputs."""
    import b200ir
    from oracle.gen_golden_cases import FREMODULE_CASES, fre_sample
    for name, dim, heads, bias, b, (hi, wi), (h, w), seed in FREMODULE_CASES:
        gold = np.load(os.path.join(GOLDEN, name + ".npz"))
        mod = b200ir.FreModule(dim, heads, bias)
        assert sorted(mod.state_dict()) == [str(k) for k in gold["keys"]]
        sd = synth_like(mod.state_dict(), seed)
        sd["para1"] = 0.5 + torch.rand(dim, 1, 1, generator=torch.Generator().manual_seed(seed + 1))
        mod.load_state_dict(sd, strict=True)
        mod = mod.cuda().eval()
        img = O.synth_image(b, hi, wi, seed=seed + 100).cuda()
        y2 = (O.synth_image(b, h, w, seed=seed + 300, c=dim) * 2 - 1).cuda()
        with torch.no_grad():
            y = mod(img, y2)
            hf, lf = mod.fft(b200ir.ops.bilinear_resize(img, h, w))
            y16 = mod(img, y2.bfloat16()).float()
        for got, key, tol in ((hf, "high", 2e-4), (lf, "low", 2e-4), (y, "y", 5e-4)):
            s, c = fre_sample(got)
            ref = gold[key]
            scale = max(1.0, float(np.abs(ref).max()))
            assert np.abs(s - ref).max() <= tol * scale, (name, key, float(np.abs(s - ref).max()))
            assert np.allclose(c, gold[key + "_sum"], rtol=1e-3, atol=1e-5), (name, key, c, gold[key + "_sum"])
        s16, _ = fre_sample(y16)
        assert np.linalg.norm(s16 - gold["y"]) <= 3e-2 * np.linalg.norm(gold["y"]), name


def test_adair_backward_kernels_vs_torch():
    """csrc/adair_bwd.cu piece by piece against torch autograd of the same formulas (fp32, and bf16 loosely)."""
    import torch.nn.functional as F
    from b200ir import autograd_ops as A
    g = torch.Generator().manual_seed(11)
    B, C, H, W = 2, 6, 20, 12
    for dtype, tol in ((torch.float32, 2e-5), (torch.bfloat16, 3e-2)):
        # --- low-pass split: fft2 -> mask -> ifft2 -> abs
        x = torch.randn(B, C, H, W, generator=g).cuda()
        counts = torch.tensor([[2, 1], [0, 3]], dtype=torch.int32).cuda()
        xr = x.to(dtype).double().requires_grad_(True)
        mask = torch.zeros(B, 1, H, W, dtype=torch.float64, device="cuda")
        for i, (h_, w_) in enumerate(counts.tolist()):
            mask[i, :, H // 2 - h_:H // 2 + h_, W // 2 - w_:W // 2 + w_] = 1
        f = torch.roll(torch.fft.fft2(xr, norm="forward"), (H // 2, W // 2), (2, 3))
        hi = torch.fft.ifft2(torch.roll(f * (1 - mask), (-(H // 2), -(W // 2)), (2, 3)), norm="forward").abs()
        lo = torch.fft.ifft2(torch.roll(f * mask, (-(H // 2), -(W // 2)), (2, 3)), norm="forward").abs()
        gh, gl = torch.randn(B, C, H, W, generator=g).cuda().to(dtype), torch.randn(B, C, H, W, generator=g).cuda().to(dtype)
        (hi * gh.double() + lo * gl.double()).sum().backward()
        xm = x.to(dtype).requires_grad_(True)
        mh, ml = A.lowfreq_split(xm, counts, 64)
        assert (mh.double() - hi).abs().max() <= tol * 10 and (ml.double() - lo).abs().max() <= tol * 10
        (mh.float() * gh.float() + ml.float() * gl.float()).sum().backward()
        assert (xm.grad.double() - xr.grad).abs().max() <= tol * 10 * max(1.0, float(xr.grad.abs().max())), dtype

        # --- FreRefine gate: low * sigmoid(conv7x7([max_c high, mean_c high])) + high * cw
        low = torch.randn(B, C, H, W, generator=g).cuda().to(dtype)
        # distinct values along the channels of a pixel (a tie would make the arg-max, hence the gradient, ambiguous)
        order = torch.rand(B, C, H, W, generator=g).argsort(dim=1).float()
        high = ((order - C / 2) * 0.5 + 0.125 * torch.rand(B, C, H, W, generator=g)).cuda().to(dtype)
        w7 = (torch.randn(1, 2, 7, 7, generator=g) * 0.2).cuda()
        cw = torch.rand(B, C, generator=g).cuda()
        gy = torch.randn(B, C, H, W, generator=g).cuda().to(dtype)
        lr, hr, wr, cr = (t.double().requires_grad_(True) for t in (low, high, w7, cw))
        sg = torch.sigmoid(F.conv2d(torch.cat([hr.max(1, keepdim=True)[0], hr.mean(1, keepdim=True)], 1), wr, padding=3))
        yr = lr * sg + hr * cr.view(B, C, 1, 1)
        (yr * gy.double()).sum().backward()
        lm, hm, wm, cm = low.clone().requires_grad_(True), high.clone().requires_grad_(True), w7.clone().requires_grad_(True), \
            cw.clone().requires_grad_(True)
        ym = A.fre_refine_gate(lm, hm, wm, cm)
        assert (ym.double() - yr).abs().max() <= tol * 10
        (ym.float() * gy.float()).sum().backward()
        for got, ref, nm in ((lm.grad, lr.grad, "dlow"), (hm.grad, hr.grad, "dhigh"), (wm.grad, wr.grad, "dw7"), (cm.grad, cr.grad, "dcw")):
            assert (got.double() - ref).abs().max() <= tol * 10 * max(1.0, float(ref.abs().max())), (dtype, nm)

        # --- AdaptiveMaxPool2d(1) and the final affine
        planes = (torch.rand(B, C, H * W, generator=g).argsort(dim=2).float() / 64 - 1).view(B, C, H, W).cuda().to(dtype)
        xr = planes.double().requires_grad_(True)
        gm = torch.randn(B, C, generator=g).cuda()
        (xr.amax(dim=(2, 3)) * gm.double()).sum().backward()
        xm = planes.clone().requires_grad_(True)
        (A.spatial_max(xm) * gm).sum().backward()
        assert torch.equal(xm.grad.double(), xr.grad.to(dtype).double())
        p1, p2 = torch.randn(C, 1, 1, generator=g).cuda(), torch.randn(C, 1, 1, generator=g).cuda()
        ar, br, q1, q2 = (t.double().requires_grad_(True) for t in (low, high, p1, p2))
        ((ar * q1 + br * q2) * gy.double()).sum().backward()
        am, bm, m1, m2 = low.clone().requires_grad_(True), high.clone().requires_grad_(True), p1.clone().requires_grad_(True), \
            p2.clone().requires_grad_(True)
        (A.affine2(am, bm, m1, m2).float() * gy.float()).sum().backward()
        for got, ref, nm in ((am.grad, ar.grad, "da"), (bm.grad, br.grad, "db"), (m1.grad, q1.grad, "dp1"), (m2.grad, q2.grad, "dp2")):
            assert (got.double() - ref).abs().max() <= tol * 10 * max(1.0, float(ref.abs().max())), (dtype, nm)


def test_adair_fremodule_training_vs_reference_golden():
    """b200ir.FreModule under autograd against the reference FreModule's loss, d loss / d feature and parameter gradients
    (oracle/gen_golden_adair_train.py; AdaIR-main/net/model.py:295-372 as AdaIR-main/train.py differentiates it)."""
    import b200ir
    from oracle.gen_golden_adair_train import fre_train_inputs, fre_train_state
    from oracle.gen_golden_cases import FREMODULE_TRAIN_CASES, fre_sample
    for name, dim, heads, bias, b, (hi, wi), (h, w), seed, rs in FREMODULE_TRAIN_CASES:
        gold = np.load(os.path.join(GOLDEN, name + ".npz"))
        for dtype in (torch.float32, torch.bfloat16):
            mod = b200ir.FreModule(dim, heads, bias)
            mod.load_state_dict(fre_train_state(mod, dim, seed, rs), strict=True)
            mod = mod.cuda().train()
            img, y2, r = fre_train_inputs(dim, b, hi, wi, h, w, seed)
            y2 = y2.cuda().to(dtype).requires_grad_(True)
            out = mod(img.cuda(), y2)
            assert out.dtype == dtype
            loss = (out.float() * r.cuda()).mean()
            loss.backward()
            f32 = dtype == torch.float32
            s, c = fre_sample(out)
            assert np.abs(s - gold["y"]).max() <= (5e-4 if f32 else 6e-2) * max(1.0, float(np.abs(gold["y"]).max())), (name, dtype)
            assert abs(float(loss.detach()) - float(gold["loss"])) <= (2e-6 if f32 else 2e-4), (name, dtype, float(loss), float(gold["loss"]))
            gs, gc = fre_sample(y2.grad)
            ref = gold["dy"]
            assert np.linalg.norm(gs - ref) <= (2e-3 if f32 else 4e-2) * np.linalg.norm(ref), (name, dtype, "dy")
            if f32:
                assert np.allclose(gc, gold["dy_sum"], rtol=2e-3, atol=1e-9), (name, gc, gold["dy_sum"])
            bad = []
            gmax = max(float(np.linalg.norm(gold[k])) for k in gold.files if k.startswith("g."))
            for k, p in mod.named_parameters():
                if "g." + k in gold.files:
                    ref = torch.from_numpy(gold["g." + k])
                    assert p.grad is not None, (name, k)
                    got = p.grad.detach().float().cpu().view(ref.shape)
                    # (some gradients are zero up to rounding in the reference, norm ~1e-9: hence the absolute term)
                    if float((got - ref).norm()) > (3e-3 if f32 else 6e-2) * float(ref.norm()) + 1e-6 * gmax:
                        bad.append((k, float((got - ref).norm()), float(ref.norm())))
                else:
                    assert p.grad is None, (name, k)      # conv, score_gen, rate_conv: untouched by the reference's backward too
            assert not bad, (name, dtype, bad)


def test_adair_network_vs_reference_golden():
    """b200ir.AdaIR (Restormer U-Net + three FreModules, AdaIR-main/net/model.py:375-497): restored image at 64x64 and at
    512x512 (non-empty spectrum mask in fre3), and the L1 training step's loss and parameter gradients."""
    import b200ir
    from oracle.gen_golden_adair_train import adair_net_inputs, adair_net_state
    from oracle.gen_golden_cases import ADAIR_NET_CASES, fre_sample, grad_sample
    for name, cfg, b, h, w, seed, train, rs in ADAIR_NET_CASES:
        gold = np.load(os.path.join(GOLDEN, name + ".npz"))
        net = b200ir.AdaIR(**cfg)
        assert sorted(net.state_dict()) == sorted(str(k) for k in gold["keys"])
        net.load_state_dict(adair_net_state(net, seed, rs), strict=True)
        net = net.cuda()
        x, clean = adair_net_inputs(b, h, w, seed)
        if not train:
            net.eval()
            with torch.no_grad():
                y32 = net.set_compute_dtype(torch.float32)(x.cuda())
                y16 = net.set_compute_dtype(torch.bfloat16)(x.cuda()).float()
            s, c = fre_sample(y32)
            assert np.abs(s - gold["y"]).max() <= 3e-4, (name, float(np.abs(s - gold["y"]).max()))
            assert np.allclose(c, gold["y_sum"], rtol=1e-4), (name, c, gold["y_sum"])
            s16, _ = fre_sample(y16)
            mse = float(((s16 - gold["y"]) ** 2).mean())
            assert 10 * np.log10(1.0 / mse) >= 40.0, (name, 10 * np.log10(1.0 / mse))
            continue
        for dtype in (torch.float32, torch.bfloat16):
            f32 = dtype == torch.float32
            net.set_compute_dtype(dtype).train()
            net.zero_grad(set_to_none=True)
            y = net(x.cuda())
            loss = b200ir.L1Loss()(y, clean.cuda())
            loss.backward()
            ref = torch.from_numpy(gold["y"])
            assert (y.detach().float().cpu() - ref).abs().max() <= (3e-4 if f32 else 8e-2), (name, dtype)
            assert abs(float(loss.detach()) - float(gold["loss"])) <= (1e-5 if f32 else 3e-3), (name, dtype, float(loss.detach()))
            bad, gmax = [], max(float(gold[k]) for k in gold.files if k.startswith("s."))
            for k, p in net.named_parameters():
                if "g." + k not in gold.files:
                    assert p.grad is None, (name, k)
                    continue
                assert p.grad is not None, (name, k)
                r = gold["g." + k].astype(np.float64) * float(gold["s." + k])
                g = grad_sample(p.grad.detach().float().cpu().numpy()).astype(np.float64)
                err = np.abs(g - r).max()
                if err > (2e-3 if f32 else 0.25) * np.abs(r).max() + 1e-5 * gmax:
                    bad.append((k, float(err), float(np.abs(r).max())))
            assert not bad, (name, dtype, bad[:8], len(bad))


def test_accelerate_adopts_adair_frequency_module():
    """dropin.accelerate() takes a module with FreModule's structure (tests/torch_models.py RefFreModule is bit-identical
    to AdaIR-main/net/model.py:295-372, checked in the build container); the adopted module keeps the state_dict and trains."""
    import b200ir
    from torch_models import RefAdaIRTail
    torch.manual_seed(6)
    net = RefAdaIRTail(32, 2)
    with torch.no_grad():
        net.fre.para1.add_(0.5 + torch.rand(32, 1, 1))
    keys = list(net.state_dict())
    img = O.synth_image(2, 24, 40, seed=9)
    ref = net(img)
    ref.square().mean().backward()
    gref = {k: q.grad.clone() for k, q in net.named_parameters() if q.grad is not None}
    net.zero_grad(set_to_none=True)
    net = net.cuda()
    net, stats = b200ir.accelerate(net, verify=True)
    assert stats == {"frequency_module": 1}, stats
    assert isinstance(net.fre, b200ir.FreModule) and list(net.state_dict()) == keys
    got = net(img.cuda())
    assert (got.detach().cpu() - ref.detach()).abs().max() <= 2e-4 * max(1.0, ref.abs().max().item())
    got.square().mean().backward()
    for k, q in net.named_parameters():
        if k in gref:
            assert q.grad is not None and _rel(q.grad.cpu(), gref[k]) <= 3e-3, (k, _rel(q.grad.cpu(), gref[k]))
        else:
            assert q.grad is None, k
