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
