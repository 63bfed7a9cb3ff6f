"""b200ir.MoCEIR (image-restoration_b200/moce.py) against the UNMODIFIED reference moce_ir.MoCEIR: checkpoint contract
on the CPU (keys, shapes, the constructor-computed `complexity` buffers), whole-model outputs and expert choices on the
GPU (tests/golden/moceir_*.npz from oracle/gen_golden_moceir.py).  SURVEY.md §8 f1 / BASELINE configs[2]."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from oracle import restormer_oracle as O
from oracle.gen_golden_cases import MOCEIR_CASES, synth_like


@pytest.mark.parametrize("case", MOCEIR_CASES, ids=[c[0] for c in MOCEIR_CASES])
def test_moceir_state_dict_contract(case):
    import b200ir
    name, cfg, b, h, w, wseed, iseed, fseed = case
    gold = np.load(os.path.join(GOLDEN, name + ".npz"))
    sd = b200ir.MoCEIR(**cfg).state_dict()
    assert sorted(sd) == [str(k) for k in gold["keys"]]
    assert [str(tuple(sd[k].shape)) for k in sorted(sd)] == [str(s) for s in gold["shapes"]]
    cx = np.stack([sd[k].numpy() for k in sorted(sd) if k.endswith("complexity")])
    assert np.allclose(cx, gold["complexity"])      # expert parameter counts -> same router complexity buffers
    net = b200ir.MoCEIR(**cfg)
    net.load_state_dict(synth_like(sd, wseed), strict=True)


def _run(case, dtype):
    import b200ir
    name, cfg, b, h, w, wseed, iseed, fseed = case
    net = b200ir.MoCEIR(**cfg)
    net.load_state_dict(synth_like(net.state_dict(), wseed), strict=True)
    net = net.cuda().eval().set_compute_dtype(dtype).set_noise_device("cpu")
    picks = []
    for m in net.modules():
        if isinstance(m, b200ir.RoutingFunction):
            m.register_forward_hook(lambda mod, inp, out: picks.append(out[1].detach().cpu().reshape(-1)))
    x = O.synth_image(b, h, w, seed=iseed)
    torch.manual_seed(fseed)
    with torch.no_grad():
        y = net(x.cuda()).cpu()
    return y, torch.stack(picks).numpy(), np.load(os.path.join(GOLDEN, name + ".npz"))


@pytest.mark.gpu
@pytest.mark.parametrize("case", MOCEIR_CASES, ids=[c[0] for c in MOCEIR_CASES])
def test_moceir_fp32_vs_reference_golden(case):
    y, picks, gold = _run(case, torch.float32)
    assert (picks == gold["picks"]).all(), (picks.tolist(), gold["picks"].tolist())
    err = np.abs(y.numpy() - gold["y"]).max()
    print("MoCEIR fp32 max-abs vs reference:", case[0], err)
    assert err <= 1e-4, err


@pytest.mark.gpu
@pytest.mark.parametrize("case", MOCEIR_CASES, ids=[c[0] for c in MOCEIR_CASES])
def test_moceir_bf16_psnr(case):
    y, picks, gold = _run(case, torch.bfloat16)
    p = O.psnr(y, torch.from_numpy(gold["y"]))
    print("MoCEIR bf16 PSNR vs reference:", case[0], p, "same experts:", bool((picks == gold["picks"]).all()))
    if (picks == gold["picks"]).all():     # a flipped near-tie in the noisy top-k is a different (valid) network
        assert p >= 50.0, p


@pytest.mark.gpu
@pytest.mark.parametrize("p,C,B,H,W", [(4, 5, 2, 16, 24), (8, 3, 1, 20, 36), (16, 4, 2, 16, 40), (32, 2, 1, 64, 32), (32, 3, 1, 16, 16)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_circconv_patch_is_the_fft_product(p, C, B, H, W, dtype):
    """csrc/moce_ops.cu circular convolution == irfft2(rfft2(q) * rfft2(k)) per zero-padded p x p patch
    (FFTAttention, moce_ir.py:386-412), and the correlation mode is its gradient."""
    from b200ir import ops
    g = torch.Generator().manual_seed(p + C + H)
    q = (torch.rand(B, C, H, W, generator=g) - 0.5).to(dtype)
    k = (torch.rand(B, C, H, W, generator=g) - 0.5).to(dtype)

    def ref(qf, kf):
        ph, pw = (p - H % p) % p, (p - W % p) % p
        qp = torch.nn.functional.pad(qf, (0, pw, 0, ph)).unflatten(2, (-1, p)).unflatten(4, (-1, p)).permute(0, 1, 2, 4, 3, 5)
        kp = torch.nn.functional.pad(kf, (0, pw, 0, ph)).unflatten(2, (-1, p)).unflatten(4, (-1, p)).permute(0, 1, 2, 4, 3, 5)
        o = torch.fft.irfft2(torch.fft.rfft2(qp) * torch.fft.rfft2(kp), s=(p, p))
        o = o.permute(0, 1, 2, 4, 3, 5).flatten(4, 5).flatten(2, 3)
        return o[:, :, :H, :W]

    qf, kf = q.float().requires_grad_(True), k.float().requires_grad_(True)
    r = ref(qf, kf)
    gy = torch.rand(r.shape, generator=g) - 0.5
    r.backward(gy)
    y = ops.circconv_patch(q.cuda(), k.cuda(), p).float().cpu()
    tol = 1e-5 if dtype == torch.float32 else 2e-2
    assert (y - r.detach()).abs().max() <= tol * max(1.0, r.abs().max().item())
    if dtype == torch.float32:
        dq = ops.circconv_patch(gy.cuda(), k.cuda(), p, mode=1).cpu()
        dk = ops.circconv_patch(gy.cuda(), q.cuda(), p, mode=1).cpu()
        sc = max(1.0, qf.grad.abs().max().item(), kf.grad.abs().max().item())
        assert (dq - qf.grad).abs().max() <= 1e-5 * sc and (dk - kf.grad).abs().max() <= 1e-5 * sc


@pytest.mark.gpu
def test_moce_small_ops():
    from b200ir import ops
    g = torch.Generator().manual_seed(5)
    a = (torch.rand(2, 6, 8, 12, generator=g) - 0.5).cuda()
    b = (torch.rand(2, 6, 8, 12, generator=g) - 0.5).cuda()
    assert torch.allclose(ops.ew(a, b, ops.EW_MUL), a * b, atol=1e-6)
    assert torch.allclose(ops.ew(a, b, ops.EW_SILU_MUL), a * torch.nn.functional.silu(b), atol=1e-6)
    assert torch.allclose(ops.ew(a, None, ops.EW_GELU), torch.nn.functional.gelu(a), atol=1e-6)
    assert torch.allclose(ops.spatial_mean(a), a.mean((-2, -1)), atol=1e-6)
    assert torch.allclose(ops.spatial_mean(a, gelu=True), torch.nn.functional.gelu(a).mean((-2, -1)), atol=1e-6)
    gg = torch.tensor([0.25, -2.0], device="cuda")
    y = ops.scale_rows(a, gg)
    ops.scale_rows(b, gg, out=y, accumulate=True)
    assert torch.allclose(y, (a + b) * gg.view(2, 1, 1, 1), atol=1e-6)
