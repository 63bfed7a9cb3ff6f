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


# ---- training mode: SparseDispatcher routing + auxiliary losses + every parameter gradient ----------------------------
from oracle.gen_golden_cases import MOCEIR_TRAIN_CASES  # noqa: E402


def _train_case(case, dtype):
    """one training forward + backward of b200ir.MoCEIR with the objective of MoCE-IR-main/src/train.py:59-72"""
    import b200ir
    name, cfg, b, h, w, wseed, iseed, fseed = case
    gold = np.load(os.path.join(GOLDEN, name + ".npz"))
    net = b200ir.MoCEIR(**cfg)
    net.load_state_dict(synth_like(net.state_dict(), wseed), strict=True)
    net = net.cuda().train().set_compute_dtype(dtype).set_noise_device("cpu")
    picks = []
    for m in net.modules():
        if isinstance(m, b200ir.RoutingFunction):
            m.register_forward_hook(lambda mod, inp, out: picks.append(out[1].detach().cpu().reshape(-1)))
    clean = O.synth_image(b, h, w, seed=iseed)
    noise = torch.randn(clean.shape, generator=torch.Generator().manual_seed(iseed + 1)) * (25.0 / 255.0)
    x = (clean + noise).clamp(0, 1)
    torch.manual_seed(fseed)
    y = net(x.cuda())
    bal = net.total_loss
    loss = b200ir.L1Loss()(y, clean.cuda()) + float(gold["balance_w"]) * bal
    loss.backward()
    torch.cuda.synchronize()
    return net, y.detach().float().cpu(), float(bal.detach()), float(loss.detach()), torch.stack(picks).numpy(), gold


@pytest.mark.gpu
@pytest.mark.parametrize("case", MOCEIR_TRAIN_CASES, ids=[c[0] for c in MOCEIR_TRAIN_CASES])
def test_moceir_training_step_fp32_vs_reference_golden(case):
    """net.train(): dispatch / combine, importance + load losses, total_loss and ALL parameter gradients against the
    unmodified reference (oracle/gen_golden_moceir_train.py).  Gradients are stored as per-tensor scale + fp16 mantissas:
    the comparison allows that quantisation (5e-4 of the tensor's largest entry)."""
    net, y, bal, loss, picks, gold = _train_case(case, torch.float32)
    assert (picks == gold["picks"]).all(), (picks.tolist(), gold["picks"].tolist())
    assert np.abs(y.numpy() - gold["y"]).max() <= 1e-4
    assert abs(bal - float(gold["total_loss"])) <= 1e-4 * abs(float(gold["total_loss"])) + 1e-6, (bal, float(gold["total_loss"]))
    assert abs(loss - float(gold["loss"])) <= 1e-5, (loss, float(gold["loss"]))
    bad, seen = [], 0
    for n, p in net.named_parameters():
        if "g." + n in gold.files:
            seen += 1
            ref = torch.from_numpy(gold["g." + n].astype(np.float32)) * float(gold["s." + n])
            assert p.grad is not None, n
            a, r = p.grad.detach().float().cpu().flatten().double(), ref.flatten().double()
            tol = 1.5e-3 * float(r.abs().max()) + 1e-9          # fp16 mantissas of the fixture + fp32 summation order
            if float((a - r).abs().max()) > tol:
                bad.append((n, float((a - r).abs().max()), float(r.abs().max())))
        else:   # an expert no sample was routed to: the reference leaves .grad None
            assert p.grad is None or float(p.grad.abs().max()) == 0.0, n
    assert seen == sum(1 for f in gold.files if f.startswith("g."))
    assert not bad, bad[:10]


@pytest.mark.gpu
def test_moceir_training_step_bf16():
    """bf16 activations: same experts, loss within 2 %, gradient direction of the large tensors preserved."""
    case = MOCEIR_TRAIN_CASES[0]
    net, y, bal, loss, picks, gold = _train_case(case, torch.bfloat16)
    if not (picks == gold["picks"]).all():
        pytest.skip("a near-tie of the noisy top-k flipped under bf16: a different (valid) routing")
    assert abs(loss - float(gold["loss"])) <= 2e-2 * abs(float(gold["loss"]))
    cos = []
    for n, p in net.named_parameters():
        if "g." + n in gold.files and p.numel() >= 256:
            r = torch.from_numpy(gold["g." + n].astype(np.float32)).flatten().double()
            a = p.grad.detach().float().cpu().flatten().double()
            cos.append(float(torch.dot(a, r) / (a.norm() * r.norm()).clamp_min(1e-300)))
    assert np.median(cos) >= 0.99 and np.mean(np.array(cos) >= 0.9) >= 0.95, (np.median(cos), sorted(cos)[:5])


@pytest.mark.gpu
def test_moceir_train_step_with_balance_loss():
    """b200ir.train_step(net, opt, ..., balance_loss_weight=w) = MoCE-IR-main/src/train.py:59-72 + AdamW: the first step's
    loss is the reference's, gradients are the golden ones (side-stream weight gradients included), parameters move."""
    import b200ir
    case = MOCEIR_TRAIN_CASES[0]
    name, cfg, b, h, w, wseed, iseed, fseed = case
    gold = np.load(os.path.join(GOLDEN, name + ".npz"))
    net = b200ir.MoCEIR(**cfg)
    net.load_state_dict(synth_like(net.state_dict(), wseed), strict=True)
    net = net.cuda().train().set_noise_device("cpu")
    opt = b200ir.AdamW(net.parameters(), lr=2e-4)
    clean = O.synth_image(b, h, w, seed=iseed)
    noise = torch.randn(clean.shape, generator=torch.Generator().manual_seed(iseed + 1)) * (25.0 / 255.0)
    x = (clean + noise).clamp(0, 1).cuda()
    before = opt.flat_p.clone()
    torch.manual_seed(fseed)
    loss = b200ir.train_step(net, opt, x, clean.cuda(), balance_loss_weight=float(gold["balance_w"]))
    assert abs(float(loss) - float(gold["loss"])) <= 1e-5
    for n, p in net.named_parameters():
        if "g." + n in gold.files:
            ref = torch.from_numpy(gold["g." + n].astype(np.float32)) * float(gold["s." + n])
            a = p.grad.detach().float().cpu()
            assert float((a - ref).abs().max()) <= 1.5e-3 * float(ref.abs().max()) + 1e-9, n
    assert not torch.equal(before, opt.flat_p)
    loss2 = b200ir.train_step(net, opt, x, clean.cuda(), balance_loss_weight=float(gold["balance_w"]))
    assert torch.isfinite(loss2)


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_dispatch_combine_kernels(dtype):
    """rows_copy (gather / gate-weighted scatter-accumulate), rowdot and the spatial-mean backward against torch."""
    from b200ir import ops
    g = torch.Generator().manual_seed(4)
    x = torch.randn(5, 6, 8, 12, generator=g).cuda().to(dtype)
    idx = torch.tensor([3, 0, 4], dtype=torch.int32).cuda()
    y = ops.rows_copy(x, torch.empty(3, 6, 8, 12, device="cuda", dtype=dtype), 3, src_idx=idx)
    assert torch.equal(y, x[idx.long()])
    sc = torch.tensor([0.5, -2.0, 1.25]).cuda()
    out = torch.ones(5, 6, 8, 12, device="cuda", dtype=dtype)
    ops.rows_copy(y, out, 3, dst_idx=idx, scale=sc, accumulate=True)
    ref = torch.ones(5, 6, 8, 12, device="cuda")
    ref.index_add_(0, idx.long(), y.float() * sc.view(-1, 1, 1, 1))
    assert (out.float() - ref).abs().max() <= (1e-6 if dtype == torch.float32 else 3e-2)
    d = ops.rowdot(y, x, 3, b_idx=idx)
    rd = (y.float() * x[idx.long()].float()).flatten(1).sum(1)
    assert (d - rd).abs().max() <= 1e-3 * rd.abs().max()
    for shp in ((4, 3, 7, 5), (2, 48, 64, 64)):      # unaligned rows (scalar loads) and rows spread over the whole cluster
        u = torch.randn(*shp, generator=g).cuda().to(dtype)
        v = torch.randn(*shp, generator=g).cuda().to(dtype)
        d = ops.rowdot(u, v, shp[0])
        rd = (u.double() * v.double()).flatten(1).sum(1)
        assert (d.double() - rd).abs().max() <= 1e-4 * (u.double() * v.double()).abs().flatten(1).sum(1).max()
        assert torch.equal(d, ops.rowdot(u, v, shp[0]))
    for gelu in (False, True):
        xs = x.float().clone().requires_grad_(True)
        m = (torch.nn.functional.gelu(xs) if gelu else xs).mean(dim=(-2, -1))
        gm = torch.randn(m.shape, generator=g).cuda()
        m.backward(gm)
        dx = ops.spatial_mean_bwd(gm, x, gelu=gelu)
        assert (dx.float() - xs.grad).abs().max() <= (1e-6 if dtype == torch.float32 else 2e-2) * max(1.0, float(xs.grad.abs().max()))
