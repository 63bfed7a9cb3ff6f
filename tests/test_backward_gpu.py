"""GPU parity of the backward (training) kernels, through the C ABI, against torch autograd over the CPU oracle —
the same gradients ``loss.backward()`` gives the reference (MoCE-IR-main/src/train.py:59-72).

Tolerances: fp32 path — every gradient within 1e-4 of the oracle's relative to that gradient's max-abs (the
BASELINE.json fp32 bar applied to gradients); bf16 path — relative L2 error <= 3e-2 per gradient tensor."""
import pytest
import torch
import torch.nn.functional as F

from oracle import restormer_oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def b200ir():
    import b200ir as m
    return m


def close32(a, ref, tol=1e-4):
    a, ref = a.detach().double().cpu(), ref.detach().double().cpu()
    assert a.shape == ref.shape, (a.shape, ref.shape)
    scale = ref.abs().max().clamp_min(1e-6).item()
    err = (a - ref).abs().max().item()
    assert err <= tol * scale, f"max-abs {err:.3e} vs scale {scale:.3e}"


def rel(a, ref):
    a, ref = a.detach().double().cpu(), ref.detach().double().cpu()
    return ((a - ref).norm() / ref.norm().clamp_min(1e-30)).item()


def leaf(t, dtype=None):
    t = t.cuda()
    if dtype is not None:
        t = t.to(dtype)
    return t.detach().requires_grad_(True)


@pytest.mark.parametrize("ln", ["none", "WithBias", "BiasFree"])
@pytest.mark.parametrize("Cin,Cout,H,W,bias,res", [(48, 144, 16, 16, False, False), (30, 17, 9, 13, True, True),
                                                   (96, 96, 24, 8, True, False), (127, 48, 16, 8, False, True)])
def test_ln_conv1x1_backward(b200ir, ln, Cin, Cout, H, W, bias, res):
    from b200ir import autograd_ops as A, _lib as L
    g = torch.Generator().manual_seed(3)
    x = torch.randn(2, Cin, H, W, generator=g)
    w = torch.randn(Cout, Cin, 1, 1, generator=g) * 0.2
    b = torch.randn(Cout, generator=g) if bias else None
    lw = torch.rand(Cin, generator=g) + 0.5
    lb = torch.randn(Cin, generator=g) * 0.1
    r = torch.randn(2, Cout, H, W, generator=g) if res else None
    gy = torch.randn(2, Cout, H, W, generator=g)

    ins = [t.clone().requires_grad_(True) if t is not None else None for t in (x, w, b, lw, lb, r)]
    xr, wr, br, lwr, lbr, rr = ins
    n = xr if ln == "none" else O.layernorm(xr, lwr, lbr if ln == "WithBias" else None)
    y = F.conv2d(n, wr, br)
    if res:
        y = y + rr
    used = [t for t in ins if t is not None]
    refg = dict(zip([id(t) for t in used], torch.autograd.grad(y, used, gy, allow_unused=True)))

    lt = {"none": L.LN_NONE, "WithBias": L.LN_WITHBIAS, "BiasFree": L.LN_BIASFREE}[ln]
    xs, ws, bs, lws, lbs, rs = [leaf(t) if t is not None else None for t in (x, w, b, lw, lb, r)]
    out = A.ln_conv1x1(xs, ws, bs, lt, lws if ln != "none" else None, lbs if ln == "WithBias" else None, rs)
    close32(out, y, 1e-5)
    out.backward(gy.cuda())
    close32(xs.grad, refg[id(xr)])
    close32(ws.grad, refg[id(wr)])
    if bias:
        close32(bs.grad, refg[id(br)])
    if res:
        close32(rs.grad, refg[id(rr)])
    if ln != "none":
        close32(lws.grad, refg[id(lwr)])
    if ln == "WithBias":
        close32(lbs.grad, refg[id(lbr)])


@pytest.mark.parametrize("C,H,W,bias", [(24, 16, 16, False), (30, 7, 13, True), (96, 40, 24, True)])
def test_dwconv_backward(b200ir, C, H, W, bias):
    from b200ir import autograd_ops as A
    g = torch.Generator().manual_seed(4)
    x = torch.randn(2, C, H, W, generator=g)
    w = torch.randn(C, 1, 3, 3, generator=g) * 0.3
    b = torch.randn(C, generator=g) if bias else None
    gy = torch.randn(2, C, H, W, generator=g)
    ins = [t.clone().requires_grad_(True) for t in (x, w) + ((b,) if bias else ())]
    y = F.conv2d(ins[0], ins[1], ins[2] if bias else None, padding=1, groups=C)
    refg = torch.autograd.grad(y, ins, gy)
    xs, ws = leaf(x), leaf(w)
    bs = leaf(b) if bias else None
    A.dwconv3x3(xs, ws, bs).backward(gy.cuda())
    close32(xs.grad, refg[0])
    close32(ws.grad, refg[1])
    if bias:
        close32(bs.grad, refg[2])


@pytest.mark.parametrize("Hd,H,W,bias", [(31, 16, 16, False), (20, 9, 7, True), (127, 24, 16, True)])
def test_gelu_gate_backward(b200ir, Hd, H, W, bias):
    from b200ir import autograd_ops as A
    g = torch.Generator().manual_seed(5)
    x = torch.randn(2, 2 * Hd, H, W, generator=g)
    w = torch.randn(2 * Hd, 1, 3, 3, generator=g) * 0.3
    b = torch.randn(2 * Hd, generator=g) * 0.2 if bias else None
    gy = torch.randn(2, Hd, H, W, generator=g)
    ins = [t.clone().requires_grad_(True) for t in (x, w) + ((b,) if bias else ())]
    t = F.conv2d(ins[0], ins[1], ins[2] if bias else None, padding=1, groups=2 * Hd)
    x1, x2 = t.chunk(2, dim=1)
    y = F.gelu(x1) * x2
    refg = torch.autograd.grad(y, ins, gy)
    xs, ws = leaf(x), leaf(w)
    bs = leaf(b) if bias else None
    A.dwconv3x3_gelu_gate(xs, ws, bs).backward(gy.cuda())
    close32(xs.grad, refg[0])
    close32(ws.grad, refg[1])
    if bias:
        close32(bs.grad, refg[2])


@pytest.mark.parametrize("C,heads,H,W,bias", [(48, 1, 16, 16, False), (64, 2, 8, 24, True), (96, 2, 16, 8, True),
                                              (96, 1, 16, 16, False),
                                              (384, 4, 8, 8, True)])     # AdaIR fre1: head dim 96 at C = 384 (chunked staging)
def test_mdta_core_backward(b200ir, C, heads, H, W, bias):
    from b200ir import autograd_ops as A
    g = torch.Generator().manual_seed(6)
    qkv = torch.randn(2, 3 * C, H, W, generator=g)
    temp = torch.rand(heads, 1, 1, generator=g) + 0.5
    wo = torch.randn(C, C, 1, 1, generator=g) * 0.2
    bo = torch.randn(C, generator=g) if bias else None
    r = torch.randn(2, C, H, W, generator=g)
    gy = torch.randn(2, C, H, W, generator=g)

    ins = [t.clone().requires_grad_(True) for t in (qkv, temp, wo, r) + ((bo,) if bias else ())]
    q, k, v = ins[0].chunk(3, dim=1)
    d = C // heads
    sh = lambda t: t.reshape(2, heads, d, H * W)
    attn = ((F.normalize(sh(q), dim=-1) @ F.normalize(sh(k), dim=-1).transpose(-2, -1)) * ins[1]).softmax(dim=-1)
    y = ins[3] + F.conv2d((attn @ sh(v)).reshape(2, C, H, W), ins[2], ins[4] if bias else None)
    refg = torch.autograd.grad(y, ins, gy)

    qs, ts, wos, rs = leaf(qkv), leaf(temp), leaf(wo), leaf(r)
    bos = leaf(bo) if bias else None
    out = A.mdta_core(qs, ts, wos, bos, heads, rs)
    close32(out, y, 1e-5)
    out.backward(gy.cuda())
    close32(qs.grad, refg[0])
    close32(ts.grad, refg[1])
    close32(wos.grad, refg[2])
    close32(rs.grad, refg[3])
    if bias:
        close32(bos.grad, refg[4])


@pytest.mark.parametrize("Cin,Cout,H,W,shuffle,bias,res", [(3, 48, 16, 16, "none", False, False),
                                                           (48, 24, 16, 24, "unshuffle", False, False),
                                                           (96, 192, 8, 8, "shuffle", False, False),
                                                           (32, 3, 16, 8, "none", True, True)])
def test_conv3x3_backward(b200ir, Cin, Cout, H, W, shuffle, bias, res):
    from b200ir import autograd_ops as A, _lib as L
    g = torch.Generator().manual_seed(7)
    x = torch.randn(2, Cin, H, W, generator=g)
    w = torch.randn(Cout, Cin, 3, 3, generator=g) * 0.1
    b = torch.randn(Cout, generator=g) if bias else None
    ins = [t.clone().requires_grad_(True) for t in (x, w) + ((b,) if bias else ())]
    y = F.conv2d(ins[0], ins[1], ins[2] if bias else None, padding=1)
    if shuffle == "unshuffle":
        y = O.pixel_unshuffle2(y)
    elif shuffle == "shuffle":
        y = O.pixel_shuffle2(y)
    r = torch.randn(y.shape, generator=g) if res else None
    if res:
        rr = r.clone().requires_grad_(True)
        y = y + rr
        ins.append(rr)
    gy = torch.randn(y.shape, generator=g)
    refg = torch.autograd.grad(y, ins, gy)
    mode = {"none": L.SHUFFLE_NONE, "unshuffle": L.UNSHUFFLE2, "shuffle": L.SHUFFLE2}[shuffle]
    xs, ws = leaf(x), leaf(w)
    bs = leaf(b) if bias else None
    rs = leaf(r) if res else None
    out = A.conv3x3(xs, ws, bs, rs, mode)
    close32(out, y, 1e-5)
    out.backward(gy.cuda())
    close32(xs.grad, refg[0])
    close32(ws.grad, refg[1])
    if bias:
        close32(bs.grad, refg[2])
    if res:
        close32(rs.grad, refg[-1])


BLOCKS = [(48, 1, 2.66, False, "WithBias", 16, 16), (96, 2, 2.66, True, "BiasFree", 16, 8),
          (64, 2, 2.0, True, "WithBias", 8, 24), (96, 1, 2.66, False, "WithBias", 16, 16)]


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("C,heads,f,bias,ln,H,W", BLOCKS)
def test_block_backward(b200ir, dtype, C, heads, f, bias, ln, H, W):
    p = O.synth_block_params(C, heads, f, bias, ln, 21)
    x = O.synth_image(2, H, W, seed=9, c=C) * 2 - 1
    gy = torch.randn(2, C, H, W, generator=torch.Generator().manual_seed(10))
    y, dx, dp = O.block_grads(x, p, heads, gy)
    blk = b200ir.TransformerBlock(C, heads, f, bias, ln).cuda()
    blk.load_state_dict(p, strict=True)
    xs = leaf(x, dtype)
    out = blk(xs)
    out.backward(gy.cuda().to(dtype))
    got = dict(blk.named_parameters())
    assert set(got) == set(dp)
    if dtype == torch.float32:
        close32(out, y, 1e-5)
        close32(xs.grad, dx)
        for n in dp:
            close32(got[n].grad, dp[n])
    else:
        assert rel(xs.grad.float(), dx) < 3e-2
        for n in dp:
            assert got[n].grad is not None and got[n].grad.dtype == torch.float32
            assert rel(got[n].grad, dp[n]) < 3e-2, n


def test_l1_loss_and_adamw(b200ir):
    from b200ir import autograd_ops as A, ops
    g = torch.Generator().manual_seed(11)
    pred, tgt = torch.rand(2, 3, 16, 24, generator=g), torch.rand(2, 3, 16, 24, generator=g)
    pr = pred.clone().requires_grad_(True)
    lr_ = O.l1_loss(pr, tgt)
    (gr,) = torch.autograd.grad(lr_ * 3.0, pr)
    ps = leaf(pred)
    ls = A.l1_loss(ps, tgt.cuda())
    (ls * 3.0).backward()
    assert abs(ls.item() - lr_.item()) < 1e-6
    close32(ps.grad, gr, 1e-6)
    # AdamW: three steps against torch.optim.AdamW itself and the oracle restatement
    n = 10007
    p0, m0, v0 = torch.randn(n, generator=g), torch.zeros(n), torch.zeros(n)
    pt = p0.clone().requires_grad_(True)
    opt = torch.optim.AdamW([pt], lr=2e-4)
    pc, mc, vc = p0.cuda(), m0.cuda(), v0.cuda()
    po, mo, vo = p0.clone(), m0.clone(), v0.clone()
    for step in range(1, 4):
        gg = torch.randn(n, generator=g)
        pt.grad = gg.clone()
        opt.step()
        ops.adamw_step(pc, gg.cuda(), mc, vc, 2e-4, 0.9, 0.999, 1e-8, 1e-2, step)
        po, mo, vo = O.adamw_reference(po, gg, mo, vo, step)
    assert (pc.cpu() - pt.detach()).abs().max() <= 1e-6
    assert (po - pt.detach()).abs().max() <= 1e-6


NET_CFG = dict(dim=16, num_blocks=[1, 2, 1, 1], num_refinement_blocks=1, heads=[1, 2, 4, 8])


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_restormer_train_steps(b200ir, dtype):
    """Three full training steps (forward, L1, backward, AdamW) of a small Restormer vs the oracle's restatement of
    the reference training_step (MoCE-IR-main/src/train.py:59-88)."""
    cfg = dict(O.DEFAULT_CFG)
    cfg.update(NET_CFG)
    sd = O.synth_state_dict(NET_CFG, seed=5)
    clean = O.synth_image(2, 32, 32, seed=21)
    noise = torch.randn(clean.shape, generator=torch.Generator().manual_seed(22)) * (25.0 / 255.0)
    degraded = (clean + noise).clamp(0, 1)       # sigma=25 synthetic denoising (degradation_utils.py:21-24)
    losses, grads, new_sd = O.train_step_reference(sd, NET_CFG, degraded, clean, steps=3)

    net = b200ir.Restormer(**cfg).cuda()
    net.load_state_dict(sd, strict=True)
    net.set_compute_dtype(dtype)
    opt = b200ir.AdamW(net.parameters(), lr=2e-4)
    got = []
    for _ in range(3):
        got.append(b200ir.train_step(net, opt, degraded.cuda(), clean.cuda()).item())
    params = dict(net.named_parameters())
    if dtype == torch.float32:
        for a, b in zip(got, losses):
            assert abs(a - b) <= 1e-5, (got, losses)
        for n, gref in grads.items():
            close32(params[n].grad, gref, 5e-4)
        for n, pref in new_sd.items():
            assert (params[n].detach().cpu() - pref).abs().max() <= 2e-5, n
    else:
        for a, b in zip(got, losses):
            assert abs(a - b) <= 2e-2 * abs(b), (got, losses)
        # bf16 activations: EVERY gradient tensor is held to a per-tensor bound, on the FIRST step (identical weights on
        # both sides; after AdamW's sign-like updates the two trajectories are different networks and a comparison of
        # their gradients no longer measures the kernels).  Real-size check: test_parity_fullsize_gpu.py (cfg4).
        _, grads1, _ = O.train_step_reference(sd, NET_CFG, degraded, clean, steps=1)
        net1 = b200ir.Restormer(**cfg).cuda()
        net1.load_state_dict(sd, strict=True)
        net1.set_compute_dtype(dtype)
        b200ir.train_step(net1, b200ir.AdamW(net1.parameters(), lr=2e-4), degraded.cuda(), clean.cuda())
        p1 = dict(net1.named_parameters())
        rows = []
        for n, gref in grads1.items():
            a, r = p1[n].grad.detach().double().cpu().flatten(), gref.double().flatten()
            rows.append((n, float(torch.dot(a, r) / (a.norm() * r.norm()).clamp_min(1e-300)), rel(p1[n].grad, gref)))
        rows.sort(key=lambda t: -t[2])
        print("worst bf16 gradient tensors (name, cosine, rel L2):", rows[:6])
        # toy sizes (4x4-pixel latent, 2 samples): each weight gradient sums over only 32-2048 pixels, so the bf16
        # rounding of the activations does not average out the way it does at the real size, where the bound is
        # cosine >= 0.999 / rel L2 <= 3e-2 (test_parity_fullsize_gpu.py::test_cfg4_train_step_bf16_vs_oracle)
        bad = [t for t in rows if t[1] < 0.995 or t[2] > 0.1]
        assert not bad, bad
    # loss must go down on a repeated batch
    assert got[-1] < got[0]


@pytest.mark.parametrize("B,Cout,Cin,H,W,per_sample,bias", [(2, 144, 48, 16, 16, False, False), (3, 48, 127, 8, 24, False, True),
                                                            (2, 510, 96, 32, 32, False, True), (2, 96, 96, 16, 16, True, True),
                                                            (1, 2042, 384, 8, 8, False, False), (2, 17, 255, 24, 24, False, False),
                                                            (4, 288, 300, 8, 16, True, False)])
def test_conv_wgrad_tc_vs_exact(b200ir, B, Cout, Cin, H, W, per_sample, bias):
    """tcgen05 weight-gradient GEMM (bf16 operands, fp32 accumulate) against the fp64 sum over the same bf16 values and
    against the CUDA-core kernel."""
    from b200ir import ops, _lib as L
    g = torch.Generator().manual_seed(31)
    dy = torch.randn(B, Cout + 5, H, W, generator=g).bfloat16().cuda()[:, 2:2 + Cout]   # channel slice of a wider buffer
    x = torch.randn(B, Cin, H, W, generator=g).bfloat16().cuda()
    if (dy.data_ptr() % 16) or (H * W * 2) % 16:
        pytest.skip("slice not 16B aligned")
    dw_tc, db_tc = ops.conv_wgrad(dy, x, 1, per_sample=per_sample, want_bias=bias, impl=L.IMPL_TC)
    dw_si, db_si = ops.conv_wgrad(dy, x, 1, per_sample=per_sample, want_bias=bias, impl=L.IMPL_SIMT)
    a = dy.double().reshape(B, Cout, -1).cpu()
    b = x.double().reshape(B, Cin, -1).cpu()
    ref = torch.einsum("bop,bcp->boc", a, b)
    if not per_sample:
        ref = ref.sum(0)
    close32(dw_tc, ref, 2e-5)
    close32(dw_si, ref, 2e-5)
    if bias:
        refb = a.sum((0, 2))
        close32(db_tc, refb, 2e-5)
        close32(db_si, refb, 2e-5)
    # determinism
    dw2, _ = ops.conv_wgrad(dy, x, 1, per_sample=per_sample, want_bias=bias, impl=L.IMPL_TC)
    assert torch.equal(dw2, dw_tc)


def test_graphed_train_step_matches_eager(b200ir):
    """CUDA-graph replay of the whole training step == the eager step, bit for bit (all reductions are fixed-order),
    including the device-side AdamW step counter."""
    cfg = dict(O.DEFAULT_CFG)
    cfg.update(NET_CFG)
    sd = O.synth_state_dict(NET_CFG, seed=5)
    clean = O.synth_image(2, 32, 32, seed=21).cuda()
    deg = (clean + 0.1 * torch.randn(clean.shape, generator=torch.Generator().manual_seed(3)).cuda()).clamp(0, 1)

    def make():
        net = b200ir.Restormer(**cfg).cuda()
        net.load_state_dict(sd, strict=True)
        net.set_compute_dtype(torch.bfloat16)
        return net, b200ir.AdamW(net.parameters(), lr=2e-4)

    net_e, opt_e = make()
    le = [b200ir.train_step(net_e, opt_e, deg, clean).item() for _ in range(4)]
    net_g, opt_g = make()
    step = b200ir.GraphedTrainStep(net_g, opt_g, deg, clean)
    lg = [step(deg, clean).item() for _ in range(4)]
    assert le == lg, (le, lg)
    assert torch.equal(opt_e.flat_p, opt_g.flat_p)
    assert opt_g.step_count == 4 and int(opt_g.step_dev.item()) == 4
    # eval forward after graphed training sees the updated weights (packed-weight caches are invalidated)
    with torch.no_grad():
        assert torch.equal(net_e(deg), net_g(deg))


@pytest.mark.parametrize("Hd,H,W,bias", [(31, 64, 64, False), (20, 24, 136, True), (8, 130, 256, True), (127, 16, 72, False),
                                         # small planes: several channels per TMA item (channel groups, ragged last group)
                                         (31, 16, 16, True), (127, 32, 32, False), (20, 40, 32, True), (5, 24, 16, True),
                                         (1021, 16, 16, False)])
def test_gelu_gate_backward_bf16_tma(b200ir, Hd, H, W, bias):
    """bf16 fast path (TMA-staged tiles, fitted cdf) of the gate backward + fused depthwise weight gradient against
    fp32 autograd on the same bf16-rounded inputs."""
    from b200ir import autograd_ops as A
    g = torch.Generator().manual_seed(15)
    x = torch.randn(2, 2 * Hd, H, W, generator=g).bfloat16()
    w = torch.randn(2 * Hd, 1, 3, 3, generator=g) * 0.3
    b = torch.randn(2 * Hd, generator=g) * 0.2 if bias else None
    gy = torch.randn(2, Hd, H, W, generator=g).bfloat16()
    ins = [x.float().requires_grad_(True), w.clone().requires_grad_(True)] + ([b.clone().requires_grad_(True)] if bias else [])
    t = F.conv2d(ins[0], ins[1], ins[2] if bias else None, padding=1, groups=2 * Hd)
    x1, x2 = t.chunk(2, dim=1)
    refg = torch.autograd.grad(F.gelu(x1) * x2, ins, gy.float())
    xs, ws = leaf(x), leaf(w)
    bs = leaf(b) if bias else None
    A.dwconv3x3_gelu_gate(xs, ws, bs).backward(gy.cuda())
    assert rel(xs.grad.float(), refg[0]) < 1e-2
    assert rel(ws.grad, refg[1]) < 1e-2
    if bias:
        assert rel(bs.grad, refg[2]) < 1e-2


@pytest.mark.parametrize("C,H,W,bias", [(24, 64, 64, False), (30, 24, 136, True), (9, 130, 256, True),
                                        (24, 16, 16, True), (30, 32, 32, False), (9, 24, 16, True), (1152, 16, 16, False)])
def test_dwconv_backward_bf16_tma(b200ir, C, H, W, bias):
    from b200ir import autograd_ops as A
    g = torch.Generator().manual_seed(16)
    x = torch.randn(2, C, H, W, generator=g).bfloat16()
    w = torch.randn(C, 1, 3, 3, generator=g) * 0.3
    b = torch.randn(C, generator=g) if bias else None
    gy = torch.randn(2, C, H, W, generator=g).bfloat16()
    ins = [x.float().requires_grad_(True), w.clone().requires_grad_(True)] + ([b.clone().requires_grad_(True)] if bias else [])
    y = F.conv2d(ins[0], ins[1], ins[2] if bias else None, padding=1, groups=C)
    refg = torch.autograd.grad(y, ins, gy.float())
    xs, ws = leaf(x), leaf(w)
    bs = leaf(b) if bias else None
    A.dwconv3x3(xs, ws, bs).backward(gy.cuda())
    assert rel(xs.grad.float(), refg[0]) < 1e-2
    assert rel(ws.grad, refg[1]) < 2e-3
    if bias:
        assert rel(bs.grad, refg[2]) < 2e-3


def test_train_steps_match_reference_golden(b200ir):
    """fp32 CUDA path vs the stored outputs of the REFERENCE's own training loop (tests/golden/train_net_*.npz, made by
    oracle/gen_golden_train.py from /root/reference/Restormer.py + torch.optim.AdamW)."""
    import os
    import numpy as np
    from conftest import GOLDEN
    from oracle.gen_golden_cases import TRAIN_NET_CASE
    name, cfgo, b, h, w, wseed, iseed, steps = TRAIN_NET_CASE
    gold = np.load(os.path.join(GOLDEN, name + ".npz"))
    cfg = dict(O.DEFAULT_CFG)
    cfg.update(cfgo)
    clean = O.synth_image(b, h, w, seed=iseed)
    noise = torch.randn(clean.shape, generator=torch.Generator().manual_seed(iseed + 1)) * (25.0 / 255.0)
    degraded = (clean + noise).clamp(0, 1)
    net = b200ir.Restormer(**cfg).cuda()
    net.load_state_dict(O.synth_state_dict(cfgo, seed=wseed), strict=True)
    opt = b200ir.AdamW(net.parameters(), lr=2e-4)
    losses = [b200ir.train_step(net, opt, degraded.cuda(), clean.cuda()).item() for _ in range(steps)]
    assert np.allclose(losses, gold["losses"], rtol=0, atol=1e-5), (losses, gold["losses"])
    for n, q in net.named_parameters():
        close32(q.grad, torch.from_numpy(gold["g_" + n]), 1e-3)
        assert (q.detach().cpu() - torch.from_numpy(gold["p_" + n])).abs().max() <= 2e-5, n


def test_block_grads_match_reference_golden(b200ir):
    import os
    import numpy as np
    from conftest import GOLDEN
    from oracle.gen_golden_cases import TRAIN_BLOCK_CASES
    for name, dim, heads, f, bias, ln, b, h, w, seed in TRAIN_BLOCK_CASES:
        gold = np.load(os.path.join(GOLDEN, name + ".npz"))
        blk = b200ir.TransformerBlock(dim, heads, f, bias, ln).cuda()
        blk.load_state_dict(O.synth_block_params(dim, heads, f, bias, ln, seed), strict=True)
        xs = leaf(O.synth_image(b, h, w, seed=seed + 100, c=dim) * 2 - 1)
        gy = torch.randn(b, dim, h, w, generator=torch.Generator().manual_seed(seed + 200))
        blk(xs).backward(gy.cuda())
        close32(xs.grad, torch.from_numpy(gold["dx"]))
        for n, q in blk.named_parameters():
            close32(q.grad, torch.from_numpy(gold["d_" + n]))


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("ln", ["WithBias", "BiasFree"])
@pytest.mark.parametrize("B,C,H,W", [(2, 48, 16, 24), (8, 96, 64, 64), (1, 30, 7, 13), (2, 384, 8, 8)])
def test_ln_bwd_with_normalised_byproduct(b200ir, dtype, ln, B, C, H, W):
    """b200ir_ln_bwd: dx (+ residual gradient), dgamma, dbeta and the LN(x) by-product against autograd of the oracle's
    LayerNorm (Restormer.py:37-39,54-57) on the same (rounded) inputs; covers the generic and the bf16x2 kernels."""
    from b200ir import ops
    L = b200ir._lib
    g = torch.Generator().manual_seed(B * C + H)
    x = (torch.randn(B, C, H, W, generator=g) + 0.5).to(dtype)
    gy = torch.randn(B, C, H, W, generator=g).to(dtype)
    r = torch.randn(B, C, H, W, generator=g).to(dtype)
    lw = torch.rand(C, generator=g) + 0.5
    lb = torch.rand(C, generator=g) - 0.5
    xr = x.float().requires_grad_(True)
    lwr, lbr = lw.clone().requires_grad_(True), lb.clone().requires_grad_(True)
    y = O.layernorm(xr, lwr, lbr if ln == "WithBias" else None)
    y.backward(gy.float())
    code = L.LN_WITHBIAS if ln == "WithBias" else L.LN_BIASFREE
    dx, dlw, dlb, xn = ops.ln_bwd(x.cuda(), gy.cuda(), code, lw.cuda(), r.cuda(), lb.cuda() if ln == "WithBias" else None,
                                  want_xn=True)
    tol = 1e-5 if dtype == torch.float32 else 1e-2
    assert rel(dx.float(), xr.grad + r.float()) < tol
    assert rel(xn.float(), y.detach()) < tol
    assert rel(dlw, lwr.grad) < (1e-5 if dtype == torch.float32 else 2e-3)
    if ln == "WithBias":
        assert rel(dlb, lbr.grad) < (1e-5 if dtype == torch.float32 else 2e-3)


def test_backward_rooted_at_a_tma_kernel(b200ir):
    """`y.backward(dy)` with one of the library's Functions as the ROOT of the graph: its backward runs in an autograd
    worker thread whose first CUDA call is then a driver-API tensor-map encode (no runtime call has bound the context to
    that thread yet) -- the encode helper binds it and retries (csrc/gemm_tc.cu make_tmap_bf16_sw)."""
    from b200ir import autograd_ops as A
    g = torch.Generator().manual_seed(31)
    B, C, H = 8, 96, 128
    qkv = (torch.randn(B, 3 * C, H, H, generator=g) * 0.5).cuda().bfloat16().requires_grad_(True)
    temp = torch.ones(1, 1, 1, device="cuda", requires_grad=True)
    w = (torch.randn(C, C, 1, 1, generator=g) / C ** 0.5).cuda().requires_grad_(True)
    dy = torch.randn(B, C, H, H, generator=g).cuda().bfloat16()
    A.mdta_core(qkv, temp, w, None, 1, None).backward(dy)
    torch.cuda.synchronize()
    assert qkv.grad is not None and torch.isfinite(qkv.grad.float()).all() and w.grad is not None
