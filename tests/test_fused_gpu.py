"""Fused LayerNorm + 1x1 conv + depthwise 3x3 (+ GELU gate) kernel (csrc/fused_ln_proj_dw.cu) against the fp32 oracle
math (Restormer.py:54-57,88-91 / :114) on the same bf16-rounded inputs, and against the two-kernel GPU path."""
import pytest
import torch
import torch.nn.functional as F

from oracle import restormer_oracle as O

pytestmark = pytest.mark.gpu


def rel_err(a, b):
    return ((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30)).item()


GATE_CASES = [
    # C, Hd, H, W, B, bias
    (48, 127, 16, 64, 1, False),
    (96, 255, 40, 128, 2, True),
    (96, 255, 9, 72, 1, False),      # ragged: 2 strips, H not a multiple of the band height
    (48, 127, 24, 256, 1, True),     # 5 strips
    (64, 128, 16, 24, 3, True),      # MoCE-IR ffn factor 2 (Hd = 2C), plane narrower than one strip
    (128, 340, 16, 64, 1, False),    # 3 channel blocks, 3 K blocks of weights
    (96, 255, 70, 56, 1, False),     # several bands per strip
    (32, 85, 8, 16, 2, True),
]


def _inputs(C, Ctot, H, W, B, bias, seed):
    g = torch.Generator().manual_seed(seed)
    x = (torch.rand(B, C, H, W, generator=g) * 2 - 0.5).bfloat16()
    w = (torch.rand(Ctot, C, generator=g) - 0.5) / C ** 0.5 * 2
    b = (torch.rand(Ctot, generator=g) - 0.5) if bias else None
    lw = 1 + 0.2 * (torch.rand(C, generator=g) - 0.5)
    lb = 0.1 * (torch.rand(C, generator=g) - 0.5)
    dw = (torch.rand(Ctot, 1, 3, 3, generator=g) - 0.5) * 0.6
    db = (torch.rand(Ctot, generator=g) - 0.5) if bias else None
    return x, w, b, lw, lb, dw, db


@pytest.mark.parametrize("ln", ["WithBias", "BiasFree"])
@pytest.mark.parametrize("C,Hd,H,W,B,bias", GATE_CASES)
def test_fused_ln_proj_dwgate(ln, C, Hd, H, W, B, bias):
    import b200ir
    from b200ir import ops
    L = b200ir._lib
    x, w, b, lw, lb, dw, db = _inputs(C, 2 * Hd, H, W, B, bias, C * 3 + Hd + W)
    xn = O.layernorm(x.float(), lw, lb if ln == "WithBias" else None)
    t = F.conv2d(xn, w.view(2 * Hd, C, 1, 1), b)
    a = F.conv2d(t, dw, db, padding=1, groups=2 * Hd)
    ref = F.gelu(a[:, :Hd]) * a[:, Hd:]
    code = L.LN_WITHBIAS if ln == "WithBias" else L.LN_BIASFREE
    xd = x.cuda()
    assert ops.ln_proj_dw_usable(xd, Hd, True)
    cu = lambda v: v.cuda() if v is not None else None
    wf = ops.pack_fused_proj_dw(w.cuda(), cu(b), code, lw.cuda(), lb.cuda() if ln == "WithBias" else None, Hd, True)
    y = ops.ln_proj_dw(xd, wf, dw.cuda(), cu(db), code, Hd, True)
    y2 = ops.ln_proj_dw(xd, wf, dw.cuda(), cu(db), code, Hd, True)
    torch.cuda.synchronize()
    assert torch.equal(y, y2)                                   # deterministic
    e = rel_err(y.float().cpu(), ref)
    assert e < 1e-2, e
    # the two-kernel path on identical inputs (stores the 2Hd-channel tensor as bf16 in between)
    packed = ops.pack_conv1x1(w.cuda(), cu(b), code, lw.cuda(), lb.cuda() if ln == "WithBias" else None)
    t2 = ops.gemm_tc(xd, packed, 2 * Hd, code) if ops.gemm_tc_usable(xd, 2 * Hd) else \
        ops.ln_conv1x1(xd, w.cuda(), cu(b), code, lw.cuda(), lb.cuda() if ln == "WithBias" else None)
    y3 = ops.dwconv3x3_gelu_gate(t2, dw.cuda(), cu(db))
    e2 = rel_err(y3.float().cpu(), ref)
    print(f"fused rel err {e:.2e}   two-kernel rel err {e2:.2e}")
    assert e <= max(1.5 * e2, 6e-3)                             # no less accurate than the path it replaces


PLAIN_CASES = [
    (48, 144, 16, 64, 1, False),
    (96, 288, 24, 128, 2, True),
    (96, 288, 9, 72, 1, False),
    (128, 384, 16, 64, 1, True),
    (32, 96, 8, 16, 2, False),
]


@pytest.mark.parametrize("ln", ["WithBias", "BiasFree"])
@pytest.mark.parametrize("C,Cout,H,W,B,bias", PLAIN_CASES)
def test_fused_ln_proj_dw_plain(ln, C, Cout, H, W, B, bias):
    import b200ir
    from b200ir import ops
    L = b200ir._lib
    x, w, b, lw, lb, dw, db = _inputs(C, Cout, H, W, B, bias, C * 5 + Cout + W)
    xn = O.layernorm(x.float(), lw, lb if ln == "WithBias" else None)
    ref = F.conv2d(F.conv2d(xn, w.view(Cout, C, 1, 1), b), dw, db, padding=1, groups=Cout)
    code = L.LN_WITHBIAS if ln == "WithBias" else L.LN_BIASFREE
    xd = x.cuda()
    assert ops.ln_proj_dw_usable(xd, Cout, False)
    cu = lambda v: v.cuda() if v is not None else None
    wf = ops.pack_fused_proj_dw(w.cuda(), cu(b), code, lw.cuda(), lb.cuda() if ln == "WithBias" else None, Cout, False)
    y = ops.ln_proj_dw(xd, wf, dw.cuda(), cu(db), code, Cout, False)
    torch.cuda.synchronize()
    e = rel_err(y.float().cpu(), ref)
    assert e < 1e-2, e


def test_fused_full_size_properties():
    """BASELINE cfg2 shape of the dominant level (B16, C=96, Hd=255, 256x256): determinism, batch independence and
    agreement with the two-kernel path (the oracle would need minutes here)."""
    import b200ir
    from b200ir import ops
    L = b200ir._lib
    C, Hd, H, W, B = 96, 255, 256, 256, 16
    x, w, b, lw, lb, dw, db = _inputs(C, 2 * Hd, H, W, B, False, 7)
    xd = x.cuda()
    wf = ops.pack_fused_proj_dw(w.cuda(), None, L.LN_WITHBIAS, lw.cuda(), lb.cuda(), Hd, True)
    y = ops.ln_proj_dw(xd, wf, dw.cuda(), None, L.LN_WITHBIAS, Hd, True)
    y1 = ops.ln_proj_dw(xd[3:4], wf, dw.cuda(), None, L.LN_WITHBIAS, Hd, True)
    packed = ops.pack_conv1x1(w.cuda(), None, L.LN_WITHBIAS, lw.cuda(), lb.cuda())
    y3 = ops.dwconv3x3_gelu_gate(ops.gemm_tc(xd, packed, 2 * Hd, L.LN_WITHBIAS), dw.cuda(), None)
    torch.cuda.synchronize()
    assert torch.equal(y[3:4], y1)
    assert rel_err(y.float(), y3.float()) < 8e-3


def test_fused_unsupported_shapes_fall_back():
    """C = 192 (weights 128 KB + three 26 KB activation stages + output staging > 227 KB of shared memory) and C = 384 are
    not taken by the fused kernel: `ln_proj_dw_usable` says so and FeedForward keeps the two-kernel path there."""
    from b200ir import ops
    for C, Hd in ((192, 510), (384, 1021), (40, 106)):
        x = torch.zeros(1, C, 16, 64, device="cuda", dtype=torch.bfloat16)
        assert not ops.ln_proj_dw_usable(x, Hd, True)
    assert not ops.ln_proj_dw_usable(torch.zeros(1, 96, 16, 64, device="cuda"), 255, True)   # fp32 activations
