"""GPU parity of every kernel (through the C ABI) against the CPU oracle on the same seeded inputs."""
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from conftest import GOLDEN
from oracle import restormer_oracle as O
from oracle.gen_golden_cases import BLOCK_CASES

pytestmark = pytest.mark.gpu

# tolerances: fp32 path max-abs <= 1e-4 (BASELINE.json north_star); bf16 path judged by PSNR >= 50 dB on
# whole-net output, per-op by relative error against the fp32 oracle of the bf16-rounded inputs.
TOL32 = 1e-4


def dev(t, dtype=None):
    t = t.cuda()
    return t.to(dtype) if dtype is not None else t


def rel_err(a, b):
    return ((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30)).item()


@pytest.fixture(scope="module")
def b200ir():
    import b200ir as m
    return m


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("C,H,W,bias", [(144, 16, 24, False), (48, 8, 8, True), (30, 7, 13, True), (96, 40, 64, False),
                                        (48, 33, 200, True), (24, 128, 128, False), (16, 70, 264, False),
                                        # narrow planes on the TMA path: several channels per work item, ragged groups
                                        (1152, 32, 32, False), (50, 16, 16, True), (7, 24, 40, True), (3, 9, 56, False)])
def test_dwconv3x3(b200ir, dtype, C, H, W, bias):
    g = torch.Generator().manual_seed(1)
    x = (torch.rand(2, C, H, W, generator=g) * 2 - 1).to(dtype)
    w = torch.rand(C, 1, 3, 3, generator=g) - 0.5
    b = (torch.rand(C, generator=g) - 0.5) if bias else None
    ref = F.conv2d(x.float(), w, b, padding=1, groups=C)
    y = b200ir.ops.dwconv3x3(dev(x), dev(w), dev(b) if bias else None).float().cpu()
    if dtype == torch.float32:
        assert (y - ref).abs().max() <= 1e-5
    else:
        assert (y - ref).abs().max() <= 2e-2 and rel_err(y, ref) < 4e-3


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("Hd,H,W,bias", [(127, 16, 16, False), (64, 8, 24, True), (255, 9, 7, True),
                                         (127, 72, 136, True), (32, 64, 64, False), (15, 130, 256, False),
                                         (1021, 32, 32, False), (31, 16, 16, True), (5, 20, 48, True)])
def test_dwconv3x3_gelu_gate(b200ir, dtype, Hd, H, W, bias):
    g = torch.Generator().manual_seed(2)
    x = (torch.rand(2, 2 * Hd, H, W, generator=g) * 4 - 2).to(dtype)
    w = torch.rand(2 * Hd, 1, 3, 3, generator=g) - 0.5
    b = (torch.rand(2 * Hd, generator=g) - 0.5) if bias else None
    t = F.conv2d(x.float(), w, b, padding=1, groups=2 * Hd)
    x1, x2 = t.chunk(2, dim=1)
    ref = F.gelu(x1) * x2
    y = b200ir.ops.dwconv3x3_gelu_gate(dev(x), dev(w), dev(b) if bias else None).float().cpu()
    if dtype == torch.float32:
        assert (y - ref).abs().max() <= 1e-5
    else:
        assert rel_err(y, ref) < 5e-3


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("ln", ["none", "WithBias", "BiasFree"])
@pytest.mark.parametrize("Cin,Cout,HW,bias,res", [(48, 144, (16, 24), False, False), (96, 510, (8, 8), True, True),
                                                  (127, 48, (5, 13), False, True), (384, 192, (8, 8), True, False)])
def test_ln_conv1x1(b200ir, dtype, ln, Cin, Cout, HW, bias, res):
    L = b200ir._lib
    g = torch.Generator().manual_seed(3)
    H, W = HW
    x = (torch.rand(2, Cin, H, W, generator=g) * 2 - 0.5).to(dtype)
    w = (torch.rand(Cout, Cin, generator=g) - 0.5) / Cin ** 0.5
    b = (torch.rand(Cout, generator=g) - 0.5) if bias else None
    lw = 1 + 0.2 * (torch.rand(Cin, generator=g) - 0.5)
    lb = 0.1 * (torch.rand(Cin, generator=g) - 0.5)
    r = (torch.rand(2, Cout, H, W, generator=g) - 0.5).to(dtype) if res else None
    xn = x.float()
    if ln != "none":
        xn = O.layernorm(xn, lw, lb if ln == "WithBias" else None)
    ref = F.conv2d(xn, w.view(Cout, Cin, 1, 1), b)
    if res:
        ref = ref + r.float()
    code = {"none": L.LN_NONE, "WithBias": L.LN_WITHBIAS, "BiasFree": L.LN_BIASFREE}[ln]
    y = b200ir.ops.ln_conv1x1(dev(x), dev(w), dev(b) if bias else None, code, dev(lw), dev(lb),
                              dev(r) if res else None).float().cpu()
    if dtype == torch.float32:
        assert (y - ref).abs().max() <= 2e-5
    else:
        assert rel_err(y, ref) < 6e-3


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("C,heads,H,W", [(48, 1, 16, 24), (96, 2, 16, 16), (96, 1, 24, 40), (64, 2, 8, 8),
                                         (192, 4, 8, 16), (384, 8, 8, 8)])
def test_mdta_core(b200ir, dtype, C, heads, H, W):
    """Gram + norms + softmax + fold + apply against Restormer.py:117-131 restated in the oracle."""
    g = torch.Generator().manual_seed(4)
    qkv = (torch.rand(2, 3 * C, H, W, generator=g) * 2 - 1).to(dtype)
    temp = 0.5 + torch.rand(heads, 1, 1, generator=g)
    wout = (torch.rand(C, C, 1, 1, generator=g) - 0.5) / C ** 0.5
    bout = torch.rand(C, generator=g) - 0.5
    res = (torch.rand(2, C, H, W, generator=g) - 0.5).to(dtype)
    q, k, v = qkv.float().chunk(3, dim=1)
    d = C // heads
    q = F.normalize(q.reshape(2, heads, d, H * W), dim=-1)
    k = F.normalize(k.reshape(2, heads, d, H * W), dim=-1)
    attn = ((q @ k.transpose(-2, -1)) * temp).softmax(dim=-1)
    out = (attn @ v.reshape(2, heads, d, H * W)).reshape(2, C, H, W)
    ref = F.conv2d(out, wout, bout) + res.float()
    from b200ir import ops
    import importlib
    A = importlib.import_module("b200ir.autograd_ops")
    with torch.no_grad():
        y = A.mdta_core(dev(qkv), dev(temp), dev(wout), dev(bout), heads, dev(res)).float().cpu()
    if dtype == torch.float32:
        assert (y - ref).abs().max() <= 2e-5
    else:
        assert rel_err(y, ref) < 6e-3


@pytest.mark.parametrize("B,C,heads,nslab", [(16, 384, 8, 2), (16, 192, 4, 4), (16, 96, 1, 9), (3, 80, 2, 3),
                                            (2, 24, 1, 1), (5, 112, 1, 2), (16, 48, 1, 18)])
@pytest.mark.parametrize("packed", [False, True])
def test_mdta_softmax_fold_shapes(b200ir, B, C, heads, nslab, packed):
    """Slab reduce + normalise + temperature softmax + W_out fold (Restormer.py:121-131) over every row-block variant of
    the register-tiled fold kernel (64 / 32 / 16 rows per CTA) and head dims that are not multiples of 16."""
    from b200ir import ops
    d = C // heads
    g = torch.Generator().manual_seed(C + heads)
    parts = torch.rand(B, heads, nslab, d * d + 2 * d, generator=g) - 0.3
    parts[..., d * d:] = parts[..., d * d:].abs() + 0.05          # squared norms are positive
    temp = 0.5 + torch.rand(heads, generator=g)
    wout = (torch.rand(C, C, generator=g) - 0.5) / C ** 0.5
    tot = parts.double().sum(2)
    G = tot[..., :d * d].reshape(B, heads, d, d)
    nq = tot[..., d * d:d * d + d].sqrt().clamp_min(1e-12)
    nk = tot[..., d * d + d:].sqrt().clamp_min(1e-12)
    attn = (G / (nq[..., :, None] * nk[..., None, :]) * temp.double()[None, :, None, None]).softmax(-1)
    m = torch.zeros(B, C, C, dtype=torch.float64)
    for h in range(heads):
        m[:, :, h * d:(h + 1) * d] = wout.double()[None, :, h * d:(h + 1) * d] @ attn[:, h]
    with torch.no_grad():
        out, a = ops.mdta_softmax_fold(dev(parts), dev(temp), dev(wout), want_attn=True, packed=packed)
    assert (a.cpu().double() - attn).abs().max() <= 2e-6
    if packed:
        o = out.float().cpu()
        assert (o[:, :C, :C].double() - m).abs().max() <= 4e-3 * m.abs().max()     # bf16 operand
        assert o[:, C:, :].abs().max() == 0 if o.shape[1] > C else True            # zero padding rows / columns
        assert o[:, :, C:].abs().max() == 0 if o.shape[2] > C else True
    else:
        assert (out.cpu().double() - m).abs().max() <= 2e-6


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("mode,Cin,Cout,H,W", [("none", 3, 48, 16, 24), ("unshuffle", 48, 24, 16, 16),
                                               ("shuffle", 96, 192, 8, 12), ("none", 96, 3, 20, 18),
                                               ("unshuffle", 20, 10, 6, 34)])
def test_conv3x3(b200ir, dtype, mode, Cin, Cout, H, W):
    L = b200ir._lib
    g = torch.Generator().manual_seed(5)
    x = (torch.rand(2, Cin, H, W, generator=g) * 2 - 1).to(dtype)
    w = (torch.rand(Cout, Cin, 3, 3, generator=g) - 0.5) / (Cin * 9) ** 0.5
    ref = F.conv2d(x.float(), w, None, padding=1)
    if mode == "unshuffle":
        ref = O.pixel_unshuffle2(ref)
    elif mode == "shuffle":
        ref = O.pixel_shuffle2(ref)
    r = (torch.rand(ref.shape, generator=g) - 0.5).to(dtype)
    ref = ref + r.float()
    code = {"none": L.SHUFFLE_NONE, "unshuffle": L.UNSHUFFLE2, "shuffle": L.SHUFFLE2}[mode]
    y = b200ir.ops.conv3x3(dev(x), dev(w), None, dev(r), code).float().cpu()
    if dtype == torch.float32:
        assert (y - ref).abs().max() <= 2e-5
    else:
        assert rel_err(y, ref) < 6e-3


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("Cin,Cout,B,H,W,bias,res", [
    (3, 48, 2, 9, 13, True, False),      # patch embed, odd width (scalar stores), ragged tile
    (3, 48, 1, 24, 128, False, False),   # two column tiles, vector stores
    (1, 20, 2, 10, 66, True, True),      # grey input, Cout not a multiple of 8
    (96, 3, 2, 16, 256, False, True),    # output conv, 16-byte staging path + `+ inp_img`
    (50, 3, 1, 11, 130, True, True),     # ragged channels / width: scalar staging
    (48, 1, 2, 8, 128, True, False),     # grey output
])
def test_conv3x3_thin(b200ir, dtype, Cin, Cout, B, H, W, bias, res):
    """The thin-input / thin-output kernels (conv3x3_thin.cu) behind b200ir_conv3x3_fwd, Restormer.py:220,257,306."""
    L = b200ir._lib
    g = torch.Generator().manual_seed(11)
    x = (torch.rand(B, Cin, H, W, generator=g) * 2 - 1).to(dtype)
    w = (torch.rand(Cout, Cin, 3, 3, generator=g) - 0.5) / (Cin * 9) ** 0.5
    bv = torch.rand(Cout, generator=g) - 0.5 if bias else None
    ref = F.conv2d(x.float(), w, bv, padding=1)
    r = None
    if res:
        r = (torch.rand(ref.shape, generator=g) - 0.5).to(dtype)
        ref = ref + r.float()
    y = b200ir.ops.conv3x3(dev(x), dev(w), dev(bv) if bias else None, dev(r) if res else None, L.SHUFFLE_NONE)
    y = y.float().cpu()
    if dtype == torch.float32:
        assert (y - ref).abs().max() <= 2e-5
    else:
        assert rel_err(y, ref) < 6e-3


@pytest.mark.parametrize("case", BLOCK_CASES, ids=[c[0] for c in BLOCK_CASES])
def test_block_vs_golden_fp32(b200ir, case):
    """TransformerBlock on the GPU (fp32 path) against the reference's stored output."""
    name, dim, heads, f, bias, ln, b, h, w, seed = case
    gold = np.load(os.path.join(GOLDEN, name + ".npz"))
    p = O.synth_block_params(dim, heads, f, bias, ln, seed)
    x = O.synth_image(b, h, w, seed=seed + 100, c=dim) * 2 - 1
    blk = b200ir.TransformerBlock(dim, heads, f, bias, ln).cuda()
    blk.load_state_dict(p, strict=True)
    with torch.no_grad():
        y = blk(x.cuda()).cpu()
        y_attn = blk.attn(blk.norm1(x.cuda())).cpu()
        y_ffn = blk.ffn(blk.norm2(x.cuda())).cpu()
        y_n1 = blk.norm1(x.cuda()).cpu()
    assert np.abs(y_n1.numpy() - gold["y_norm1"]).max() <= TOL32
    assert np.abs(y_attn.numpy() - gold["y_attn"]).max() <= TOL32
    assert np.abs(y_ffn.numpy() - gold["y_ffn"]).max() <= TOL32
    assert np.abs(y.numpy() - gold["y_block"]).max() <= TOL32


@pytest.mark.parametrize("case", BLOCK_CASES, ids=[c[0] for c in BLOCK_CASES])
def test_block_bf16(b200ir, case):
    name, dim, heads, f, bias, ln, b, h, w, seed = case
    gold = np.load(os.path.join(GOLDEN, name + ".npz"))
    p = O.synth_block_params(dim, heads, f, bias, ln, seed)
    x = O.synth_image(b, h, w, seed=seed + 100, c=dim) * 2 - 1
    blk = b200ir.TransformerBlock(dim, heads, f, bias, ln).cuda()
    blk.load_state_dict(p, strict=True)
    with torch.no_grad():
        y = blk(x.cuda().bfloat16()).float().cpu()
    assert rel_err(y, torch.from_numpy(gold["y_block"])) < 1e-2
