"""SURVEY.md §8 f1, first pieces of the MoCE-IR decoder path: depthwise K x K conv and the two-source MDTA
(CrossAttention, moce_ir.py:325-368) against the oracle and the reference's stored outputs."""
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from conftest import GOLDEN
from oracle import restormer_oracle as O
from oracle.gen_golden_cases import CROSS_CASES

pytestmark = pytest.mark.gpu


def rel_err(a, b):
    return ((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30)).item()


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("K,C,B,H,W,bias", [(7, 64, 2, 16, 24, True), (7, 9, 1, 33, 70, False), (5, 12, 2, 8, 8, True),
                                            (3, 20, 1, 40, 32, True), (7, 96, 2, 64, 64, False)])
def test_dwconv_kxk(dtype, K, C, B, H, W, bias):
    from b200ir import ops
    g = torch.Generator().manual_seed(K * 100 + C)
    x = (torch.rand(B, C, H, W, generator=g) * 2 - 1).to(dtype)
    w = (torch.rand(C, 1, K, K, generator=g) - 0.5) / K
    b = (torch.rand(C, generator=g) - 0.5) if bias else None
    ref = F.conv2d(x.float(), w, b, padding=K // 2, groups=C)
    y = ops.dwconv_kxk(x.cuda(), w.cuda(), b.cuda() if bias else None).float().cpu()
    if dtype == torch.float32:
        assert (y - ref).abs().max() <= 1e-5
    else:
        assert rel_err(y, ref) < 4e-3


def test_dwconv_kxk_writes_channel_slice():
    """out= may be a channel slice of a wider buffer (how CrossAttention assembles its qkv tensor)."""
    from b200ir import ops
    g = torch.Generator().manual_seed(3)
    x = (torch.rand(2, 10, 12, 16, generator=g) - 0.5).cuda()
    w = (torch.rand(10, 1, 7, 7, generator=g) - 0.5).cuda()
    buf = torch.full((2, 25, 12, 16), 7.0, device="cuda")
    ops.dwconv_kxk(x, w, None, out=buf[:, 5:15])
    ref = F.conv2d(x, w, None, padding=3, groups=10)
    assert (buf[:, 5:15] - ref).abs().max() <= 1e-5
    assert (buf[:, :5] == 7).all() and (buf[:, 15:] == 7).all()


@pytest.mark.parametrize("case", CROSS_CASES, ids=[c[0] for c in CROSS_CASES])
def test_cross_attention_vs_reference_golden(case):
    """b200ir.CrossAttention (fp32 CUDA-core path, then bf16 tensor-core path) against the reference module's output."""
    import b200ir
    name, dim, heads, bias, b, h, w, seed = case
    mod = b200ir.CrossAttention(dim, heads, bias).cuda().eval()
    mod.load_state_dict(O.synth_cross_attention_params(dim, heads, bias, seed), strict=True)
    x = (O.synth_image(b, h, w, seed=seed + 100, c=dim) * 2 - 1).cuda()
    y = (O.synth_image(b, h, w, seed=seed + 300, c=dim) * 2 - 1).cuda()
    gold = torch.from_numpy(np.load(os.path.join(GOLDEN, name + ".npz"))["y"])
    with torch.no_grad():
        out32 = mod(x, y).cpu()
        out16 = mod(x.bfloat16(), y.bfloat16()).float().cpu()
    assert (out32 - gold).abs().max() <= 1e-4
    assert rel_err(out16, gold) < 2e-2


@pytest.mark.parametrize("case", CROSS_CASES, ids=[c[0] for c in CROSS_CASES])
def test_cross_attention_backward_vs_reference_golden(case):
    """Gradients w.r.t. both inputs and every parameter (fp32 path) against the reference module's own autograd."""
    import b200ir
    name, dim, heads, bias, b, h, w, seed = case
    mod = b200ir.CrossAttention(dim, heads, bias).cuda()
    mod.load_state_dict(O.synth_cross_attention_params(dim, heads, bias, seed), strict=True)
    x = (O.synth_image(b, h, w, seed=seed + 100, c=dim) * 2 - 1).cuda().requires_grad_(True)
    y = (O.synth_image(b, h, w, seed=seed + 300, c=dim) * 2 - 1).cuda().requires_grad_(True)
    gy = torch.randn(b, dim, h, w, generator=torch.Generator().manual_seed(seed + 200)).cuda()
    gold = np.load(os.path.join(GOLDEN, name + ".npz"))
    out = mod(x, y)
    out.backward(gy)
    assert (out.detach().cpu() - torch.from_numpy(gold["y"])).abs().max() <= 1e-4

    def close(a, ref, what):
        ref = torch.from_numpy(ref)
        scale = ref.abs().max().clamp_min(1e-6)
        assert (a.detach().cpu() - ref).abs().max() <= 5e-4 * scale, what

    close(x.grad, gold["dx"], "dx")
    close(y.grad, gold["dy"], "dy")
    for n, q in mod.named_parameters():
        close(q.grad, gold["d_" + n], n)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("K,C,B,H,W,bias", [(7, 24, 2, 16, 24, True), (7, 9, 1, 33, 70, False), (5, 12, 2, 8, 8, True),
                                            (7, 96, 2, 64, 64, True)])
def test_dwconv_kxk_backward(dtype, K, C, B, H, W, bias):
    from b200ir import autograd_ops as A
    g = torch.Generator().manual_seed(K * 10 + C)
    x = torch.randn(B, C, H, W, generator=g).to(dtype)
    w = torch.randn(C, 1, K, K, generator=g) * 0.2
    b = torch.randn(C, generator=g) if bias else None
    gy = torch.randn(B, C, H, W, generator=g).to(dtype)
    ins = [x.float().clone().requires_grad_(True), w.clone().requires_grad_(True)] + ([b.clone().requires_grad_(True)] if bias else [])
    refg = torch.autograd.grad(F.conv2d(ins[0], ins[1], ins[2] if bias else None, padding=K // 2, groups=C), ins, gy.float())
    xs = x.cuda().requires_grad_(True)
    ws = w.cuda().requires_grad_(True)
    bs = b.cuda().requires_grad_(True) if bias else None
    A.dwconv_kxk(xs, ws, bs).backward(gy.cuda())
    tol = 1e-5 if dtype == torch.float32 else 1e-2
    assert rel_err(xs.grad.float().cpu(), refg[0]) < tol
    assert rel_err(ws.grad.cpu(), refg[1]) < (1e-5 if dtype == torch.float32 else 2e-3)
    if bias:
        assert rel_err(bs.grad.cpu(), refg[2]) < (1e-5 if dtype == torch.float32 else 2e-3)


def test_cross_attention_larger_vs_oracle():
    """A decoder-sized call (C=96, 2 heads, 64x64) on the tensor-core path against the CPU oracle."""
    import b200ir
    dim, heads = 96, 2
    mod = b200ir.CrossAttention(dim, heads, True).cuda().eval()
    p = O.synth_cross_attention_params(dim, heads, True, 7)
    mod.load_state_dict(p, strict=True)
    x = O.synth_image(2, 64, 64, seed=1, c=dim) * 2 - 1
    y = O.synth_image(2, 64, 64, seed=2, c=dim) * 2 - 1
    ref = O.cross_attention(x.bfloat16().float(), y.bfloat16().float(), p, "", heads)
    with torch.no_grad():
        out = mod(x.cuda().bfloat16(), y.cuda().bfloat16()).float().cpu()
    assert rel_err(out, ref) < 2e-2
