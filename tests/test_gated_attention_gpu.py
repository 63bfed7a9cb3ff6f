"""gated MDTA (head-wise / element-wise sigmoid gate on attn.v) against the reference experiment
module's stored outputs and gradients (tests/golden/gated_mdta_*.npz, oracle/gen_golden_gated.py)."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from oracle import restormer_oracle as O
from oracle.gen_golden_cases import GATED_CASES

pytestmark = pytest.mark.gpu


def rel_err(a, b):
    return ((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30)).item()


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("B,C,G,H,W", [(2, 32, 32, 8, 12), (2, 48, 2, 16, 8), (1, 30, 5, 7, 9), (2, 96, 1, 16, 16)])
def test_sigmoid_gate_fwd_bwd(dtype, B, C, G, H, W):
    from b200ir import autograd_ops as A
    g = torch.Generator().manual_seed(C + G)
    u = torch.randn(B, C, H, W, generator=g).to(dtype)
    gt = torch.randn(B, G, H, W, generator=g).to(dtype)
    gy = torch.randn(B, C, H, W, generator=g).to(dtype)
    ur, gr = u.float().clone().requires_grad_(True), gt.float().clone().requires_grad_(True)
    ref = ur * torch.sigmoid(gr.repeat_interleave(C // G, dim=1))
    ref.backward(gy.float())
    us, gs = u.cuda().requires_grad_(True), gt.cuda().requires_grad_(True)
    y = A.sigmoid_gate(us, gs)
    y.backward(gy.cuda())
    tol = 1e-5 if dtype == torch.float32 else 8e-3
    assert rel_err(y.detach().float().cpu(), ref.detach()) < tol
    assert rel_err(us.grad.float().cpu(), ur.grad) < tol
    assert rel_err(gs.grad.float().cpu(), gr.grad) < tol


@pytest.mark.parametrize("case", GATED_CASES, ids=[c[0] for c in GATED_CASES])
def test_gated_mdta_vs_reference_golden(case):
    import b200ir
    name, dim, heads, bias, gate_type, b, h, w, seed = case
    mod = b200ir.GatedMDTA(dim, heads, bias, gate_type=gate_type).cuda()
    mod.load_state_dict(O.synth_gated_attention_params(dim, heads, bias, gate_type, seed), strict=True)
    x = (O.synth_image(b, h, w, seed=seed + 100, c=dim) * 2 - 1).cuda().requires_grad_(True)
    gy = torch.randn(b, dim, h, w, generator=torch.Generator().manual_seed(seed + 200)).cuda()
    gold = np.load(os.path.join(GOLDEN, name + ".npz"))
    y = mod(x)
    y.backward(gy)
    assert (y.detach().cpu() - torch.from_numpy(gold["y"])).abs().max() <= 1e-4

    def close(a, ref, what):
        ref = torch.from_numpy(ref)
        assert (a.detach().cpu() - ref).abs().max() <= 5e-4 * ref.abs().max().clamp_min(1e-6), what

    close(x.grad, gold["dx"], "dx")
    for n, q in mod.named_parameters():
        close(q.grad, gold["d_" + n], n)
    # bf16 tensor-core path, inference
    with torch.no_grad():
        y16 = mod(x.detach().bfloat16()).float().cpu()
    assert rel_err(y16, torch.from_numpy(gold["y"])) < 2e-2


def test_gated_block_matches_composition():
    """GatedTransformerBlock (LayerNorm folded into the qkv GEMM, residual in the project_out epilogue) == oracle
    composition x + gated_attention(LN(x)); x + ffn(LN(x))."""
    import b200ir
    dim, heads = 32, 2
    blk = b200ir.GatedTransformerBlock(dim, heads, 2.66, True, "WithBias", gate_type="headwise").cuda().eval()
    torch.manual_seed(3)
    for p in blk.parameters():
        p.data.add_(0.1 * torch.randn_like(p))
    x = (O.synth_image(2, 16, 16, seed=9, c=dim) * 2 - 1)
    sd = {k: v.detach().cpu() for k, v in blk.state_dict().items()}
    a = O.layernorm(x, sd["norm1.body.weight"], sd["norm1.body.bias"])
    x1 = x + O.gated_attention(a, sd, "attn.", heads, "headwise")
    ref = x1 + O.feedforward(O.layernorm(x1, sd["norm2.body.weight"], sd["norm2.body.bias"]), sd, "ffn.")
    with torch.no_grad():
        got = blk(x.cuda()).cpu()
    assert (got - ref).abs().max() <= 1e-4
