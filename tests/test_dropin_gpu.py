"""b200ir.accelerate(): in-place swap of the Restormer-family pieces of an existing model for the CUDA kernels, and
the MoCE-IR encoder flavour (names norms.0 / norms.1 / mixer / ffn) against outputs of the reference moce_ir.py."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from oracle import restormer_oracle as O
from oracle.gen_golden_cases import MOCE_CASES, moce_block_params

pytestmark = pytest.mark.gpu


def _close(a, ref, tol=1e-4):
    a, ref = a.detach().double().cpu(), ref.detach().double().cpu()
    assert a.shape == ref.shape
    assert (a - ref).abs().max() <= tol * max(1.0, ref.abs().max().item()), (a - ref).abs().max()


def test_moce_encoder_group_matches_reference_golden():
    import b200ir
    for name, dim, heads, f, bias, ln, nblk, b, h, w, seed in MOCE_CASES:
        gold = np.load(os.path.join(GOLDEN, name + ".npz"))
        grp = b200ir.EncoderResidualGroup(dim, heads, nblk, f, ln, bias).cuda()
        grp.load_state_dict(moce_block_params(dim, heads, f, bias, ln, nblk, seed), strict=True)
        x = (O.synth_image(b, h, w, seed=seed + 100, c=dim) * 2 - 1).cuda().requires_grad_(True)
        gy = torch.randn(b, dim, h, w, generator=torch.Generator().manual_seed(seed + 200)).cuda()
        y = grp(x)
        y.backward(gy)
        _close(y, torch.from_numpy(gold["y"]))
        _close(x.grad, torch.from_numpy(gold["dx"]))
        for n, q in grp.named_parameters():
            ref = torch.from_numpy(gold["d_" + n])
            assert (q.grad.cpu() - ref).abs().max() <= 1e-4 * ref.abs().max().clamp_min(1e-3), (name, n)
        assert grp.loss == 0


def test_accelerate_swaps_in_place_and_keeps_the_contract():
    import b200ir
    from torch_models import RefMiniNet
    torch.manual_seed(0)
    net = RefMiniNet(16)
    for p in net.parameters():  # non-trivial LayerNorm / temperature parameters
        if p.dim() == 1 or p.dim() == 3:
            p.data.add_(0.3 * torch.randn(p.shape))
    img = O.synth_image(2, 32, 48, seed=3)
    with torch.no_grad():
        ref = net(img)
    keys = list(net.state_dict().keys())
    params = {n: p for n, p in net.named_parameters()}
    net = net.cuda()
    net, stats = b200ir.accelerate(net, verify=True)
    print(stats)
    # 1 Restormer block + 2 MoCE encoder blocks; the gated block is rejected as a block and as an attention, its
    # LayerNorms / FeedForward still match individually; decoder pieces are swapped one by one
    assert stats["block"] == 1 and stats["encoder_block"] == 2, stats
    assert stats["rejected"] == 2, stats
    assert stats["attention"] == 1 and stats["feedforward"] == 2 and stats["layernorm"] == 3, stats
    assert stats["downsample"] == 1 and stats["upsample"] == 1, stats
    assert isinstance(net.enc[0], b200ir.TransformerBlock) and isinstance(net.mid[0], b200ir.EncoderBlock)
    assert type(net.gated).__name__ == "RefTransformerBlock"
    assert list(net.state_dict().keys()) == keys                                  # checkpoint contract unchanged
    assert all(dict(net.named_parameters())[n] is not None for n in params)
    with torch.no_grad():
        got = net(img.cuda())
    _close(got, ref, 2e-4)
    # training through the swapped modules: gradients reach the ORIGINAL parameter objects
    net.zero_grad()
    loss = (net(img.cuda()) - 0.5).abs().mean()
    loss.backward()
    assert all(p.grad is not None for p in net.parameters())


def test_accelerate_adopts_cross_attention():
    """A module with CrossAttention's structure (q / q_dwconv 3x3 / kv / kv_dwconv 7x7 / project_out) is swapped for
    b200ir.CrossAttention, forward and backward keep working through the ORIGINAL parameters."""
    import b200ir
    from torch_models import RefDecoderMixerNet
    torch.manual_seed(1)
    net = RefDecoderMixerNet(32, 2)
    net.mixer.temperature.data.add_(0.3 * torch.randn(2, 1, 1))
    x = O.synth_image(2, 16, 24, seed=5, c=32) * 2 - 1
    xr = x.clone().requires_grad_(True)
    ref = net(xr)
    ref.abs().mean().backward()
    gref = {n: p.grad.clone() for n, p in net.named_parameters()}
    net.zero_grad()
    net = net.cuda()
    net, stats = b200ir.accelerate(net, verify=True)
    assert stats == {"cross_attention": 1}, stats
    assert isinstance(net.mixer, b200ir.CrossAttention)
    xg = x.cuda().requires_grad_(True)
    got = net(xg)
    _close(got.detach(), ref.detach(), 2e-4)
    got.abs().mean().backward()
    _close(xg.grad, xr.grad, 1e-3)
    for n, p in net.named_parameters():
        _close(p.grad, gref[n], 1e-3)


def test_accelerate_adopts_gated_mdta():
    import b200ir
    from torch_models import RefGatedMDTA
    torch.manual_seed(2)
    net = torch.nn.Sequential(RefGatedMDTA(32, 2, True, "headwise"), RefGatedMDTA(32, 4, False, "elementwise"))
    x = O.synth_image(2, 16, 24, seed=6, c=32) * 2 - 1
    with torch.no_grad():
        ref = net(x)
    net = net.cuda()
    net, stats = b200ir.accelerate(net, verify=True)
    assert stats == {"gated_attention": 2}, stats
    assert isinstance(net[0], b200ir.GatedMDTA) and net[1].gate_type == "elementwise"
    with torch.no_grad():
        _close(net(x.cuda()), ref, 2e-4)
